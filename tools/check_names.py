"""Poor man's pyflakes (none is installed here): report names a function loads that are bound nowhere in the function, its
module, or builtins.  Run before a GPU call: a NameError in bench.py costs a whole box."""
import ast
import builtins
import sys


def check(path):
    tree = ast.parse(open(path).read())
    module_names = set(dir(builtins)) | {"__file__", "__name__", "__doc__"}
    for node in ast.walk(tree):
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            for a in node.names:
                module_names.add((a.asname or a.name).split(".")[0])
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)):
            module_names.add(node.name)
        elif isinstance(node, (ast.Assign, ast.AnnAssign, ast.For, ast.With, ast.If, ast.Try)):
            for n in ast.walk(node):
                if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Store):
                    module_names.add(n.id)
    bad = 0

    def own_nodes(fn):
        """Nodes of `fn` itself; nested functions / lambdas / comprehensions are walked too, nested defs are not."""
        stack = list(ast.iter_child_nodes(fn))
        while stack:
            n = stack.pop()
            yield n
            if not isinstance(n, (ast.FunctionDef, ast.ClassDef)):
                stack.extend(ast.iter_child_nodes(n))

    def visit(fn, outer):
        nonlocal bad
        local = set(outer)
        nested = []
        for n in own_nodes(fn):
            if isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
                local.add(n.id)
            elif isinstance(n, (ast.FunctionDef, ast.ClassDef)):
                local.add(n.name)
                nested.append(n)
            elif isinstance(n, (ast.Import, ast.ImportFrom)):
                for a in n.names:
                    local.add((a.asname or a.name).split(".")[0])
            elif isinstance(n, ast.ExceptHandler) and n.name:
                local.add(n.name)
            elif isinstance(n, ast.arg):
                local.add(n.arg)
        for n in own_nodes(fn):
            if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Load) and n.id not in local:
                print("%s:%d: undefined name %r in %s" % (path, n.lineno, n.id, getattr(fn, "name", "<module>")))
                bad += 1
        for n in nested:
            if isinstance(n, ast.FunctionDef):
                visit(n, local)
            else:                                   # class body: methods see the enclosing function scope, not the class scope
                for m in n.body:
                    if isinstance(m, ast.FunctionDef):
                        visit(m, local)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef):
            visit(node, module_names)
        elif isinstance(node, ast.ClassDef):
            for m in node.body:
                if isinstance(m, ast.FunctionDef):
                    visit(m, module_names)
    return bad


if __name__ == "__main__":
    sys.exit(1 if sum(check(p) for p in sys.argv[1:]) else 0)
