"""`Tensor.toString` formatting (Tensors.scala:776-811): nested brackets, elements printed like Scala's Float.toString."""
import numpy as np


def float_to_string(x) -> str:
    """java.lang.Float.toString: shortest digits that round-trip, at least one fraction digit, computerised scientific
    notation outside [1e-3, 1e7)."""
    x = np.float32(x)
    if np.isnan(x):
        return "NaN"
    if np.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0:
        return "-0.0" if np.signbit(x) else "0.0"
    mag = abs(float(x))
    if 1e-3 <= mag < 1e7:
        s = np.format_float_positional(x, unique=True, trim="0")
        return s + "0" if s.endswith(".") else s
    mant, exp = np.format_float_scientific(x, unique=True, trim="0", exp_digits=1).split("e")
    if "." not in mant:
        mant += ".0"
    elif mant.endswith("."):
        mant += "0"
    return "%sE%d" % (mant, int(exp))


def format_tensor(shape, flat) -> str:
    shape = list(shape)
    if not shape:
        if len(flat) != 1:
            raise ValueError("shape [] does not match %d elements" % len(flat))
        return float_to_string(flat[0])
    if len(shape) == 1:
        if shape[0] != len(flat):
            raise ValueError("shape does not match the data size")
        return "[" + ",".join(float_to_string(v) for v in flat) + "]"
    step = len(flat) // shape[0] if shape[0] else 0
    return "[" + ",".join(format_tensor(shape[1:], flat[i * step:(i + 1) * step]) for i in range(shape[0])) + "]"
