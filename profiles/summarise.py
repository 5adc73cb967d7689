#!/usr/bin/env python
"""Turns an ncu report (gpurun_out/*.ncu-rep, scratch) into the committed summaries under profiles/.
usage: python profiles/summarise.py gpurun_out/prof_c2_r1.ncu-rep profiles/r1_c2 [launches.csv]"""
import csv
import json
import subprocess
import sys
from collections import Counter, OrderedDict

KEYS = OrderedDict([
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput % of ncu peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm throughput %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor memory (TMEM) pipe active % of elapsed"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "tensor HMMA sub-pipe active cycles (avg per SM)"),
    ("sm__cycles_elapsed.avg", "SM cycles elapsed (avg)"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait / issue"),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not_selected / issue"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_pipe_throttle / issue"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier / issue"),
])


def ncu(args):
    return subprocess.run(["ncu"] + args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout


def to_bytes(value, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(value) * scale.get(unit, 1)


def main():
    rep, out = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(ncu(["-i", rep, "--page", "raw", "--csv"]).splitlines()))
    hdr, units = rows[0], rows[1]
    md = ["# ncu summary of `%s`" % rep.split("/")[-1], "",
          "Captured with `ncu --set full --clock-control none --import-source on` under gpurun on a B200 (sm_100a).",
          "Per-launch values; `dram throughput %` is relative to ncu's own nominal peak, not MEASURED_PEAKS.json.", ""]
    traffic = {}
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = d["Kernel Name"]
        md.append("## %s (launch id %s)" % (name, d.get("ID", "?")))
        md.append("")
        md.append("| metric | value |")
        md.append("|---|---|")
        for k, label in KEYS.items():
            if k in d and d[k] != "":
                md.append("| %s | %s %s |" % (label, d[k], units[hdr.index(k)]))
        rd = to_bytes(d["dram__bytes_read.sum"], units[hdr.index("dram__bytes_read.sum")])
        wr = to_bytes(d["dram__bytes_write.sum"], units[hdr.index("dram__bytes_write.sum")])
        dur = float(d["gpu__time_duration.sum"])
        dunit = units[hdr.index("gpu__time_duration.sum")]
        dur_s = dur * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1}.get(dunit, 1e-9)
        md.append("| dram traffic (read + write) | %.4f GB -> %.0f GB/s under the profiler |" % ((rd + wr) / 1e9, (rd + wr) / dur_s / 1e9))
        md.append("")
        key = name.split("(")[0].replace("void ", "").strip()
        traffic.setdefault(key, []).append(rd + wr)
    # opcode histogram of the first kernel from the source page
    src = list(csv.reader(ncu(["-i", rep, "--page", "source", "--csv"]).splitlines()))
    blocks, cur = [], None
    for r in src:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            blocks.append(cur)
        elif r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and r:
            cur["rows"].append(r)
    seen = set()
    for b in blocks:
        if b["name"] in seen or "hdr" not in b:
            continue
        seen.add(b["name"])
        h = b["hdr"]
        ia, isrc = h.index("Instructions Executed"), h.index("Source")
        c, tot = Counter(), 0
        for r in b["rows"]:
            n = int(r[ia] or 0)
            t = r[isrc].split()
            if not t:
                continue
            op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
            c[op.split(".")[0]] += n
            tot += n
        md.append("### SASS opcode mix of %s (warp instructions executed: %d)" % (b["name"], tot))
        md.append("")
        md.append("| opcode | share |")
        md.append("|---|---|")
        for op, n in c.most_common(14):
            md.append("| %s | %.1f %% |" % (op, 100.0 * n / max(tot, 1)))
        md.append("")
    open(out + "_summary.md", "w").write("\n".join(md) + "\n")
    json.dump({k: sum(v) / len(v) for k, v in traffic.items()}, open(out + "_traffic.json", "w"), indent=1)
    print("wrote", out + "_summary.md", out + "_traffic.json")


if __name__ == "__main__":
    main()
