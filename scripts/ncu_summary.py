#!/usr/bin/env python
"""Summarise an `ncu --set full` report into profiles/: key counters (JSON) + top stall lines (text).

    python scripts/ncu_summary.py gpurun_out/prof_exact_tc.ncu-rep exact_tc r01 [kernel-name-regex]

With several captured launches (or a regex) the launch with the LONGEST duration among the matches is summarised.
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

rep, name, rnd = sys.argv[1], sys.argv[2], sys.argv[3]
kfilter = ["--kernel-name", "regex:" + sys.argv[4]] if len(sys.argv) > 4 else []
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"] + kfilter, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
_di = hdr.index("gpu__time_duration.sum")
_launches = [r for r in rows[2:] if len(r) == len(hdr)]
_pick = max(range(len(_launches)), key=lambda i: float(_launches[i][_di].replace(",", "") or 0))
vals = _launches[_pick]
_idcol = hdr.index("ID")
kfilter_src = kfilter
want = re.compile(
    r"^(Kernel Name|gpu__time_duration\.sum|dram__bytes_read\.sum|dram__bytes_write\.sum|dram__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"lts__t_bytes\.sum|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
    r"sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_active|sm__pipe_tensor_subpipe_hmma_cycles_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__warps_active\.avg\.per_cycle_active|launch__registers_per_thread|launch__grid_size|launch__block_size|"
    r"launch__shared_mem_per_block_dynamic|smsp__inst_executed\.sum|smsp__issue_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum|sm__cycles_elapsed\.avg|sm__cycles_active\.avg|smsp__cycles_active\.avg|"
    r"sm__inst_executed_pipe_lsu\.avg\.pct_of_peak_sustained_active|sm__inst_executed_pipe_alu\.avg\.pct_of_peak_sustained_active|"
    r"sm__inst_executed_pipe_fma\.avg\.pct_of_peak_sustained_active|l1tex__throughput\.avg\.pct_of_peak_sustained_elapsed)$")
out = {}
for h, u, v in zip(hdr, units, vals):
    if want.match(h):
        try:
            out[h] = {"value": float(v.replace(",", "")), "unit": u}
        except ValueError:
            out[h] = {"value": v, "unit": u}


def to_bytes(e):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return e["value"] * mult.get(e["unit"], 1)


summary = {"kernel": out.get("Kernel Name", {}).get("value"), "report": os.path.basename(rep), "metrics": out}
if "dram__bytes_read.sum" in out:
    summary["dram_bytes_per_launch"] = to_bytes(out["dram__bytes_read.sum"]) + to_bytes(out["dram__bytes_write.sum"])
path = os.path.join(ROOT, "profiles", f"{rnd}_ncu_summary.json")
allsum = json.load(open(path)) if os.path.exists(path) else {}
allsum[name] = summary
json.dump(allsum, open(path, "w"), indent=1, sort_keys=True)

# the source page holds one block per matching launch (a "Kernel Name" row, a header row, the SASS rows): take the picked one
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + kfilter, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
blk = starts[_pick] if _pick < len(starts) else starts[0]
end = starts[starts.index(blk) + 1] if starts.index(blk) + 1 < len(starts) else len(rows)
hdr = rows[blk + 1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[blk + 2:end] if len(r) == len(hdr)]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
agg = {h: sum(int(r[ix[h]] or 0) for r in data) for h in stall_cols}
with open(os.path.join(ROOT, "profiles", f"{rnd}_{name}_stalls.txt"), "w") as f:
    f.write(f"# {summary['kernel']}\n# total warp-state samples {tot}; SASS instructions {len(data)}\n")
    f.write("# stall reasons (all warps, incl. the idle/producer warps parked on barriers):\n")
    for h, v in sorted(agg.items(), key=lambda x: -x[1])[:10]:
        f.write(f"#   {h:28s} {v:8d}  {100.0 * v / max(tot, 1):5.1f}%\n")
    f.write("# top 40 instructions by samples: samples, executed, top stall, SASS\n")
    for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:40]:
        st = max(stall_cols, key=lambda h: int(r[ix[h]] or 0))
        f.write(f"{r[ix['# Samples']]:>8s} {r[ix['Instructions Executed']]:>12s} {st:24s} {r[ix['Source']][:100]}\n")
print(json.dumps({k: v["value"] for k, v in out.items()}, indent=1))
