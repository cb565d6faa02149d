#!/bin/bash
# r02 closing run: full GPU suite, smoke(), default bench (+ reference arm), launch list of the bench command
mkdir -p gpurun_out
echo "=== suite"; python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "=== smoke"; python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
echo "=== bench"; python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02h.log 2> gpurun_out/bench_r02h.err; echo rc=$?; tail -c 400 gpurun_out/bench_r02h.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02h_ref.log 2>&1; echo "ref rc=$?"
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/bench_r02h.log") if x.startswith("{")][-1])
p=d.pop("paths",{})
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["config"]["kernel_ms"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"), d["e2e"].get("serial_pageable_value"), d["roofline"]["frac"], d["parity"]["ids_equal"], d.get("clocks"))
for k,v in p.items():
    print("==",k,{kk:v.get(kk) for kk in ("value","ms_per_step","kernel_ms","error")}, "frac", (v.get("roofline") or {}).get("frac"), "e2e", (v.get("e2e") or {}).get("value"), "parity", str(v.get("parity"))[:120])
rt=p["adc_scan"]["recall_target"]
for k,v in rt.items():
    if isinstance(v,list):
        for e in v: print(k[:12], e)
PY
python bench.py --steps 2 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_plain.log 2>&1 && \
timeout -k 10 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
