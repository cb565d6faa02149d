#!/bin/bash
# r02 final evidence run: full GPU suite, default bench (+ reference arm), ncu of the shipped PQ encode kernel, launch list
mkdir -p gpurun_out
echo "=== suite"; python -m pytest tests -q -m gpu -x 2>&1 | tail -4
echo "=== bench"; python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02f.log 2> gpurun_out/bench_r02f.err; echo rc=$?; tail -c 600 gpurun_out/bench_r02f.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02f_ref.log 2>&1; echo "ref rc=$?"
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/bench_r02f.log") if x.startswith("{")][-1])
p=d.pop("paths",{})
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["config"]["kernel_ms"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"), d["e2e"].get("serial_pageable_value"), d["roofline"]["frac"], d["roofline"].get("sustained"), d["parity"]["ids_equal"])
for k,v in p.items():
    print("==",k,{kk:v.get(kk) for kk in ("value","ms_per_step","kernel_ms","error")}, "frac", (v.get("roofline") or {}).get("frac"), "parity", str(v.get("parity"))[:160], v.get("frac_of_pinned_h2d"))
PY
NCU="ncu --set full --clock-control none --import-source on -f"
timeout -k 10 900 $NCU -k regex:pq_encode_tc16_kernel -s 12 -c 1 -o gpurun_out/r02_pq_encode16 python bench.py --workload pq_encode > gpurun_out/ncu_pq16.log 2>&1; echo "ncu pq rc=$?"
python bench.py --steps 2 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_plain.log 2>&1 && \
timeout -k 10 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
