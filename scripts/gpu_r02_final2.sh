#!/bin/bash
# r02 final evidence run (after the 16-bit ADC collect): full GPU suite, default bench (+ reference arm), ncu of the ADC collect kernel
mkdir -p gpurun_out
echo "=== suite"; python -m pytest tests -q -m gpu -x 2>&1 | tail -4
echo "=== bench"; python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02g.log 2> gpurun_out/bench_r02g.err; echo rc=$?; tail -c 600 gpurun_out/bench_r02g.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02g_ref.log 2>&1; echo "ref rc=$?"
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/bench_r02g.log") if x.startswith("{")][-1])
p=d.pop("paths",{})
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["config"]["kernel_ms"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"), d["e2e"].get("serial_pageable_value"), d["roofline"]["frac"], d["roofline"].get("sustained"), d["parity"]["ids_equal"])
for k,v in p.items():
    print("==",k,{kk:v.get(kk) for kk in ("value","ms_per_step","kernel_ms","error")}, "frac", (v.get("roofline") or {}).get("frac"), "parity", str(v.get("parity"))[:160], v.get("frac_of_pinned_h2d"))
PY
NCU="ncu --set full --clock-control none --import-source on -f"
timeout -k 10 900 $NCU -k regex:adc_coarse_collect -s 1 -c 1 -o gpurun_out/r02_adc_coarse python scripts/dev_adc_coarse.py 10000000 1024 48 10 > gpurun_out/ncu_adc.log 2>&1; echo "ncu adc rc=$?"
