#!/bin/bash
mkdir -p gpurun_out
echo "=== suite"; python -m pytest tests -q -m gpu -x 2>&1 | tail -6
echo "=== bench"; python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02b.log 2> gpurun_out/bench_r02b.err; echo rc=$?; tail -c 800 gpurun_out/bench_r02b.err
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/bench_r02b.log") if x.startswith("{")][-1])
p=d.pop("paths",{})
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["config"]["kernel_ms"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"), d["roofline"]["frac"], d["roofline"].get("sustained"), d["parity"]["ids_equal"])
for k,v in p.items():
    print("==",k,{kk:v.get(kk) for kk in ("value","ms_per_step","kernel_ms","error")}, "frac", (v.get("roofline") or {}).get("frac"), "parity", v.get("parity"), v.get("sweep_gb_per_s"), v.get("frac_of_pinned_h2d"))
PY
