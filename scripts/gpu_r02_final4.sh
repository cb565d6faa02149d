#!/bin/bash
# r02 closing bench line (after the last ADC collect-target change): PQ tests, default bench, reference arm
mkdir -p gpurun_out
python -m pytest tests/test_pq_gpu.py tests/test_config_size_gpu.py -q -m gpu -x 2>&1 | tail -2
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02k.log 2> gpurun_out/bench_r02k.err; echo rc=$?; tail -c 300 gpurun_out/bench_r02k.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02k_ref.log 2>&1; echo "ref rc=$?"
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/bench_r02k.log") if x.startswith("{")][-1])
p=d.pop("paths",{})
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["config"]["kernel_ms"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"), d["roofline"]["frac"], d["parity"]["ids_equal"])
for k,v in p.items():
    print("==",k,{kk:v.get(kk) for kk in ("value","ms_per_step","kernel_ms","error")}, "frac", (v.get("roofline") or {}).get("frac"), "parity", str(v.get("parity"))[:90])
rt=p["adc_scan"]["recall_target"]
for k,v in rt.items():
    if isinstance(v,list):
        for e in v: print(k[:12], e)
PY
