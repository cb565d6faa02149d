#!/bin/bash
mkdir -p gpurun_out
echo "=== exact tests"; python -m pytest tests/test_exact_tc_gpu.py tests/test_exact_gpu.py tests/test_config_size_gpu.py tests/test_ivf_gpu.py -q -m gpu -x 2>&1 | tail -12
echo "=== bench headline"; python bench.py --steps 20 --warmup 5 --paths none > gpurun_out/bench_head.log 2> gpurun_out/bench_head.err; echo rc=$?; tail -c 800 gpurun_out/bench_head.err
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/bench_head.log") if x.startswith("{")][-1])
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["config"]["kernel_ms"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"), d["roofline"]["frac"], d["roofline"].get("sustained"), d["parity"]["ids_equal"], d["config"]["uncertified_queries_last_step"])
PY
echo "=== hnsw knobs"; python scripts/dev_hnsw_knobs.py 8 9 2>&1 | tail -3
