#!/bin/bash
mkdir -p gpurun_out
echo "=== suite"; python -m pytest tests -q -m gpu -x 2>&1 | tail -12
echo "=== bench headline"; python bench.py --steps 20 --warmup 5 --paths none > gpurun_out/bench_head.log 2> gpurun_out/bench_head.err; echo rc=$?; tail -c 800 gpurun_out/bench_head.err
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/bench_head.log") if x.startswith("{")][-1])
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["config"]["kernel_ms"], d["e2e"]["value"], d["e2e"].get("serial_value"), d["roofline"]["frac"], d["roofline"].get("sustained"), d["parity"])
PY
echo "=== ubench"; ./scripts/ubench/minmax
echo "=== ncu hnsw"
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:hnsw_search_kernel -s 3 -c 1 -f -o gpurun_out/r02_hnsw_search \
   python bench.py --workload hnsw_search > gpurun_out/ncu_hnsw.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_hnsw.log | cut -c1-300
