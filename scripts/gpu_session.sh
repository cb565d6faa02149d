#!/bin/bash
# One gpurun call = one session: tests (each file under its own timeout), bench, ncu launch list.
# Usage: scripts/gpu_session.sh [tests|bench|ncu|all]...   logs land in gpurun_out/
set -u
mkdir -p gpurun_out
what="${*:-all}"
run() { # name timeout cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name ($(date +%T))"
  timeout -k 10 "$t" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "--- $name rc=$rc"; tail -n 25 "gpurun_out/$name.log"
  return $rc
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.used --format=csv > gpurun_out/smi.log 2>&1
if [[ "$what" == *tests* || "$what" == *all* ]]; then
  run smoke 600 python __graft_entry__.py smoke
  run suite_one_process 1200 python -m pytest tests -q -m gpu -x --timeout 600     # the way the driver runs it
  for f in ${TESTS:-test_exact_tc_gpu test_exact_gpu test_pq_gpu test_hnsw_gpu test_ivf_gpu test_service_gpu test_exchange_gpu}; do
    [ -f tests/$f.py ] && run $f 900 python -m pytest tests/$f.py -q -m gpu -x --timeout 600
  done
fi
if [[ "$what" == *bench* || "$what" == *all* ]]; then
  run bench 900 python bench.py --steps 5 --warmup 3
  run bench_ref 600 python bench.py --impl reference --steps 2 --warmup 1
fi
if [[ "$what" == *ncu* || "$what" == *all* ]]; then
  if python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_plain.log 2>&1; then
    timeout -k 10 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
      --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_run.log 2>&1
    echo "--- ncu launch list rc=$?"; tail -n 5 gpurun_out/launches.csv
  fi
fi
if [[ "$what" == *full* ]]; then
  # one --set full capture of the dominant kernel (KERNEL regex, default the tensor-core scan)
  timeout -k 10 1200 ncu --set full --clock-control none --import-source on -k "regex:${KERNEL:-exact_tc_kernel}" \
    -s ${SKIP:-3} -c 1 -f -o "gpurun_out/${PROF:-prof_exact_tc}" ${CMD:-python bench.py --steps 2 --warmup 3} > gpurun_out/ncu_full.log 2>&1
  echo "--- ncu full rc=$?"; tail -n 5 gpurun_out/ncu_full.log
fi
echo "=== done ($(date +%T))"
