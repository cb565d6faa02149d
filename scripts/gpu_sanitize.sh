#!/bin/bash
# compute-sanitizer memcheck + racecheck over the lock-free kernels on small shapes -> gpurun_out/sanitizer_*.log
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for fam in ${FAMILIES:-exact_tc pq_encode_tc exchange hnsw adc static}; do
  for tool in memcheck racecheck; do
    log=gpurun_out/sanitizer_${tool}_${fam}.log
    timeout -k 10 ${SAN_TIMEOUT:-420} $CS --tool $tool --print-limit 20 --kernel-name kns=3vdb --launch-timeout 0 \
        python scripts/sanitize_small.py $fam > $log 2>&1
    echo "--- $tool $fam rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|results ok' $log | tr '\n' ' ')"
  done
done
