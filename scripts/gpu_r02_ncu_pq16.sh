#!/bin/bash
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
timeout -k 10 900 $NCU -k regex:pq_encode_tc16_kernel -s 2 -c 1 -o gpurun_out/r02_pq_encode16 python scripts/dev_pq16.py 2000000 48 256 1 rand > gpurun_out/ncu_pq16.log 2>&1; echo "rc=$?"
tail -3 gpurun_out/ncu_pq16.log
