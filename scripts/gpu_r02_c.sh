#!/bin/bash
mkdir -p gpurun_out
echo "=== ubench minmax"; ./scripts/ubench/minmax | tail -8
echo "=== ubench gather"; ./scripts/ubench/gather
echo "=== hnsw knobs"; python scripts/dev_hnsw_knobs.py 2>&1 | tail -8
echo "=== ncu tc_post"
timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:tc_post_kernel -s 3 -c 1 -f -o gpurun_out/r02_tc_post \
   python bench.py --steps 3 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_tc_post.log 2>&1; echo "ncu rc=$?"
