#!/bin/bash
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
timeout -k 10 1200 $NCU -k regex:pq_encode_tc_kernel -s 12 -c 1 -o gpurun_out/r02_pq_encode python bench.py --workload pq_encode > gpurun_out/ncu_pq.log 2>&1; echo "rc=$?"
timeout -k 10 1200 $NCU -k regex:adc_scan_kernel -s 4 -c 4 -o gpurun_out/r02_adc_scan python bench.py --workload adc_scan > gpurun_out/ncu_adc.log 2>&1; echo "rc=$?"
echo "=== plain full bench for the record"
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02c.log 2> gpurun_out/bench_r02c.err; echo "rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02c_ref.log 2>&1; echo "rc=$?"
