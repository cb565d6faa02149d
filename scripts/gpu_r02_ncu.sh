#!/bin/bash
# r02 evidence: one `ncu --set full` capture per shipped hot kernel (run only after the plain commands exited 0 without ncu)
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
set -x
timeout -k 10 900 $NCU -k regex:'exact_tc_kernel|tc_select_kernel|tc_rerank_bins_kernel|tc_final_kernel' -s 12 -c 4 -o gpurun_out/r02_exact_path \
    python bench.py --steps 3 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_exact.log 2>&1; echo "rc=$?"
timeout -k 10 900 $NCU -k regex:hnsw_search_kernel -s 3 -c 1 -o gpurun_out/r02_hnsw_search \
    python bench.py --workload hnsw_search > gpurun_out/ncu_hnsw.log 2>&1; echo "rc=$?"
timeout -k 10 1200 $NCU -k regex:'pq_encode_tc_kernel|adc_scan_kernel' -s 3 -c 8 -o gpurun_out/r02_pq_path \
    python bench.py --workload adc_scan > gpurun_out/ncu_pq.log 2>&1; echo "rc=$?"
# launch list of the default bench command (per-launch times are cold-cache and serialised: shares, not absolutes)
python bench.py --steps 2 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_plain.log 2>&1 && \
timeout -k 10 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --paths none --no-sustained > gpurun_out/ncu_list.log 2>&1; echo "rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_launches_bench.csv
