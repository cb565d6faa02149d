#!/bin/bash
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
timeout -k 10 900 $NCU -k regex:adc_coarse_collect -s 1 -c 1 -o gpurun_out/r02_adc_coarse16 python scripts/dev_adc_coarse.py 40000000 1024 16 10 > gpurun_out/ncu_adc16.log 2>&1; echo "rc=$?"
tail -3 gpurun_out/ncu_adc16.log
