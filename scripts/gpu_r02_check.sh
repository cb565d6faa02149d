#!/bin/bash
# r02 session: the no-alloc test alone in a fresh process, the whole gpu suite the way the driver runs it, staging bench.
mkdir -p gpurun_out
echo "=== no-alloc alone"; python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -8
echo "=== full suite"; python -m pytest tests -q -m gpu -x 2>&1 | tail -15
echo "=== staging bench"
python bench.py --steps 5 --warmup 3 --paths staging --no-sustained > gpurun_out/bench_staging.log 2> gpurun_out/bench_staging.err; echo rc=$?
tail -c 1500 gpurun_out/bench_staging.err
python -c "
import json
d=json.loads([x for x in open('gpurun_out/bench_staging.log') if x.startswith('{')][-1])
print(json.dumps(d['paths'])[:2500])"
