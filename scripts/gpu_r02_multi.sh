#!/bin/bash
# multi-GPU bench line the way the driver launches it: gpu_r02_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_r02f_${N}gpu.log 2> gpurun_out/bench_r02f_${N}gpu.err; echo rc=$?
tail -c 400 gpurun_out/bench_r02f_${N}gpu.err
python - <<PY
import json
d=json.loads([x for x in open("gpurun_out/bench_r02f_${N}gpu.log") if x.startswith("{")][-1])
p=d.pop("paths",{})
print({k:d.get(k) for k in ("value","queries_per_s_real","ms_per_step","n_gpus")}, "e2e", d["e2e"]["value"], d["parity"])
for k,v in p.items():
    print("==",k,{kk:v.get(kk) for kk in ("value","ms_per_step","error")}, str(v.get("parity"))[:200])
PY
