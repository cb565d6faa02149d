#!/usr/bin/env python
"""Benchmark of the search hot path (BASELINE.json metric: QPS @ recall@10 >= 0.95, 1M x 384, k = 10).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload exact|adc|hnsw]

One "step" = one pass of the hot path over one batch of synthetic queries.  Default workload is
BASELINE.json configs[1]: exact brute-force cosine top-10 over 1M x 384 fp32 rows per GPU with a
4096-query batch.  For N > 1 (torchrun, one rank per GPU, NCCL) every rank owns its own 1M-row shard
(contiguous id range), all ranks answer the same query batch, and the per-shard top-10 lists are
all-gathered and merged on the GPU -- weak scaling, `value` = queries/s x shards.

Prints ONE JSON line (rank 0).  `value` is timed with inputs resident in HBM; `e2e` is the same metric
through the host-facing ExactIndex.search() call with host buffers (H2D + D2H inside the timed
region).  `--impl reference` times the CPU port of the reference's path (oracle/, numpy on all host
cores) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "QPS @ recall@10>=0.95, 1Mx384 k=10"
N_ROWS, DIM, N_QUERIES, K = 1_000_000, 384, 4096, 10
SEED_DB, SEED_Q = 1234, 4321


def _ncu_traffic(kernel: str):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` summary
    (profiles/r01_ncu_summary.json: dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    p = os.path.join(ROOT, "profiles", "r01_ncu_summary.json")
    try:
        return json.load(open(p))[kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_burst": d.get("bf16_tflops", 1590.0),
                "bf16_sustained": d.get("bf16_tflops_sustained", 1400.0), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region (NVML every 5 ms; falls back to
    polling nvidia-smi when pynvml is unavailable)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        self.index, self._stop, self._t = index, threading.Event(), None
        self.sm, self.mx, self.reasons, self.power = [], 0.0, set(), 0.0
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = None
            try:
                import torch
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
            except Exception:
                pass
            try:
                self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if isinstance(uuid, str) else uuid)
            except Exception:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
        except Exception:
            self._nvml = None

    def _sample_nvml(self):
        n = self._nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
        try:
            self.power = max(self.power, n.nvmlDeviceGetPowerUsage(self._h) / 1000.0)
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self._h)
            for name, bit in (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                              ("sw_power_cap", 0x4)):
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
        if out.returncode == 0 and out.stdout.strip():
            r = [c.strip() for c in out.stdout.strip().split(",")]
            self.sm.append(float(r[0]))
            self.mx = max(self.mx, float(r[1]))
            for name, v in zip(self.NAMES, r[3:7]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)

    def _run(self):
        while not self._stop.is_set():
            try:
                self._sample_nvml() if self._nvml else self._sample_smi()
            except Exception:
                pass
            self._stop.wait(0.005 if self._nvml else 0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "power_w_max": round(self.power, 1) or None,
                "source": "nvml" if self._nvml else "nvidia-smi"}


def _dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def _cpu_cosine_baseline(X_host, Q_host, seconds_budget=20.0):
    """Oracle port of the reference's path (numpy sgemm + largest-k, all host cores) on a bounded
    sample of the query batch against the full 1M-row shard."""
    from oracle import oracle as O
    xn = np.sqrt(np.einsum("ij,ij->i", X_host, X_host)).astype(np.float32)
    ids = np.arange(X_host.shape[0], dtype=np.int64)
    nq = 8
    t0 = time.perf_counter()
    O.exact_cosine_topk_gemm(X_host, ids, Q_host[:nq], K, xn=xn)
    dt = time.perf_counter() - t0
    nq2 = int(max(8, min(256, nq * (seconds_budget * 0.6) / max(dt, 1e-3))))
    t0 = time.perf_counter()
    O.exact_cosine_topk_gemm(X_host, ids, Q_host[:nq2], K, xn=xn)
    dt = time.perf_counter() - t0
    return {"value": nq2 / dt, "unit": "queries/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"{nq2} of {Q_host.shape[0]} queries x {X_host.shape[0]} rows, numpy sgemm + argpartition "
                      f"(oracle.exact_cosine_topk_gemm), {dt:.2f} s"}


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the reference is pure Python and its
    tree does not exist on the GPU box) on the same metric / config, rank 0 only."""
    rank, _, world = _dist_env()
    if rank != 0:
        return
    # the ours arm is weak scaling: at N GPUs every query is answered over N x 1M rows, so the CPU arm scans that many
    rows = N_ROWS * max(1, int(args.gpus))
    try:                                              # torchrun exports OMP_NUM_THREADS=1: give BLAS the host's cores back
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=os.cpu_count())
    except Exception:                                 # noqa: BLE001
        pass
    X = np.empty((rows, DIM), dtype=np.float32)
    for r in range(max(1, int(args.gpus))):           # the same per-shard seeds as the GPU arm's shards
        X[r * N_ROWS:(r + 1) * N_ROWS] = np.random.default_rng(SEED_DB + r).standard_normal((N_ROWS, DIM), dtype=np.float32)
    Q = np.random.default_rng(SEED_Q).standard_normal((N_QUERIES, DIM), dtype=np.float32)
    from oracle import oracle as O
    xn = np.sqrt(np.einsum("ij,ij->i", X, X)).astype(np.float32)
    ids = np.arange(rows, dtype=np.int64)
    nq = 32                                           # bounded sample per step
    for _ in range(args.warmup):
        O.exact_cosine_topk_gemm(X, ids, Q[:nq], K, xn=xn)
    t0 = time.perf_counter()
    for s in range(args.steps):
        O.exact_cosine_topk_gemm(X, ids, Q[(s * nq) % (N_QUERIES - nq):][:nq], K, xn=xn)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    qps = nq / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"exact cosine top-{K}, {rows} x {DIM} fp32, {nq}-query sample of the "
                               f"{N_QUERIES}-query batch per step (CPU port of the reference path)"},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": os.cpu_count(), "kind": "port",
                         "sample": f"{nq} queries x {rows} rows per step, numpy sgemm + argpartition"},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist

    from vector_db_b200 import _lib, ops, sharded
    from vector_db_b200.indexing import ExactIndex

    rank, local_rank, world = _dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    peaks = _peaks()

    # ---- synthetic shard (random-normal fp32, seeded per rank), resident in HBM
    g = torch.Generator(device=dev).manual_seed(SEED_DB + rank)
    X = torch.randn((N_ROWS, DIM), device=dev, generator=g, dtype=torch.float32)
    gq = torch.Generator(device=dev).manual_seed(SEED_Q)
    Qd = torch.randn((N_QUERIES, DIM), device=dev, generator=gq, dtype=torch.float32)
    index = ExactIndex(DIM, "cosine", device=dev, engine=args.engine)
    index.add_device(torch.arange(rank * N_ROWS, (rank + 1) * N_ROWS, dtype=torch.int64, device=dev), X)
    del X
    X = index.store.vecs[:N_ROWS]
    Q_pin = Qd.cpu().pin_memory()

    # shard exchange: one fused kernel over NVLink peer memory (csrc/exchange.cu); --exchange nccl selects the
    # packed all-gather + merge kernel instead, which is also what a box without peer mapping ends up with
    exchange, peer = "none", None
    if world > 1:
        exchange = args.exchange
        if exchange == "peer":
            try:
                peer = sharded.PeerExchange(N_QUERIES * K, dev)
            except Exception as e:                                  # noqa: BLE001 -- no symmetric memory here
                print(f"bench.py: peer exchange unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
                exchange = "nccl"
        flag = torch.tensor([1 if exchange == "peer" else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)                 # all ranks must take the same path
        if int(flag.item()) == 0:
            exchange, peer = "nccl", None

    def merge(d, i):
        if world == 1:
            return d, i
        if peer is not None:
            return peer.exchange_merge(d, i, descending=True)
        gd, gi = sharded.gather_topk(d, i)                      # one packed NCCL all-gather
        return sharded.merge_shards(gd, gi, descending=True)

    def step_device():
        return merge(*index.search_device(Qd, K))

    def step_e2e():
        q = Q_pin.to(dev, non_blocking=True)
        d, i = merge(*index.search_device(q, K))
        return d.cpu(), i.cpu()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        out = step_device()
    barrier()
    engine = "exact_tc" if ops.tc_supported(N_QUERIES, N_ROWS, DIM, K, "cosine") and args.engine != "f32" else "exact_f32"

    # ---- timed region: device-resident inputs
    _lib.profile_enable(True)
    for slot in ("exact_f32", "exact_tc", "rerank", "finalize"):
        _lib.profile_read(slot)
    launches0 = _lib.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        ev0.record()
        for _ in range(args.steps):
            out = step_device()
        ev1.record()
        barrier()
    ms_total = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_step = float(ms_total.item()) / args.steps
    launches = _lib.launch_count() - launches0
    prof = {s: _lib.profile_read(s) for s in ("exact_f32", "exact_tc", "rerank", "finalize")}
    _lib.profile_enable(False)

    # ---- end to end: pinned host queries in, host results out, every step, through the streaming search API
    # (ExactIndex.search_batches: H2D of batch n+1 and D2H of batch n-1 on copy streams under the kernels of batch n)
    def run_e2e(n):
        last = None
        for last in index.search_batches((Q_pin for _ in range(n)), K, post=merge if world > 1 else None):
            pass
        return last

    run_e2e(2)
    barrier()
    t0 = time.perf_counter()
    res = run_e2e(args.steps)
    torch.cuda.synchronize(dev)
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_step = float(e2e_s.item()) / args.steps
    e2e_serial_t0 = time.perf_counter()                              # the one-call-at-a-time form, for comparison
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize(dev)
    e2e_serial_step = (time.perf_counter() - e2e_serial_t0) / args.steps

    if peer is not None:
        peer.check()                                                # raises if any exchange gave up waiting for a peer
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- recall of the returned ids against the CPU oracle on a sample (untimed), N = 1 only
    cpu_baseline, recall = None, None
    if world == 1:
        from oracle import oracle as O
        X_host, Q_host = X.cpu().numpy(), Qd.cpu().numpy()
        ns = 16
        _, Iref = O.exact_cosine_topk_gemm(X_host, np.arange(N_ROWS, dtype=np.int64), Q_host[:ns], K)
        got = out[1][:ns].cpu().numpy()
        recall = float(np.mean([len(set(got[j]) & set(Iref[j])) / K for j in range(ns)]))
        cpu_baseline = _cpu_cosine_baseline(X_host, Q_host)

    n_launch, k_ms = prof[engine]
    flops = 2.0 * N_QUERIES * N_ROWS * DIM
    kernel_ms = k_ms / max(n_launch, 1)
    achieved = flops / (kernel_ms * 1e-3) / 1e12 if n_launch else None
    peak = peaks["bf16_burst"] if ms_step < 50 else peaks["bf16_sustained"]
    qps = N_QUERIES * world / (ms_step * 1e-3)
    line = {
        "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if engine == "exact_tc" else "f32", "data": "synthetic",
        "config": {
            "workload": f"exact brute-force cosine top-{K}, {N_ROWS} x {DIM} fp32 rows per GPU, {N_QUERIES}-query "
                        f"batch (BASELINE.json configs[1])",
            "engine": engine, "rows_total": N_ROWS * world, "recall_at_10_vs_oracle": recall,
            "uncertified_queries_last_step": int(index.last_uncertified),
            "value_definition": "queries/s x shards: every query is answered over all shards; each GPU scans its "
                                "own 1M rows, per-shard top-10 exchanged and merged on the GPU",
            "shard_exchange": {"none": "single GPU", "peer": "one fused kernel: NVLink peer-memory stores + arrival "
                               "counters + merge (csrc/exchange.cu)",
                               "nccl": "packed NCCL all-gather + merge kernel"}[exchange],
            "l2_policy": "inputs (1.5 GB of rows per GPU) are larger than the 126 MB L2",
            "kernel_ms": {s: (round(v[1] / max(v[0], 1), 4), v[0]) for s, v in prof.items() if v[0]},
        },
        "clocks": clocks.summary(),
        "gpu_launches": int(launches),
        "e2e": {"value": N_QUERIES * world / e2e_step, "unit": "queries/s",
                "h2d_bytes_per_step": int(Q_pin.numel() * 4),
                "d2h_bytes_per_step": int(N_QUERIES * K * (4 + 8)),
                "api": "ExactIndex.search_batches: host batches in, host results out, copies on their own streams "
                       "under the previous / next batch's kernels (depth 2)",
                "serial_value": N_QUERIES * world / e2e_serial_step,
                "serial_api": "one blocking call per batch: pinned H2D, kernels, D2H"},
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": _ncu_traffic(engine),
                     "kernel": engine, "kernel_ms": kernel_ms,
                     "peak_source": peaks["source"] + (" burst (kernel timed in a <50 ms step)" if ms_step < 50 else " sustained"),
                     "algorithmic_flops_per_launch": flops},
    }
    if cpu_baseline is not None:
        line["cpu_baseline"] = cpu_baseline
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--engine", default="auto", choices=["auto", "f32", "tc"])
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N>1: fused peer-memory exchange+merge kernel, or NCCL all-gather + merge kernel")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
