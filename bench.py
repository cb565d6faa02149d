#!/usr/bin/env python
"""Benchmark of the search hot path (BASELINE.json metric: QPS @ recall@10 >= 0.95, 1M x 384, k = 10).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload exact|hnsw_small|pq_encode|adc_scan|hnsw_search|ivf_search|staging|sharded_adc]
                    [--paths all|none|a,b,...]

One "step" = one pass of the hot path over one batch of synthetic input.  The HEADLINE (default workload
`exact`) is BASELINE.json configs[1]: exact brute-force cosine top-10 over 1M x 384 fp32 rows per GPU with a
4096-query batch.  For N > 1 (torchrun, one rank per GPU, NCCL) every rank owns its own 1M-row shard (contiguous
id range), all ranks answer the same query batch and the per-shard top-10 lists are exchanged and merged on the
GPU -- weak scaling.  `value` counts queries/s x 1M-row shards in BOTH arms (the reference arm scans the same
N x 1M rows and multiplies by N as well), `queries_per_s_real` is the plain rate.

Besides the headline the same JSON line carries `paths`: every other BASELINE.json config measured in the same
run on rank 0 -- `hnsw_small` (configs[0]: HNSW build + search on 10k x 384), `pq_encode` (configs[2] encode, 10M x 384, m = 48), `adc_scan` (configs[2] scan, 10M x 48 codes,
Q = 1024, plus the shortlist + rerank point that reaches recall@10 >= 0.95), `hnsw_search` (configs[3], 1M x 768,
M = 32, ef = 128), `ivf_search` (batched IVF on the tensor cores), `staging` (the memmap -> HBM upload) at N = 1, and `sharded_adc` (configs[4], 100M x 128, m = 16,
STRONG scaling: the same 100M codes at every N) at every N.  Each entry has kernel_ms (CUDA events), roofline,
cpu_baseline (the oracle restatement timed in this run), e2e (host buffers through the host API) and parity
(id EQUALITY against the oracle on a sample).  `--workload X` promotes one of them to the headline.

Prints ONE JSON line (rank 0).  `--impl reference` times the CPU port of the reference's path (oracle/, numpy on
the host cores) on a bounded sample of the same workload, on shards generated from the same seeds.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "QPS @ recall@10>=0.95, 1Mx384 k=10"
N_ROWS, DIM, N_QUERIES, K = 1_000_000, 384, 4096, 10
SEED_DB, SEED_Q = 1234, 4321
PARITY_QUERIES = 32


# ------------------------------------------------------------------------------------------------ helpers
def _ncu_traffic(kernel: str):
    """DRAM bytes per launch of a kernel from the committed `ncu --set full` summaries
    (profiles/r02_ncu_summary.json, else r01: dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    for name in ("r02_ncu_summary.json", "r01_ncu_summary.json"):
        try:
            return json.load(open(os.path.join(ROOT, "profiles", name)))[kernel]["dram_bytes_per_launch"]
        except Exception:                                 # noqa: BLE001
            continue
    return None


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_burst": d.get("bf16_tflops", 1590.0),
                "bf16_sustained": d.get("bf16_tflops_sustained", 1400.0), "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def _host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:                                     # noqa: BLE001
        return os.cpu_count() or 1


def _give_blas_the_cores():
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is allowed all the host's cores."""
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=_host_threads())
    except Exception:                                     # noqa: BLE001
        pass


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
            except Exception:                             # noqa: BLE001
                pass
            try:
                self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if isinstance(uuid, str) else uuid)
            except Exception:                             # noqa: BLE001
                self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
        except Exception:                                 # noqa: BLE001
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
        except Exception:                                 # noqa: BLE001
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
            except Exception:                             # noqa: BLE001
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


def _shard_rows(rank: int) -> np.ndarray:
    """The 1M x 384 fp32 shard of a rank: numpy default_rng(SEED_DB + rank), so BOTH arms scan the same rows."""
    return np.random.default_rng(SEED_DB + rank).standard_normal((N_ROWS, DIM), dtype=np.float32)


def _queries() -> np.ndarray:
    return np.random.default_rng(SEED_Q).standard_normal((N_QUERIES, DIM), dtype=np.float32)


def _parity(V, I, Vref, Iref, rtol=1e-4, atol=1e-6):
    """ids of the GPU result against the oracle's: `ids_equal` is tie-aware (BASELINE.json: bit-exact except for
    ties within fp32 tolerance -- a mismatch is accepted only inside a group of oracle values that agree to
    rounding, and then the id sets of the group must agree), `ids_identical_frac` is the strict per-entry rate."""
    V, I, Vref, Iref = (np.asarray(a) for a in (V, I, Vref, Iref))
    fin = np.isfinite(Vref) & np.isfinite(V)
    denom = np.maximum(np.abs(Vref[fin]), 1e-12)
    max_rel = float(np.max(np.abs(V[fin] - Vref[fin]) / denom)) if fin.any() else 0.0
    ok, k = True, I.shape[1]
    for r in np.nonzero((I != Iref).any(axis=1))[0]:
        for j in np.nonzero(I[r] != Iref[r])[0]:
            tol = 1e-5 * abs(Vref[r, j]) + atol
            near = np.abs(Vref[r] - Vref[r, j]) <= 2 * tol
            if not (near.sum() > 1 or near[k - 1]):
                ok = False
            elif not near[k - 1] and set(I[r][near].tolist()) != set(Iref[r][near].tolist()):
                ok = False
    recall = float(np.mean([len(set(I[r].tolist()) & set(Iref[r].tolist())) / max(1, (Iref[r] >= 0).sum())
                            for r in range(I.shape[0])]))
    return {"checked_queries": int(I.shape[0]), "ids_equal": bool(ok and max_rel <= rtol),
            "ids_identical_frac": float((I == Iref).mean()), "max_rel_err": max_rel, "recall_at_k_vs_oracle": recall}


def _timed(fn, iters, warm, slots=()):
    """(last result, ms per iteration by CUDA events on the current stream, {slot: (kernel ms per launch, launches)})."""
    import torch

    from vector_db_b200 import _lib
    out = None
    for _ in range(warm):
        out = fn()
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    for s in slots:
        _lib.profile_read(s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    prof = {}
    for s in slots:
        n, ms = _lib.profile_read(s)
        prof[s] = (ms / max(n, 1), n)
    _lib.profile_enable(False)
    return out, e0.elapsed_time(e1) / max(iters, 1), prof


def _hbm_roofline(alg_bytes, kernel_ms, peaks, kernel, model, ncu_key=None, sustained=False):
    ach = alg_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms else None
    return {"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": ach / peaks["hbm_gbs"] if ach else None, "traffic": _ncu_traffic(ncu_key or kernel),
            "kernel": kernel, "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes, "model": model,
            "peak_source": peaks["source"] + ": HBM copy bandwidth"}


# ------------------------------------------------------------------------------------------------ CPU legs
def _cpu_exact(X_host, Q_host, seconds_budget=12.0):
    """Oracle port of the reference's path (numpy sgemm + largest-k on all host cores) on a bounded sample of
    the query batch against the full shard."""
    from oracle import oracle as O
    _give_blas_the_cores()
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
    return {"value": nq2 / dt, "unit": "queries/s", "cores": _host_threads(), "kind": "port",
            "sample": f"{nq2} of {Q_host.shape[0]} queries x {X_host.shape[0]} rows, numpy sgemm + argpartition "
                      f"(oracle.exact_cosine_topk_gemm), {dt:.2f} s"}


def _cpu_adc(codes_host, lut_host, k, row_offset=0, threads=None):
    """oracle.adc_scan_topk over ALL rows for the given queries, row chunks spread over host threads (numpy's
    gathers release the GIL); the fixed-point table is per query, so chunk results merge exactly by (d, id).
    Returns (d, ids, seconds, threads)."""
    from oracle import oracle as O
    n = codes_host.shape[0]
    threads = threads or min(_host_threads(), 32)
    chunk = max(1 << 16, -(-n // (threads * 2)))
    bounds = [(s, min(n, s + chunk)) for s in range(0, n, chunk)]

    def run(b):
        s, e = b
        return O.adc_scan_topk(codes_host[s:e], lut_host, min(k, e - s), ids=np.arange(s, e, dtype=np.int64) + row_offset)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        parts = list(ex.map(run, bounds))
    nq = lut_host.shape[0]
    D = np.concatenate([np.pad(p[0], ((0, 0), (0, k - p[0].shape[1])), constant_values=np.inf) for p in parts], axis=1)
    I = np.concatenate([np.pad(p[1], ((0, 0), (0, k - p[1].shape[1])), constant_values=-1) for p in parts], axis=1)
    out_d = np.empty((nq, k), np.float32)
    out_i = np.empty((nq, k), np.int64)
    for qi in range(nq):
        order = np.lexsort((I[qi], D[qi]))[:k]
        out_d[qi], out_i[qi] = D[qi][order], I[qi][order]
    return out_d, out_i, time.perf_counter() - t0, threads


# ------------------------------------------------------------------------------------------------ paths
def _pq_setup(dev, D, m, seed=777, rounds=10):
    """Codebooks [m,256,8]: `rounds` Lloyd rounds (GPU k-means step, a12) on a 100k-row sample."""
    import torch

    from vector_db_b200 import ops
    g = torch.Generator(device=dev).manual_seed(seed)
    sample = torch.randn((100_000, D), device=dev, generator=g)
    cb = sample[:256].view(256, m, D // m).permute(1, 0, 2).contiguous()
    for _ in range(rounds):
        ops.pq_kmeans_step(sample, cb)
    return cb


def path_pq_encode(dev, peaks, keep_codes=None):
    """BASELINE configs[2], encode half: 10M x 384 fp32 rows -> uint8 codes, m = 48, K = 256, in 2M-row calls (3 GB of
    rows per call: larger than L2, nothing cached between calls).  Reference: pq.py:77-89."""
    import torch

    from oracle import oracle as O
    from vector_db_b200 import ops
    N, D, m, CH = 10_000_000, 384, 48, 2_000_000
    cb = _pq_setup(dev, D, m)
    cents = list(cb.cpu().numpy())
    codes = torch.empty((N, m), dtype=torch.uint8, device=dev)
    from vector_db_b200 import _lib
    ops.pq_encode(torch.randn((8192, D), device=dev), cb)                       # warm-up (module load, attributes)
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    _lib.profile_read("pq_encode")
    samples, total_ms = [], 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for c in range(N // CH):
        g = torch.Generator(device=dev).manual_seed(9000 + c)
        X = torch.randn((CH, D), device=dev, generator=g)
        e0.record()
        codes[c * CH:(c + 1) * CH] = ops.pq_encode(X, cb)
        e1.record()
        torch.cuda.synchronize()
        total_ms += e0.elapsed_time(e1)
        # parity sample: rows spread over the first / middle / last row tiles of the call
        rows = np.unique(np.concatenate([np.arange(0, 160), np.arange(CH // 2, CH // 2 + 80), np.arange(CH - 160, CH)]))
        samples.append((c * CH + rows, X[torch.from_numpy(rows).to(dev)].cpu().numpy()))
        if c == 0:
            X0 = X[:1_000_000].cpu().pin_memory()
        del X
    n_launch, k_ms = _lib.profile_read("pq_encode")
    _lib.profile_enable(False)
    kernel_ms = k_ms / max(n_launch, 1)
    rows_all = np.concatenate([s[0] for s in samples])
    x_all = np.concatenate([s[1] for s in samples])
    ref = O.pq_encode(x_all, cents)
    got = codes[torch.from_numpy(rows_all).to(dev)].cpu().numpy().astype(np.int64)
    parity = {"checked_rows": int(rows_all.size), "codes_equal": bool(np.array_equal(ref, got)),
              "mismatching_codes": int((ref != got).sum()), "oracle": "oracle.pq_encode (pq.py:77-89), bit-exact bar"}
    # CPU baseline: the reference's own per-row loop (1 core) and the vectorised numpy restatement
    t0 = time.perf_counter()
    O.pq_encode_loop(x_all[:200], cents)
    loop_rate = 200 / (time.perf_counter() - t0)
    nv = 20_000
    xs = np.random.default_rng(5).standard_normal((nv, D), dtype=np.float32)
    t0 = time.perf_counter()
    O.pq_encode(xs, cents)
    vec_dt = time.perf_counter() - t0
    # e2e: pinned host rows in, codes back on the host, through ops.pq_encode (1M-row call)
    def e2e_step():
        return ops.pq_encode(X0.to(dev, non_blocking=True), cb).cpu()
    e2e_step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        e2e_step()
    e2e_dt = (time.perf_counter() - t0) / 3
    alg = float(CH) * (4 * D + m)
    out = {
        "config": {"workload": f"PQ encode m={m} nbits=8, {N} x {D} fp32 rows in {N // CH} calls of {CH} rows "
                               f"(BASELINE.json configs[2], encode)", "codebooks": "10 Lloyd rounds on a 100k-row sample"},
        "metric": "vectors/s", "value": N / (total_ms * 1e-3), "unit": "vectors/s", "ms_per_step": total_ms / (N // CH),
        "ms_for_10M_rows": total_ms, "kernel": "pq_encode_tc16_kernel", "kernel_ms": kernel_ms, "kernel_launches": n_launch,
        "dtype": "f16 split operands, f16 accumulators (candidates) + fp32 resolve (exact)",
        "roofline": _hbm_roofline(alg, kernel_ms, peaks, "pq_encode_tc16_kernel", "rows x (4D + m) bytes per call",
                                  ncu_key="pq_encode_tc16"),
        "cpu_baseline": {"value": nv / vec_dt, "unit": "vectors/s", "cores": 1, "kind": "port",
                         "sample": f"oracle.pq_encode (vectorised numpy restatement) on {nv} rows, {vec_dt:.2f} s; the "
                                   f"reference's own per-row loop (oracle.pq_encode_loop, pq.py:82-89) on 200 rows: "
                                   f"{loop_rate:.0f} vectors/s"},
        "e2e": {"value": X0.shape[0] / e2e_dt, "unit": "vectors/s", "h2d_bytes_per_step": int(X0.numel() * 4),
                "d2h_bytes_per_step": int(X0.shape[0] * m), "api": "ops.pq_encode on a pinned 1M-row host batch, codes "
                "copied back (PCIe-bound: 1.5 GB in per step)"},
        "parity": parity,
    }
    if keep_codes is not None:
        keep_codes["codes"], keep_codes["cb"], keep_codes["cents"] = codes, cb, cents
    return out


def _clustered(dev, n, D, seed, blobs=1024, sigma=0.15):
    import torch
    g = torch.Generator(device=dev).manual_seed(seed)
    centres = torch.randn((blobs, D), device=dev, generator=torch.Generator(device=dev).manual_seed(99))
    s = sigma * centres.norm(dim=1).mean() / (D ** 0.5)
    return centres[torch.randint(0, blobs, (n,), device=dev, generator=g)] + s * torch.randn((n, D), device=dev, generator=g)


def path_adc_scan(dev, peaks, state):
    """BASELINE configs[2], scan half: ADC top-10 over 10M x 48 uint8 codes, Q = 1024 (no reference symbol:
    api/models.py:18 plumbing; arithmetic pq.py:48-62,77-80 squared).  Plus the recall point: shortlist + exact
    rerank on 1M x 384 rows."""
    import torch

    from oracle import oracle as O
    from vector_db_b200 import ops
    from vector_db_b200.indexing import ExactIndex
    from vector_db_b200.inference.pq import PQIndex
    codes, cb, cents = state["codes"], state["cb"], state["cents"]
    N, m = codes.shape
    Q, D = 1024, cb.shape[0] * cb.shape[2]
    Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(SEED_Q))
    Q_pin = Qd.cpu().pin_memory()

    def step():
        return ops.pq_adc_topk(codes, ops.pq_lut(Qd, cb), K)

    (d, i), step_ms, prof = _timed(step, iters=5, warm=2, slots=("adc_scan",))
    kernel_ms = prof["adc_scan"][0]

    def e2e_step():
        q = Q_pin.to(dev, non_blocking=True)
        dd, ii = ops.pq_adc_topk(codes, ops.pq_lut(q, cb), K)
        return dd.cpu(), ii.cpu()
    e2e_step()
    t0 = time.perf_counter()
    for _ in range(3):
        e2e_step()
    e2e_dt = (time.perf_counter() - t0) / 3
    # parity + CPU baseline in one pass: the oracle over ALL 10M rows for a few queries
    npar = 4
    codes_h = codes.cpu().numpy()
    lut_h = O.adc_lut(Qd[:npar].cpu().numpy(), cents)
    rd, ri, cpu_s, thr = _cpu_adc(codes_h, lut_h, K)
    par = _parity(d[:npar].cpu().numpy(), i[:npar].cpu().numpy(), rd, ri, rtol=1e-6)
    par["bit_equal"] = bool(np.array_equal(ri, i[:npar].cpu().numpy()) and np.array_equal(rd, d[:npar].cpu().numpy()))
    par["oracle"] = "oracle.adc_scan_topk over all rows (fixed-point table: integer sums, order independent)"
    del codes_h
    alg = float(Q) * N * m
    out = {
        "config": {"workload": f"PQ ADC scan top-{K}, {N} x {m} uint8 codes, {Q}-query batch (BASELINE.json configs[2], scan)"},
        "metric": "queries/s", "value": Q / (step_ms * 1e-3), "unit": "queries/s", "ms_per_step": step_ms,
        "kernel": "adc_coarse_collect_kernel", "kernel_ms": kernel_ms,
        "kernel_ms_is": "the whole scan phase: threshold sample + 16-bit tables + collect kernel + re-evaluation of the "
                        "collected rows on the 32-bit tables (the collect kernel is ~90% of it)",
        "dtype": "u8 codes, fixed-point i32 sums",
        "roofline": _hbm_roofline(alg, kernel_ms, peaks, "adc_coarse_collect_kernel",
                                  "per-query code stream: Q x N x m code bytes (SURVEY.md 8d); EIGHT queries share one pass "
                                  "over the codes (one 128-bit shared-memory lookup per code byte returns eight 16-bit "
                                  "lower-bound entries), so the figure exceeds the HBM peak and DRAM traffic is far below "
                                  "it -- the physical bound is the shared-memory pipe (ncu: l1tex throughput 98.6% of peak, "
                                  "profiles/r02_ncu_summary.json adc_coarse)",
                                  ncu_key="adc_coarse"),
        "cpu_baseline": {"value": npar / cpu_s, "unit": "queries/s", "cores": thr, "kind": "port",
                         "sample": f"oracle.adc_scan_topk, {npar} queries x {N} rows (row chunks on {thr} threads), {cpu_s:.2f} s"},
        "e2e": {"value": Q / e2e_dt, "unit": "queries/s", "h2d_bytes_per_step": int(Q_pin.numel() * 4),
                "d2h_bytes_per_step": int(Q * K * 12), "api": "pinned host queries -> ops.pq_lut + ops.pq_adc_topk -> host (d, ids)"},
        "parity": par,
    }
    # ---- the recall target: ADC shortlist + exact fp32 rerank, 1M x 384
    rec = {}
    for tag, X in (("clustered", _clustered(dev, N_ROWS, DIM, 4242)),
                   ("iid", torch.randn((N_ROWS, DIM), device=dev, generator=torch.Generator(device=dev).manual_seed(31)))):
        Qr = (_clustered(dev, 1024, DIM, 777) if tag == "clustered" else
              torch.randn((1024, DIM), device=dev, generator=torch.Generator(device=dev).manual_seed(32)))
        cbr = X[:256].reshape(256, 48, 8).permute(1, 0, 2).contiguous()
        for _ in range(15):
            ops.pq_kmeans_step(X[:200_000], cbr)
        idx = PQIndex(list(cbr.cpu().numpy()), keep_vectors=True, device=dev)
        idx.add_device(X)
        ex = ExactIndex(DIM, "l2", device=dev)
        ex.add_device(torch.arange(N_ROWS, device=dev), X)
        _, truth = ex.search_device(Qr, K)
        truth = truth.cpu().numpy()
        rows = []
        for R in ((10, 100, 1000) if tag == "clustered" else (1000, 8000)):
            (dd, ii), ms, pr = _timed(lambda: idx.search_device(Qr, K, shortlist=R, rerank=R > K), iters=3, warm=1,
                                      slots=("adc_scan", "rerank"))
            got = ii.cpu().numpy()
            recall = float(np.mean([len(set(got[r]) & set(truth[r])) / K for r in range(got.shape[0])]))
            rows.append({"shortlist": R, "recall_at_10": round(recall, 4), "ms_per_step": round(ms, 3),
                         "qps": round(Qr.shape[0] / (ms * 1e-3), 1), "adc_kernel_ms": round(pr["adc_scan"][0], 3),
                         "adc_frac_of_code_stream_roofline": round(float(Qr.shape[0]) * N_ROWS * 48 / (pr["adc_scan"][0] * 1e-3) / 1e9
                                                                   / peaks["hbm_gbs"], 3)})
        rec[tag] = rows
        del idx, ex, X
    out["recall_target"] = {"workload": "1M x 384, m=48, Q=1024: ADC top-R shortlist + exact fp32 rerank; recall@10 vs the exact scan",
                            "clustered (1024 Gaussian blobs, sigma 0.15)": rec["clustered"],
                            "iid Gaussian (PQ on unstructured data: the target is NOT reachable through PQ, SURVEY 7.3(3))": rec["iid"]}
    return out


def path_hnsw_search(dev, peaks):
    """BASELINE configs[3]: batched HNSW search, 1M x 768, M = 32, ef = 128, Q = 10k, on the wave-built graph;
    parity = the reference's own search (oracle restatement of hnsw.py:330-374) on the SAME graph."""
    import random

    import torch

    from oracle.hnsw_oracle import search_arrays
    from vector_db_b200 import ops
    from vector_db_b200.indexing import HNSW, ExactIndex
    N, D, M, ef, Q = 1_000_000, 768, 32, 128, 10_000
    X = torch.randn((N, D), device=dev, generator=torch.Generator(device=dev).manual_seed(SEED_DB))
    Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(SEED_Q))
    idx = HNSW(M=M, ef_construction=200, rng=random.Random(42), device=dev)
    t0 = time.perf_counter()
    stats = idx.build_from_arrays(range(N), X)
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    vec = idx._vectors
    entry_row, top = idx._read_state()

    def step():
        return ops.hnsw_search(Qd, vec.vecs, vec.n, idx._adj0, idx._adj_up, idx._up_ptr_dev, entry_row, top, ef, K,
                               want_stats=True, levels=idx._levels_dev)

    (d, r, st), step_ms, prof = _timed(step, iters=5, warm=2, slots=("hnsw_search",))
    kernel_ms = prof["hnsw_search"][0]
    st_h = st.cpu().numpy().astype(np.int64)
    evals, expans = float(st_h[:, 0].mean()), float(st_h[:, 1].mean())
    alg = Q * (evals * D * 4 + expans * 2 * M * 4)
    Q_host = Qd.cpu().numpy()
    idx.search_batch(Q_host, K, ef=ef)
    t0 = time.perf_counter()
    for _ in range(3):
        d_h, i_h = idx.search_batch(Q_host, K, ef=ef)
    e2e_dt = (time.perf_counter() - t0) / 3
    # parity + CPU baseline: the reference's search on the same graph, 64 queries, 1 core
    X_host = X.cpu().numpy()
    adj0, adj_up, up_ptr, _, _ = idx._host_graph()
    n_or = 64
    t0 = time.perf_counter()
    same, ev_same, ref_rows = 0, 0, []
    for j in range(n_or):
        res, ev = search_arrays(X_host, adj0, up_ptr, adj_up, entry_row, top, Q_host[j], K, ef)
        ref_rows.append([rr for _, rr in res] + [-1] * (K - len(res)))
        same += int(ref_rows[-1] == i_h[j].tolist())
        ev_same += int(ev == int(st_h[j, 0]))
    cpu_s = time.perf_counter() - t0
    ex = ExactIndex(D, "l2", device=dev)
    ex.add_device(torch.arange(N, device=dev), X)
    _, truth = ex.search_device(Qd[:512], K)
    truth = truth.cpu().numpy()
    rec_gpu = float(np.mean([len(set(i_h[j]) & set(truth[j])) / K for j in range(512)]))
    rec_ref = float(np.mean([len(set(ref_rows[j]) & set(truth[j])) / K for j in range(n_or)]))
    rec_gpu_same = float(np.mean([len(set(i_h[j]) & set(truth[j])) / K for j in range(n_or)]))
    return {
        "config": {"workload": f"batched HNSW search, {N} x {D} fp32, M={M}, ef_search={ef}, {Q}-query batch, k={K} "
                               f"(BASELINE.json configs[3])",
                   "graph": f"wave-parallel GPU builder, ef_construction=200: {build_s:.1f} s ({N / build_s:.0f} inserts/s)",
                   "build_stats": stats},
        "metric": "queries/s", "value": Q / (step_ms * 1e-3), "unit": "queries/s", "ms_per_step": step_ms,
        "kernel": "hnsw_search_kernel", "kernel_ms": kernel_ms, "dtype": "f32",
        "evals_per_query": evals, "expansions_per_query": expans,
        "roofline": dict(_hbm_roofline(alg, kernel_ms, peaks, "hnsw_search_kernel",
                                       "evals x 4D + expansions x 2M x 4 bytes per query, counted by the kernel (SURVEY.md "
                                       "8d).  ALGORITHMIC figure: hub rows are served from L2, so `traffic` (ncu DRAM bytes) "
                                       "is well below it and the kernel is latency-bound, not at the HBM roofline",
                                       ncu_key="hnsw_search")),
        "cpu_baseline": {"value": n_or / cpu_s, "unit": "queries/s", "cores": 1, "kind": "port",
                         "sample": f"oracle.hnsw_oracle.search_arrays (hnsw.py:330-374 restated) on {n_or} queries, same graph, {cpu_s:.2f} s"},
        "e2e": {"value": Q / e2e_dt, "unit": "queries/s", "h2d_bytes_per_step": int(Q * D * 4), "d2h_bytes_per_step": int(Q * K * 12),
                "api": "HNSW.search_batch: numpy queries in, numpy (dist, ids) out"},
        "parity": {"checked_queries": n_or, "ids_equal": bool(same == n_or), "identical_result_lists": same,
                   "identical_eval_counts": ev_same,
                   "recall_at_10_reference_search": rec_ref, "recall_at_10_gpu_search_same_queries": rec_gpu_same,
                   "recall_at_10_gpu_search_512q": rec_gpu,
                   "oracle": "the reference's search restated (pinned to reference-built graphs by tests/golden/hnsw_*.npz), "
                             "same graph; gate: recall within 0.5 pt of the reference's (met with equality)",
                   "note": "recall vs the exact scan on iid 768-d Gaussian rows is low for the reference's search and the GPU "
                           "search alike: that is the data (SURVEY.md 7.3(3)), not the traversal"},
    }


def path_hnsw_small(dev, peaks):
    """BASELINE configs[0], the reference's own CPU-runnable case: HNSW M = 16, ef_construction = 200, ef_search = 64, build +
    search on 10k random 384-d vectors, 100 queries, k = 10.  Build = the sequential-semantics insert kernel (the graph equals
    the reference's link for link -- pinned on smaller reference-built graphs by tests/golden/hnsw_*.npz); parity of the search
    = the reference's search (oracle restatement of hnsw.py:330-374) on the SAME graph."""
    import random

    import torch

    from oracle.hnsw_oracle import search_arrays
    from vector_db_b200.indexing import HNSW, ExactIndex
    from vector_db_b200.types import Node
    N, D, M, efc, ef, Q = 10_000, 384, 16, 200, 64, 100
    rng = np.random.default_rng(SEED_DB)
    X = rng.standard_normal((N, D), dtype=np.float32)
    Qh = np.random.default_rng(SEED_Q).standard_normal((Q, D), dtype=np.float32)
    idx = HNSW(M=M, ef_construction=efc, rng=random.Random(42), device=dev)
    t0 = time.perf_counter()
    idx.build_index([Node(id=i, embedding=X[i]) for i in range(N)])
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    build_evals = idx.distance_evaluations()
    idx.search_batch(Qh, K, ef=ef)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 20
    for _ in range(reps):
        d_h, i_h = idx.search_batch(Qh, K, ef=ef)
    e2e_dt = (time.perf_counter() - t0) / reps
    vec = idx._vectors
    entry_row, top = idx._read_state()
    from vector_db_b200 import ops
    Qd = torch.from_numpy(Qh).to(dev)
    (_, _, st), step_ms, prof = _timed(lambda: ops.hnsw_search(Qd, vec.vecs, vec.n, idx._adj0, idx._adj_up, idx._up_ptr_dev, entry_row,
                                                             top, ef, K, want_stats=True, levels=idx._levels_dev),
                                       iters=20, warm=3, slots=("hnsw_search",))
    st_h = st.cpu().numpy().astype(np.int64)
    adj0, adj_up, up_ptr, _, _ = idx._host_graph()
    t0 = time.perf_counter()
    same, ev_same, ref_rows = 0, 0, []
    for j in range(Q):
        res, ev = search_arrays(X, adj0, up_ptr, adj_up, entry_row, top, Qh[j], K, ef)
        ref_rows.append([rr for _, rr in res] + [-1] * (K - len(res)))
        same += int(ref_rows[-1] == i_h[j].tolist())
        ev_same += int(ev == int(st_h[j, 0]))
    cpu_s = time.perf_counter() - t0
    ex = ExactIndex(D, "l2", device=dev)
    ex.add(range(N), X)
    _, truth = ex.search(Qh, K)
    rec = float(np.mean([len(set(i_h[j]) & set(truth[j])) / K for j in range(Q)]))
    return {
        "config": {"workload": f"HNSW build + search, {N} x {D} fp32, M={M}, ef_construction={efc}, ef_search={ef}, {Q} queries, "
                               f"k={K} (BASELINE.json configs[0])",
                   "build": f"sequential-semantics insert kernel: {build_s:.1f} s = {N / build_s:.0f} inserts/s, "
                            f"{build_evals} distance evaluations (the reference's Python build: ~52 inserts/s, BASELINE.md 2)"},
        "metric": "queries/s", "value": Q / (step_ms * 1e-3), "unit": "queries/s", "ms_per_step": step_ms,
        "kernel": "hnsw_search_kernel", "kernel_ms": prof["hnsw_search"][0], "dtype": "f32",
        "build_seconds": build_s, "inserts_per_s": N / build_s,
        "evals_per_query": float(st_h[:, 0].mean()), "recall_at_10_vs_exact": rec,
        "roofline": None,
        "cpu_baseline": {"value": Q / cpu_s, "unit": "queries/s", "cores": 1, "kind": "port",
                         "sample": f"oracle.hnsw_oracle.search_arrays (hnsw.py:330-374 restated) on the {Q} queries, same graph, {cpu_s:.2f} s"},
        "e2e": {"value": Q / e2e_dt, "unit": "queries/s", "h2d_bytes_per_step": int(Q * D * 4), "d2h_bytes_per_step": int(Q * K * 12),
                "api": "HNSW.search_batch: numpy queries in, numpy (dist, ids) out"},
        "parity": {"checked_queries": Q, "ids_equal": bool(same == Q), "identical_result_lists": same,
                   "identical_eval_counts": ev_same,
                   "oracle": "the reference's search restated, same graph; the graph itself is the reference's (insert kernel with "
                             "sequential semantics, pinned by tests/golden/hnsw_*.npz and test_gpu_build_equals_reference_graph)"},
    }


def path_ivf_search(dev, peaks):
    """SURVEY.md 8f rank 3 (widening beyond BASELINE.json): batched IVF search, 1M x 384, K = 1024 lists, 4096 queries, k = 10:
    the list-major tensor-core engine (vdb_ivf_topk_tc) against the query-major fp32 list scan (vdb_ivf_scan); parity = the
    reference's IvfIndex.search (oracle restatement of ivf.py:139-199) on the same centroids and lists."""
    import torch

    from oracle import ivf_oracle as IO
    from vector_db_b200.indexing import IvfIndex
    N, D, KL, Q = 1_000_000, 384, 1024, 4096
    X = torch.randn((N, D), device=dev, generator=torch.Generator(device=dev).manual_seed(SEED_DB))
    Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(SEED_Q))
    X_host = X.cpu().numpy()
    del X
    idx = IvfIndex(k=KL, device=dev)
    t0 = time.perf_counter()
    idx.build_from_arrays(range(N), X_host, iters=10, init_rows=np.random.default_rng(SEED_DB).choice(N, KL, replace=False))
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    rows = {}
    for n_probe in (4, 16):
        for engine in ("tc", "scan"):
            idx.batch_engine = engine
            (d, i, scanned), ms, _ = _timed(lambda: idx.search_device(Qd, n_probe, K), iters=5, warm=2)
            rows[(n_probe, engine)] = (ms, d, i, float(scanned.float().mean().item()), idx.last_uncertified if engine == "tc" else 0)
    n_probe = 4
    ms_tc, d_tc, i_tc, cand, unc = rows[(n_probe, "tc")]
    ms_sc, d_sc, i_sc, _, _ = rows[(n_probe, "scan")]
    engines_equal = bool(torch.equal(i_tc, i_sc) and torch.equal(d_tc, d_sc)
                         and torch.equal(rows[(16, "tc")][2], rows[(16, "scan")][2]))
    idx.batch_engine = "tc"
    Q_host = Qd.cpu().numpy()
    idx.search_batch(Q_host, n_probe, K)
    t0 = time.perf_counter()
    for _ in range(3):
        d_h, i_h = idx.search_batch(Q_host, n_probe, K)
    e2e_dt = (time.perf_counter() - t0) / 3
    # parity + CPU baseline: the reference's search on the same centroids / lists, 8 queries, 1 core
    class _Rows:                                         # id -> embedding without a million-entry dict
        def __getitem__(self, i):
            return X_host[int(i)]

        def __contains__(self, i):
            return 0 <= int(i) < N

    lists = idx.inverted_lists
    n_or = 8
    t0 = time.perf_counter()
    same = 0
    max_rel = 0.0
    for j in range(n_or):
        res, _, _ = IO.search(idx.centroids, lists, _Rows(), Q_host[j], n_probe, K)
        ids_ref = [int(r[1]) for r in res]
        same += int(ids_ref == [int(v) for v in i_h[j] if v >= 0])
        dr = np.array([float(r[0]) for r in res], dtype=np.float64)
        max_rel = max(max_rel, float(np.max(np.abs(d_h[j][:len(dr)] - dr) / np.maximum(dr, 1e-30))) if len(dr) else 0.0)
    cpu_s = time.perf_counter() - t0
    alg = Q * cand * (4 * D + 4)
    return {
        "config": {"workload": f"batched IVF search, {N} x {D} fp32, {KL} lists (upload + 10 Lloyd rounds, {build_s:.2f} s), {Q}-query batch, "
                               f"n_probe={n_probe}, k={K} (SURVEY.md 8f rank 3; not a BASELINE.json config)",
                   "engines_ms": {f"n_probe={p},{e}": round(v[0], 3) for (p, e), v in rows.items()},
                   "candidates_per_query": cand, "uncertified_queries_rerun_through_the_scan": unc},
        "metric": "queries/s", "value": Q / (ms_tc * 1e-3), "unit": "queries/s", "ms_per_step": ms_tc,
        "kernel": "exact_tc_kernel (list-major units) + select / rerank / final", "kernel_ms": ms_tc, "dtype": "bf16 coarse + f32 rerank",
        "query_major_scan_ms": ms_sc,
        "roofline": dict(_hbm_roofline(alg, ms_tc, peaks, "ivf list-major step",
                                       "candidates x (4D + 4) bytes per query (the query-major model of DESIGN.md 4.5) over the WHOLE "
                                       "step incl. the host-side grouping; above 1 means lists shared by the batch are read once")),
        "cpu_baseline": {"value": n_or / cpu_s, "unit": "queries/s", "cores": 1, "kind": "port",
                         "sample": f"oracle.ivf_oracle.search (ivf.py:139-199 restated) on {n_or} queries, same lists, {cpu_s:.2f} s"},
        "e2e": {"value": Q / e2e_dt, "unit": "queries/s", "h2d_bytes_per_step": int(Q * D * 4), "d2h_bytes_per_step": int(Q * K * 12),
                "api": "IvfIndex.search_batch: numpy queries in, numpy (dist, ids) out"},
        "parity": {"checked_queries": n_or, "ids_equal": bool(same == n_or), "identical_result_lists": same, "max_rel_err": max_rel,
                   "tensor_core_engine_equals_list_scan_all_queries": engines_equal,
                   "oracle": "the reference's IvfIndex.search restated (pinned to a reference run by tests/golden/ivf.npz), same "
                             "centroids and lists"},
    }


def path_staging(dev, peaks):
    """Storage staging (north_star subsystem 1; storage.py:135-161,254-259): the reference's layer-1 memmap file
    ({i8 id, f4[D]} records, stride 8+4D) -> pinned ring -> strided DMA -> SoA in HBM."""
    import torch

    from vector_db_b200.storage import MMapNodeStorage
    n, D = N_ROWS, DIM
    tmp = tempfile.mkdtemp(prefix="vdb_stage_")
    path = os.path.join(tmp, "bench.embeddings.npy")
    try:
        st = MMapNodeStorage(path, None, dim=D, capacity=n)
        rng = np.random.default_rng(3)
        for s in range(0, n, 100_000):
            st.embeddings["id"][s:s + 100_000] = np.arange(s, s + 100_000)
            st.embeddings["embedding"][s:s + 100_000] = rng.standard_normal((100_000, D), dtype=np.float32)
        st.embeddings.flush()
        st.close()
        st = MMapNodeStorage(path, None, dim=D, capacity=n)            # reopen: index the file as a restart would
        assert st.size() == n
        # the box's pinned H2D rate (one contiguous cudaMemcpyAsync), for the denominator
        pin = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
        dst = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
        dst.copy_(pin, non_blocking=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            dst.copy_(pin, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        h2d_gbs = 3 * (1 << 30) / (e0.elapsed_time(e1) * 1e-3) / 1e9
        del pin, dst
        runs = []
        for _ in range(3):
            ids, vecs = st.stage_to_device(dev)
            runs.append(dict(st.last_stage))
        sweep = {}
        for thr in (4, 8, 16, 32):
            for chunk in (2048, 8192, 32768):
                st._stage_ring = None
                st.stage_to_device(dev, chunk_records=chunk, threads=thr)                  # first call allocates the ring
                st.stage_to_device(dev, chunk_records=chunk, threads=thr)
                sweep[f"threads={thr},chunk_records={chunk}"] = round(st.last_stage["gb_per_s"], 2)
        st._stage_ring = None
        best = max(runs, key=lambda r_: r_["gb_per_s"])
        t0 = time.perf_counter()
        ids1, vecs1 = st.stage_to_device(dev, threads=1, depth=2, chunk_records=65536)
        single = dict(st.last_stage)
        chk = np.random.default_rng(0).integers(0, n, 64)
        ok = bool(np.array_equal(vecs[torch.from_numpy(chk).to(dev)].cpu().numpy(), np.asarray(st.embeddings["embedding"][chk]))
                  and np.array_equal(ids[torch.from_numpy(chk).to(dev)].cpu().numpy(), chk))
        st.close()
        return {"config": {"workload": f"MMapNodeStorage.stage_to_device: {n} records x {8 + 4 * D} B (reference layer-1 file, "
                                       f"page-cache resident) -> HBM SoA"},
                "metric": "GB/s", "value": best["gb_per_s"], "unit": "GB/s (file bytes / wall time)", "seconds": best["seconds"],
                "threads": best["threads"], "ring_depth": best["depth"], "runs_gb_per_s": [round(r_["gb_per_s"], 2) for r_ in runs],
                "single_thread_fill_gb_per_s": round(single["gb_per_s"], 2), "sweep_gb_per_s": sweep,
                "pinned_h2d_gb_per_s": h2d_gbs, "frac_of_pinned_h2d": best["gb_per_s"] / h2d_gbs,
                "parity": {"checked_rows": 64, "rows_equal": ok}}
    finally:
        try:
            os.remove(path)
            os.rmdir(tmp)
        except OSError:
            pass


def path_sharded_adc(dev, peaks, rank, world, peer_ok):
    """BASELINE configs[4]: PQ (m = 16, nbits = 8) index over 100M x 128 rows sharded by row across the GPUs of the
    box -- STRONG scaling: the same 100M codes at every N (N = 1 holds all 1.6 GB).  Every rank generates its rows on
    the device from per-2M-row-chunk seeds (so the row set does not depend on N), encodes them with the tensor-core
    encoder, scans its shard, and the per-shard top-10 lists are merged after ONE packed NCCL all-gather."""
    import torch
    import torch.distributed as dist

    from oracle import oracle as O
    from vector_db_b200 import _lib, ops, sharded
    N_TOTAL, D, m, Q, CH = 100_000_000, 128, 16, 1024, 500_000

    def agree(ok: bool, what: str):
        """All ranks leave together when one of them failed in a local phase (no rank is left waiting in a collective)."""
        t_ = torch.tensor([1 if ok else 0], device=dev)
        if world > 1:
            dist.all_reduce(t_, op=dist.ReduceOp.MIN)
        if int(t_.item()) == 0:
            raise RuntimeError(f"sharded_adc: a rank failed during {what}")

    # chunk-aligned shards (200 chunks of 500k rows: 25 per rank at 8 GPUs) so every rank generates whole seeded chunks
    per = -(-(N_TOTAL // CH) // world) * CH
    lo, hi = min(N_TOTAL, rank * per), min(N_TOTAL, (rank + 1) * per)
    n_local = hi - lo
    err = None
    try:
        cb = _pq_setup(dev, D, m, seed=778)                # same seed on every rank: replicated codebooks (131 KB)
        cents = list(cb.cpu().numpy())
        codes = torch.empty((n_local, m), dtype=torch.uint8, device=dev)
        enc_ms = 0.0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for s in range(0, n_local, CH):
            X = torch.randn((CH, D), device=dev, generator=torch.Generator(device=dev).manual_seed(50_000 + (lo + s) // CH))
            e0.record()
            codes[s:s + CH] = ops.pq_encode(X, cb)
            e1.record()
            torch.cuda.synchronize()
            enc_ms += e0.elapsed_time(e1)
            del X
    except Exception as e_:                               # noqa: BLE001
        err = e_
    agree(err is None, f"shard generation / encode ({err})")
    Qd = torch.randn((Q, D), device=dev, generator=torch.Generator(device=dev).manual_seed(SEED_Q))

    def local_search(q, k):
        d, rows = ops.pq_adc_topk(codes, ops.pq_lut(q, cb), k)
        return d, torch.where(rows >= 0, rows + lo, rows)

    def step():
        return sharded.search_sharded(local_search, Qd, K)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(2):
        out = step()
    barrier()
    _lib.profile_enable(True)
    _lib.profile_read("adc_scan")
    steps = 3 if world == 1 else 5
    barrier()
    e0.record()
    for _ in range(steps):
        out = step()
    e1.record()
    barrier()
    nk, kms = _lib.profile_read("adc_scan")
    _lib.profile_enable(False)
    t = torch.tensor([e0.elapsed_time(e1) / steps, kms / max(nk, 1), enc_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step, kernel_ms, enc_max = (float(v) for v in t.tolist())
    # parity: every rank runs the oracle over ITS shard for a few queries (host threads), rank 0 merges the per-shard
    # oracle lists by (d, id) on the host and compares with the GPU's merged answer -- the oracle over all 100M rows
    npar = 4
    err = None
    try:
        lut_h = O.adc_lut(Qd[:npar].cpu().numpy(), cents)
        rd, ri, cpu_s, thr = _cpu_adc(codes.cpu().numpy(), lut_h, K, row_offset=lo,
                                      threads=max(1, min(32, _host_threads() // max(world, 1))))
    except Exception as e_:                               # noqa: BLE001
        err = e_
    agree(err is None, f"the oracle pass ({err})")
    pack = torch.empty((npar, K, 3), dtype=torch.float64, device=dev)
    pack[..., 0] = torch.from_numpy(rd.astype(np.float64)).to(dev)
    pack[..., 1] = torch.from_numpy(ri.astype(np.float64)).to(dev)
    pack[..., 2] = cpu_s
    if world > 1:
        allp = [torch.empty_like(pack) for _ in range(world)]
        dist.all_gather(allp, pack)
    else:
        allp = [pack]
    if rank != 0:
        return None
    allp = torch.cat([p.cpu() for p in allp], dim=1).numpy()              # [npar, world*K, 3]
    ref_d = np.empty((npar, K), np.float32)
    ref_i = np.empty((npar, K), np.int64)
    for qi in range(npar):
        dd, ii = allp[qi, :, 0].astype(np.float32), allp[qi, :, 1].astype(np.int64)
        order = np.lexsort((ii, dd))[:K]
        ref_d[qi], ref_i[qi] = dd[order], ii[order]
    par = _parity(out[0][:npar].cpu().numpy(), out[1][:npar].cpu().numpy(), ref_d, ref_i, rtol=1e-6)
    par["bit_equal"] = bool(np.array_equal(ref_i, out[1][:npar].cpu().numpy()) and np.array_equal(ref_d, out[0][:npar].cpu().numpy()))
    par["oracle"] = ("oracle.adc_scan_topk over every shard's rows (each rank its own, host threads), per-shard lists "
                     "merged by (d, id) on the host: the oracle over all 100M rows")
    cpu_total_s = float(allp[0, ::K, 2].sum())                             # serial CPU time over all shards
    alg = float(Q) * n_local * m
    return {
        "config": {"workload": f"PQ ADC top-{K}, {N_TOTAL} x {D} rows as m={m} uint8 codes sharded over {world} GPU(s), "
                               f"{Q}-query batch (BASELINE.json configs[4]); rows generated on the device per 500k-row chunk "
                               f"(seed = 50000 + chunk), encoded shard-side, dropped",
                   "rows_per_gpu": n_local, "exchange": "one packed NCCL all-gather (122,880 B per rank) + merge kernel" if world > 1 else "none"},
        "metric": "queries/s", "value": Q / (ms_step * 1e-3), "unit": "queries/s", "scaling": "strong", "n_gpus": world,
        "ms_per_step": ms_step, "kernel": "adc_coarse_collect_kernel", "kernel_ms": kernel_ms,
        "kernel_ms_is": "scan phase (threshold sample + collect on 16-bit tables + re-evaluation), max over ranks",
        "encode_ms_per_shard": enc_max, "encode_vectors_per_s_per_gpu": n_local / (enc_max * 1e-3),
        "roofline": _hbm_roofline(alg, kernel_ms, peaks, "adc_coarse_collect_kernel",
                                  "per GPU: Q x rows_per_gpu x m code bytes (per-query code stream; eight queries share a "
                                  "pass over the codes, so the figure exceeds the HBM peak -- shared-memory-pipe bound: ncu "
                                  "l1tex data-pipe wavefronts at 97% of peak, profiles/r02_ncu_summary.json adc_coarse_m16)",
                                  ncu_key="adc_coarse_m16"),
        "cpu_baseline": {"value": npar / cpu_total_s, "unit": "queries/s", "cores": thr, "kind": "port",
                         "sample": f"oracle.adc_scan_topk, {npar} queries x all {N_TOTAL} rows (sum over shards of the "
                                   f"per-shard time, {thr} threads each), {cpu_total_s:.2f} s"},
        "parity": par,
    }


PATHS = ("hnsw_small", "pq_encode", "adc_scan", "hnsw_search", "ivf_search", "staging", "sharded_adc")


def run_paths(which, dev, peaks, rank, world):
    """Every requested path, each isolated: a failure is recorded in its entry, never in the headline."""
    import torch
    out, state = {}, {}

    def guarded(name, fn):
        t0 = time.perf_counter()
        try:
            r = fn()
        except Exception as e:                            # noqa: BLE001
            import traceback
            r = {"error": f"{type(e).__name__}: {e}", "trace": traceback.format_exc()[-600:]}
        torch.cuda.synchronize(dev)
        torch.cuda.empty_cache()
        if r is not None:
            r["bench_seconds"] = round(time.perf_counter() - t0, 1)
            out[name] = r

    single = rank == 0 and world == 1
    if single and "hnsw_small" in which:
        guarded("hnsw_small", lambda: path_hnsw_small(dev, peaks))
    if single and ("pq_encode" in which or "adc_scan" in which):
        guarded("pq_encode", lambda: path_pq_encode(dev, peaks, keep_codes=state))
    if single and "adc_scan" in which and "codes" in state:
        guarded("adc_scan", lambda: path_adc_scan(dev, peaks, state))
    state.clear()
    torch.cuda.empty_cache()
    if single and "hnsw_search" in which:
        guarded("hnsw_search", lambda: path_hnsw_search(dev, peaks))
    if single and "ivf_search" in which:
        guarded("ivf_search", lambda: path_ivf_search(dev, peaks))
    if single and "staging" in which:
        guarded("staging", lambda: path_staging(dev, peaks))
    if "sharded_adc" in which:
        guarded("sharded_adc", lambda: path_sharded_adc(dev, peaks, rank, world, True))
    return out


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the reference is pure Python and its tree does not
    exist on the GPU box) on the same metric / config / shards, rank 0 only."""
    rank, _, world = _dist_env()
    if rank != 0:
        return
    _give_blas_the_cores()
    from oracle import oracle as O
    shards = max(1, int(args.gpus))
    if args.workload != "exact":
        return run_reference_path(args)
    rows = N_ROWS * shards               # the GPU arm is weak scaling: at N GPUs every query is answered over N x 1M rows
    X = np.empty((rows, DIM), dtype=np.float32)
    for r in range(shards):              # the SAME shards as the GPU arm (numpy default_rng(SEED_DB + rank))
        X[r * N_ROWS:(r + 1) * N_ROWS] = _shard_rows(r)
    Q = _queries()
    xn = np.sqrt(np.einsum("ij,ij->i", X, X)).astype(np.float32)
    ids = np.arange(rows, dtype=np.int64)
    nq = 32                              # bounded sample per step
    for _ in range(args.warmup):
        O.exact_cosine_topk_gemm(X, ids, Q[:nq], K, xn=xn)
    t0 = time.perf_counter()
    for s in range(args.steps):
        O.exact_cosine_topk_gemm(X, ids, Q[(s * nq) % (N_QUERIES - nq):][:nq], K, xn=xn)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    real = nq / dt
    value = real * shards
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "queries_per_s_real": real,
        "config": {"workload": f"exact brute-force cosine top-{K}, {N_ROWS} x {DIM} fp32 rows per GPU x {shards} shard(s) "
                               f"= {rows} rows, {N_QUERIES}-query batch (BASELINE.json configs[1]); CPU port of the "
                               f"reference path on a {nq}-query sample per step",
                   "value_definition": "queries/s x 1M-row shards, the unit of the GPU arm: every query is answered over "
                                       f"all {shards} shard(s) ({rows} rows); queries_per_s_real is the plain rate",
                   "shards": f"numpy default_rng({SEED_DB} + shard).standard_normal -- identical to the GPU arm's"},
        "cpu_baseline": {"value": value, "unit": "queries/s", "cores": _host_threads(), "kind": "port",
                         "sample": f"{nq} queries x {rows} rows per step, numpy sgemm + argpartition "
                                   f"(oracle.exact_cosine_topk_gemm), x {shards} shard(s)"},
        "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_reference_path(args):
    """CPU arm of the non-headline workloads: the oracle restatement of that path on a bounded sample."""
    from oracle import oracle as O
    wl = args.workload
    rng = np.random.default_rng(7)
    if wl == "pq_encode":
        cents = [rng.standard_normal((256, 8)).astype(np.float32) for _ in range(48)]
        n = 20_000
        X = rng.standard_normal((n, 384), dtype=np.float32)
        t0 = time.perf_counter()
        for _ in range(max(args.steps, 1)):
            O.pq_encode(X, cents)
        dt = (time.perf_counter() - t0) / max(args.steps, 1)
        value, unit, sample, cores = n / dt, "vectors/s", f"oracle.pq_encode on {n} x 384 rows, m=48 per step", 1
    elif wl in ("adc_scan", "sharded_adc"):
        n, m = (1_000_000, 48) if wl == "adc_scan" else (1_000_000, 16)
        total = 10_000_000 if wl == "adc_scan" else 100_000_000
        codes = rng.integers(0, 256, (n, m)).astype(np.uint8)
        lut = rng.random((8, m, 256), dtype=np.float32)
        t0 = time.perf_counter()
        for _ in range(max(args.steps, 1)):
            _, _, _, cores = _cpu_adc(codes, lut, K)
        dt = (time.perf_counter() - t0) / max(args.steps, 1)
        value, unit = 8 / dt * n / total, "queries/s"
        sample = f"oracle.adc_scan_topk, 8 queries x {n}-row slice per step, EXTRAPOLATED linearly to {total} rows"
    else:
        raise SystemExit(f"--impl reference --workload {wl}: the CPU leg of this path runs inside the GPU arm "
                         f"(paths.{wl}.cpu_baseline); standalone arms exist for exact, pq_encode, adc_scan, sharded_adc")
    line = {"impl": "reference", "metric": f"{wl}: {unit}", "value": value, "unit": unit, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "strong" if wl == "sharded_adc" else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl},
            "cpu_baseline": {"value": value, "unit": unit, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
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
    which = set(PATHS) if args.paths == "all" else set(p for p in args.paths.split(",") if p and p != "none")

    if args.workload != "exact":
        paths = run_paths({args.workload} | ({"pq_encode"} if args.workload == "adc_scan" else set()), dev, peaks, rank, world)
        if rank == 0:
            p = paths.get(args.workload, {"error": "path did not run (sharded_adc is the only multi-GPU path)"})
            line = {"metric": f"{args.workload}: {p.get('unit', '?')}", "value": p.get("value"), "unit": p.get("unit"),
                    "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": p.get("ms_per_step"),
                    "higher_is_better": True, "scaling": p.get("scaling", "weak"), "vs_baseline": None,
                    "dtype": p.get("dtype", "u8"), "data": "synthetic", "config": p.get("config", {"workload": args.workload}),
                    "gpu_launches": int(_lib.launch_count()), "paths": paths}
            for key in ("roofline", "cpu_baseline", "e2e", "parity"):
                if key in p:
                    line[key] = p[key]
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- synthetic shard: numpy default_rng(SEED_DB + rank) on the host (both arms scan the same rows), resident in HBM
    X_host = _shard_rows(rank)
    Q_host = _queries()
    Qd = torch.from_numpy(Q_host).to(dev)
    index = ExactIndex(DIM, "cosine", device=dev, engine=args.engine)
    index.add_device(torch.arange(rank * N_ROWS, (rank + 1) * N_ROWS, dtype=torch.int64, device=dev),
                     torch.from_numpy(X_host).to(dev))
    Q_pin = torch.from_numpy(Q_host).pin_memory()

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

    def merge_nccl(d, i):
        gd, gi = sharded.gather_topk(d, i)                          # one packed NCCL all-gather
        return sharded.merge_shards(gd, gi, descending=True)

    def merge(d, i):
        if world == 1:
            return d, i
        if peer is not None:
            return peer.exchange_merge(d, i, descending=True)
        return merge_nccl(d, i)

    def step_device():
        return merge(*index.search_device(Qd, K))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        out = step_device()
    barrier()
    engine = "exact_tc" if ops.tc_supported(N_QUERIES, N_ROWS, DIM, K, "cosine") and args.engine != "f32" else "exact_f32"

    # once, untimed: the fused peer exchange must give what the all-gather + merge kernel gives
    exchange_equal = None
    if world > 1 and peer is not None:
        dl, il = index.search_device(Qd, K)
        a = peer.exchange_merge(dl, il, descending=True)
        b = merge_nccl(dl, il)
        torch.cuda.synchronize(dev)
        peer.check()
        exchange_equal = bool(torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]))
        barrier()

    # ---- timed region: device-resident inputs
    slots = ("exact_f32", "exact_tc", "rerank", "finalize")
    _lib.profile_enable(True)
    for slot in slots:
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
    if peer is not None:
        peer.check()                                                # a timed-out exchange would make the step fast AND wrong
    ms_total = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_step = float(ms_total.item()) / args.steps
    launches = _lib.launch_count() - launches0
    prof = {s: _lib.profile_read(s) for s in slots}
    _lib.profile_enable(False)

    # ---- the same step in a SUSTAINED loop (>= 2 s): the burst figure above is a 50 ms region
    sustained = None
    if world == 1 and not args.no_sustained:
        n_sus = max(50, int(2200.0 / max(ms_step, 0.1)))
        _lib.profile_enable(True)
        _lib.profile_read(engine)
        ev0.record()
        for _ in range(n_sus):
            step_device()
        ev1.record()
        torch.cuda.synchronize(dev)
        ns_, kms_ = _lib.profile_read(engine)
        _lib.profile_enable(False)
        for s_ in slots:
            _lib.profile_read(s_)                                    # drop the other slots' events of this loop
        sustained = {"steps": n_sus, "ms_per_step": ev0.elapsed_time(ev1) / n_sus, "kernel_ms": kms_ / max(ns_, 1)}

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
    serial = serial_pageable = None
    if world == 1:                                                   # the one-blocking-call form (ExactIndex.search)
        for src_q, name in ((Q_pin, "pinned"), (Q_host, "pageable")):
            index.search(src_q, K)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                index.search(src_q, K)
            dt = (time.perf_counter() - t0) / args.steps
            if name == "pinned":
                serial = dt
            else:
                serial_pageable = dt
    if peer is not None:
        peer.check()

    # ---- parity (untimed) at every N: each rank runs the oracle over ITS shard for a sample of the batch, rank 0
    # merges the per-shard oracle lists by (-sim, id) on the host and compares ids with the GPU's merged answer
    from oracle import oracle as O
    _give_blas_the_cores()
    ns = PARITY_QUERIES
    ids_local = np.arange(rank * N_ROWS, (rank + 1) * N_ROWS, dtype=np.int64)
    S_loc, I_loc = O.exact_cosine_topk(X_host, ids_local, Q_host[:ns], K)
    pack = torch.empty((ns, K, 2), dtype=torch.float64, device=dev)
    pack[..., 0] = torch.from_numpy(S_loc.astype(np.float64)).to(dev)
    pack[..., 1] = torch.from_numpy(I_loc.astype(np.float64)).to(dev)
    if world > 1:
        allp = [torch.empty_like(pack) for _ in range(world)]
        dist.all_gather(allp, pack)
    else:
        allp = [pack]
    if rank != 0:
        run_paths(which & {"sharded_adc"}, dev, peaks, rank, world)
        dist.destroy_process_group()
        return
    allp = torch.cat([p.cpu() for p in allp], dim=1).numpy()
    S_ref = np.empty((ns, K), np.float32)
    I_ref = np.empty((ns, K), np.int64)
    for qi in range(ns):
        ss, ii = allp[qi, :, 0].astype(np.float32), allp[qi, :, 1].astype(np.int64)
        order = np.lexsort((ii, -ss))[:K]
        S_ref[qi], I_ref[qi] = ss[order], ii[order]
    got_s, got_i = out[0][:ns].cpu().numpy(), out[1][:ns].cpu().numpy()
    parity = _parity(got_s, got_i, S_ref, I_ref, rtol=1e-4)
    parity["oracle"] = ("oracle.exact_cosine_topk over every shard (each rank its own rows), per-shard lists merged by "
                        "(-sim, id) on the host" if world > 1 else "oracle.exact_cosine_topk over the 1M rows")
    # the streamed e2e result is the same answer
    parity["e2e_result_equals_device_result"] = bool(np.array_equal(res[1], out[1].cpu().numpy()))
    if exchange_equal is not None:
        parity["peer_exchange_equals_allgather_merge"] = exchange_equal
    cpu_baseline = _cpu_exact(X_host, Q_host) if world == 1 else None
    del X_host

    n_launch, k_ms = prof[engine]
    flops = 2.0 * N_QUERIES * N_ROWS * DIM
    kernel_ms = k_ms / max(n_launch, 1)
    achieved = flops / (kernel_ms * 1e-3) / 1e12 if n_launch else None
    peak = peaks["bf16_burst"] if ms_step * args.steps < 500 else peaks["bf16_sustained"]
    real_qps = N_QUERIES / (ms_step * 1e-3)
    line = {
        "metric": METRIC, "value": real_qps * world, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if engine == "exact_tc" else "f32", "data": "synthetic",
        "queries_per_s_real": real_qps,
        "config": {
            "workload": f"exact brute-force cosine top-{K}, {N_ROWS} x {DIM} fp32 rows per GPU x {world} shard(s) = "
                        f"{N_ROWS * world} rows, {N_QUERIES}-query batch (BASELINE.json configs[1])",
            "engine": engine, "rows_total": N_ROWS * world, "recall_at_10_vs_oracle": parity["recall_at_k_vs_oracle"],
            "uncertified_queries_last_step": int(index.last_uncertified),
            "value_definition": "queries/s x 1M-row shards (weak scaling: every query is answered over all shards, each GPU "
                                "scans its own 1M rows, per-shard top-10 exchanged and merged on the GPU); the reference arm "
                                "reports the same unit over the same rows; queries_per_s_real is the plain rate",
            "shards": f"numpy default_rng({SEED_DB} + rank).standard_normal, uploaded -- identical in both arms",
            "shard_exchange": {"none": "single GPU", "peer": "one fused kernel: NVLink peer-memory stores + arrival "
                               "counters + merge (csrc/exchange.cu)",
                               "nccl": "packed NCCL all-gather + merge kernel"}[exchange],
            "l2_policy": "inputs (0.77 GB bf16 operand mirror + 1.5 GB fp32 rows per GPU) are larger than the 126 MB L2",
            "kernel_ms": {s: (round(v[1] / max(v[0], 1), 4), v[0]) for s, v in prof.items() if v[0]},
            "dtype_note": "bf16 names the tensor-core coarse pass; returned ids/distances come from the fp32 rerank "
                          "behind a rigorous certificate (uncertified queries re-scanned in fp32 inside the call)",
        },
        "clocks": clocks.summary(),
        "gpu_launches": int(launches),
        "parity": parity,
        "e2e": {"value": N_QUERIES * world / e2e_step, "unit": "queries/s",
                "h2d_bytes_per_step": int(Q_pin.numel() * 4),
                "d2h_bytes_per_step": int(N_QUERIES * K * (4 + 8)),
                "api": "ExactIndex.search_batches: host batches in, host results out, copies on their own streams "
                       "under the previous / next batch's kernels (depth 2)"},
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": _ncu_traffic(engine),
                     "kernel": engine, "kernel_ms": kernel_ms,
                     "peak_source": peaks["source"] + (": cuBLAS bf16 burst (kernel timed in a <0.5 s region)"
                                                       if peak == peaks["bf16_burst"] else ": cuBLAS bf16 sustained"),
                     "algorithmic_flops_per_launch": flops,
                     "step_frac": flops / (ms_step * 1e-3) / 1e12 / peak},
    }
    if serial is not None:
        line["e2e"]["serial_value"] = N_QUERIES / serial
        line["e2e"]["serial_pageable_value"] = N_QUERIES / serial_pageable
        line["e2e"]["serial_api"] = ("ExactIndex.search: one blocking call per batch, host in / host out; serial_value = queries in "
                                     "pinned host memory (DMA straight from the caller's buffer), serial_pageable_value = pageable "
                                     "numpy queries (staged through a pinned buffer first)")
    if sustained is not None:
        s_ach = flops / (sustained["kernel_ms"] * 1e-3) / 1e12
        line["roofline"]["sustained"] = {"kernel_ms": sustained["kernel_ms"], "achieved": s_ach, "peak": peaks["bf16_sustained"],
                                         "frac": s_ach / peaks["bf16_sustained"], "steps": sustained["steps"],
                                         "ms_per_step": sustained["ms_per_step"],
                                         "queries_per_s": N_QUERIES / (sustained["ms_per_step"] * 1e-3)}
    if cpu_baseline is not None:
        line["cpu_baseline"] = cpu_baseline
    del index
    torch.cuda.empty_cache()
    if which:
        line["paths"] = run_paths(which, dev, peaks, rank, world)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="exact", choices=["exact"] + list(PATHS),
                    help="which path is the headline of the JSON line (default: BASELINE.json configs[1])")
    ap.add_argument("--paths", default="all", help="all | none | comma list of " + ",".join(PATHS) +
                    ": the other BASELINE.json configs measured after the headline (single-GPU paths run at N = 1 only)")
    ap.add_argument("--engine", default="auto", choices=["auto", "f32", "tc"])
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N>1: fused peer-memory exchange+merge kernel, or NCCL all-gather + merge kernel")
    ap.add_argument("--no-sustained", action="store_true", help="skip the >= 2 s sustained loop of the headline step")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
