/* vdb_b200.h -- C ABI of the B200-native search core (libvdb_b200.so).
 *
 * Drop-in boundary for the hot path of GProgrammer1/Vector-DB.  The reference has no FFI of its
 * own (pure Python); each entry point below names the reference call site (file:line, relative to
 * the reference repo root) whose arithmetic it replaces.  The Python shim in vector_db_b200/
 * binds these with ctypes; INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - Plain pointers and sizes only; no torch / C++ types.  Pointers marked [dev] are CUDA device
 *     pointers (the shim passes tensor.data_ptr()), [host] are host pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All work is
 *     enqueued on it; nothing synchronises unless stated.
 *   - The library keeps no persistent allocations: scratch memory is passed in as `ws`/`ws_bytes`
 *     and sized by the matching *_workspace() query.
 *   - Return 0 on success, a VDB_ERR_* code otherwise; vdb_last_error() returns a thread-local
 *     human-readable message.  There is NO CPU fallback: without a CUDA device every compute entry
 *     point fails with VDB_ERR_CUDA.
 *   - Top-k outputs are ordered by (distance, index) ascending -- the tuple order of
 *     tests/indexing/test_hnsw.py:128-134 and src/vector_db/indexing/hnsw.py:366 -- and padded
 *     with (+inf, -1) when fewer than k results exist.  L2 distances are NON-squared
 *     (src/vector_db/util/distance.py:3-4).  Row indices are positions in the device arrays; when
 *     an `ids` table is given, outputs are translated to external ids.
 */
#ifndef VDB_B200_H
#define VDB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VDB_OK 0
#define VDB_ERR_INVALID 1    /* bad argument (shim raises ValueError) */
#define VDB_ERR_CUDA 2       /* CUDA runtime failure / no device (RuntimeError) */
#define VDB_ERR_WORKSPACE 3  /* workspace too small (RuntimeError) */
#define VDB_ERR_UNSUPPORTED 4/* shape outside the kernels' envelope (ValueError) */
#define VDB_ERR_OVERFLOW 5   /* a bounded on-chip structure overflowed; retry with the bigger variant */

#define VDB_METRIC_L2 0
#define VDB_METRIC_COSINE 1

const char* vdb_last_error(void);
int vdb_abi_version(void);
/* number of kernels this library launched from the calling process so far (bench.py gpu_launches) */
uint64_t vdb_launch_count(void);

/* Per-kernel timing for bench.py: while enabled, every launch of a dominant kernel is bracketed by
 * CUDA events on its own stream.  vdb_profile_read() waits for the slot's events, returns the
 * number of launches and their summed duration, and clears the slot.
 * Slots: 0 exact fp32 scan, 1 exact tensor-core scan, 2 ADC scan, 3 PQ encode, 4 HNSW search,
 * 5 rerank, 6 top-k finalise/merge, 7 HNSW insert, 8 IVF list scan. */
void vdb_profile_enable(int on);
int vdb_profile_read(int slot, int* launches, float* total_ms);

/* ---- a9/a10 storage staging ---------------------------------------------------------------
 * De-interleave `n` records of the reference's layer-1 memmap layout
 * {int64 id; float32 embedding[D]} (stride 8+4D, src/vector_db/storage.py:135-144) from PINNED
 * host memory into SoA device arrays with two strided DMA copies.  */
int vdb_stage_aos_h2d(const void* host_records /*[host,pinned]*/, int64_t n, int D,
                      int64_t* ids /*[dev] n*/, float* vecs /*[dev] n*D*/, void* stream);

/* The same upload from PAGEABLE memory -- the memmap of the reference's `.embeddings.npy` file itself
 * (storage.py:135-144).  The records are cut into chunks of buffer_bytes; host worker t (of `threads`, each with its own
 * CUDA stream) takes chunks t, t + threads, ...: memcpy out of the page cache into one of its pinned buffers, one
 * contiguous DMA into the matching device buffer, a kernel that splits the id / embedding columns (a strided DMA of
 * 8-byte rows is 5x slower than the contiguous copy).  `depth` buffers in each ring, at least two per worker (threads is
 * cut to depth / 2).  Returns after every copy and kernel has finished; `stream` is made to wait for them. */
int vdb_stage_aos_pageable_h2d(const void* records /*[host,pageable] n records*/, int64_t n, int D,
                               int64_t* ids /*[dev] n or NULL*/, float* vecs /*[dev] n*D*/,
                               void* pinned_ring /*[host,pinned] depth*buffer_bytes*/,
                               void* device_ring /*[dev] depth*buffer_bytes*/,
                               size_t buffer_bytes, int depth, int threads, void* stream);

/* Per-row squared norms and the bf16 mirror used by the tensor-core scan.
 * Either output may be NULL. */
int vdb_prepare_rows(const float* vecs /*[dev] N*D*/, int64_t N, int D,
                     float* sqnorm /*[dev] N*/, void* bf16_mirror /*[dev] N*D bf16*/, void* stream);

/* ---- a1/a3 exact scan, fp32 direct-difference form ------------------------------------------
 * Replaces the per-pair loop euclidean_vector_distance + sort of
 * tests/indexing/test_hnsw.py:128-134 / src/vector_db/indexing/ivf.py:176-199.
 * metric L2: out_d = sqrt(sum((q-x)^2)), ascending.  metric COSINE (no reference implementation;
 * largest-first order of src/vector_db/util/math.py:3-5): out_d = q.x/(|q||x|), descending.
 * `ids` may be NULL (outputs are row indices).  `row_mask` (optional bitmap, bit r of word r/32)
 * restricts results to set rows (the filter_ids of hnsw.py:96,127).  N < 2^32. */
size_t vdb_exact_topk_f32_workspace(int Q, int64_t N, int k);
int vdb_exact_topk_f32(const float* q /*[dev] Q*D*/, int Q, const float* db /*[dev] N*D*/,
                       const int64_t* ids /*[dev] N or NULL*/, const uint32_t* row_mask /*[dev] or NULL*/,
                       int64_t N, int D, int k, int metric,
                       float* out_d /*[dev] Q*k*/, int64_t* out_i /*[dev] Q*k*/,
                       void* ws, size_t ws_bytes, void* stream);

/* Allocation-free variant of vdb_exact_topk_f32.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same inputs, outputs and results as vdb_exact_topk_f32, but the scan's scratch (Q x slices x k
 * partial keys of 8 bytes) lives in a file-scope `static __device__` array of vdb_exact_topk_f32_static_max_bytes()
 * = 8 MiB inside the library (held by every GPU that loads it, until exit) instead of a caller workspace; the call
 * makes no cudaMalloc / cudaFree / allocating library call.  Shapes whose scratch exceeds the limit are rejected
 * with VDB_ERR_UNSUPPORTED and a message naming the need (Q = 4096, k = 10 needs 1 MB; Q x k up to ~100k fits in the
 * worst case of 4 x SMs slices per query group).  All calls on a device SHARE the array, which keeps the previous
 * call's contents (the entry point re-initialises what it reads): calls that overlap on the GPU -- different streams,
 * different threads -- race; serialise them on one stream. */
size_t vdb_exact_topk_f32_static_max_bytes(void);
int vdb_exact_topk_f32_static(const float* q /*[dev] Q*D*/, int Q, const float* db /*[dev] N*D*/,
                              const int64_t* ids /*[dev] N or NULL*/, const uint32_t* row_mask /*[dev] or NULL*/,
                              int64_t N, int D, int k, int metric,
                              float* out_d /*[dev] Q*k*/, int64_t* out_i /*[dev] Q*k*/, void* stream);

/* Exact fp32 rerank of per-query candidate rows (cand[q][r], -1 = empty; duplicates allowed and
 * collapsed): the direct-difference distance of util/distance.py:3-4 and the (dist,id) order.
 * out_kth (optional) receives the k-th returned distance per query (+inf if fewer). */
int vdb_rerank_f32(const float* q, int Q, const float* db, const int64_t* ids, int64_t N, int D,
                   const int32_t* cand /*[dev] Q*R*/, int R, int k, int metric,
                   float* out_d, int64_t* out_i, void* stream);

/* ---- a3 exact scan on tcgen05 tensor cores ------------------------------------------------
 * Coarse pass: bf16 x bf16 -> fp32 GEMM tiles (TMA-fed tcgen05.mma, accumulators in TMEM) over a
 * per-metric OPERAND MIRROR of the rows, built so that the accumulator is the ranking score
 * itself: L2 rows hold bf16(-2x) plus a three-way bf16 split of |x|^2 in three extra K columns
 * (score = |x|^2 - 2 q.x), cosine rows hold bf16(-x/|x|) (score = -q.x/|x|).  The fused epilogue
 * reduces every accumulator row to per-bin minima (bin = VDB_TC_BIN consecutive rows) and keeps
 * the best bins per (query, db slice) in registers.  vdb_exact_topk_tc merges the slices, reranks
 * every row of the selected bins in fp32 (the reference's direct-difference arithmetic) and
 * certifies the result with a rigorous bf16 error bound, eps = s (|q~-q| A + |q| B + 2^-13 |q| A), s = 2 for L2 and
 * 1 for cosine (two Cauchy-Schwarz terms, one per rounded operand, on MEASURED residual norms); the queries it cannot
 * certify are re-scanned by the fp32 kernel inside the same call.
 * Requirements: D % 8 == 0, one query tile's operand fits shared memory (D <= 512), k <= 16;
 * vdb_exact_topk_tc_supported() answers for a shape.
 * Mirror: one bf16 buffer of vdb_tc_mirror_elems(N, D, metric) elements, 128-byte aligned: the N x D operand
 * rows, then (L2) a packed [ceil4(N)][16] block holding the bf16 split of |x|^2 of every row. */
#define VDB_TC_BIN 8
int vdb_exact_topk_tc_supported(int Q, int64_t N, int D, int k, int metric);
size_t vdb_tc_mirror_elems(int64_t N, int D, int metric);
int vdb_tc_prepare_rows(const float* vecs /*[dev] N*D*/, int64_t N, int D, int metric,
                        float* sqnorm /*[dev] N or NULL*/, void* mirror /*[dev] vdb_tc_mirror_elems bf16*/,
                        float* stats /*[dev] 2 floats out: A = max |x~| and B = max |x~ - x| over the rows, x~ the bf16
                           operand row (x for L2, x/|x| for cosine) -- the measured rounding residual the certificate
                           of vdb_exact_topk_tc is built on*/,
                        void* stream);
size_t vdb_exact_topk_tc_workspace(int Q, int64_t N, int D, int k, int metric);
int vdb_exact_topk_tc(const float* q /*[dev] Q*D fp32*/, int Q,
                      const float* db /*[dev] N*D fp32*/, const void* mirror /*[dev] see above*/,
                      const float* db_stats /*[dev] the 2 floats vdb_tc_prepare_rows wrote for this mirror */,
                      const int64_t* ids /*[dev] or NULL*/, int64_t N, int D, int k, int metric,
                      float* out_d, int64_t* out_i,
                      int32_t* out_uncertified /*[dev] Q+1: count, then the indices of the queries the certificate rejected;
                         those queries were re-scanned by the exact fp32 kernel inside this call (device-side list, no host
                         round trip), so out_d / out_i are exact for every query on return*/,
                      void* ws, size_t ws_bytes, void* stream);
/* ---- f3: list-major IVF on the same tensor-core pipeline (src/vector_db/indexing/ivf.py:170-199 for a BATCH of queries) --------
 * The query-major list scan (vdb_ivf_scan) reads every probed list once per query; at batch sizes of thousands the lists are
 * shared, and the work is a set of small dense contractions: (queries probing list l) x (rows of list l).
 *   lm_vecs / lm_ids: the rows in LIST-MAJOR order, every list padded to whole 256-row tiles (padding: id -1);
 *   lm_mirror / lm_stats: vdb_tc_prepare_rows(L2) of the RESIDUALS row - centroid[list] in the same order (padding rows:
 *     zero residual and a huge |x|^2 entry in the norm block) -- the query operand holds q - centroid[list] per pair, so the
 *     accumulator is the squared distance and the bf16 error scales with the residual norms, not with |q||x|;
 *   row_query[Qa]: the query of every row of the query operand -- one row per (query, probed list) PAIR, the pairs of a list
 *     contiguous and padded to whole 128-row tiles (-1 = padding); pair_rows[Q][n_probe]: the operand row of a query's p-th
 *     probe (-1 = that list is empty); unit_table[n_units][4] = (pair tile, first row tile, end row tile, 0): one unit per
 *     pair tile of a non-empty list.
 * The coarse pass keeps bin lists per pair; selection, trim, exact fp32 rerank and certificate then run per QUERY over the
 * lists of all its pairs, exactly as in vdb_exact_topk_tc.  out_i = lm_ids[position] (or the list-major position when
 * lm_ids is NULL), ordered by (distance, list-major position).  Queries the certificate rejects are LISTED in
 * out_uncertified (count, then indices) and must be re-run by the caller through vdb_ivf_scan (a brute-force rescan would
 * look at lists the query did not probe).  n_probe <= 32, k <= 16, D % 8 == 0, D <= 512. */
size_t vdb_ivf_topk_tc_workspace(int Q, int64_t Qa, int64_t N, int D, int k, int n_units);
int vdb_ivf_topk_tc(const float* q /*[dev] Q*D*/, int Q, const float* lm_vecs /*[dev] N*D*/, const void* lm_mirror,
                    const float* lm_stats /*[dev] 2*/, const int64_t* lm_ids /*[dev] N or NULL*/, int64_t N /* padded rows */,
                    int D, int k, int n_probe, const int32_t* pair_rows /*[dev] Q*n_probe*/,
                    const int32_t* row_query /*[dev] Qa*/, const int32_t* row_list /*[dev] Qa: the pair's list*/,
                    const float* centroids /*[dev] K*D*/, int64_t Qa, const int32_t* unit_table /*[dev] n_units*4*/,
                    int n_units, const int32_t* meta /*[dev] vdb_ivf_group_pairs' totals (then Qa / n_units are upper bounds) or NULL*/,
                    const float* list_stats /*[dev] K*2 per-list (max |bf16 r|, max |bf16 r - r|) or NULL*/,
                    float* out_d /*[dev] Q*k*/, int64_t* out_i /*[dev] Q*k*/,
                    int32_t* out_uncertified /*[dev] Q+1*/, void* ws, size_t ws_bytes, void* stream);

/* Device-side preparation / completion of vdb_ivf_topk_tc (no host round trip):
 *   vdb_ivf_group_pairs: (query, probe) pairs grouped by list -> operand-row maps, unit table and meta = {operand rows, units};
 *     list_tile_ptr [K+1] = first 256-row tile of every list in the list-major copy (empty lists have no tiles); Qa_max >=
 *     (ceil(Q n_probe / 128) + min(K, Q n_probe)) * 128 rows, unit_table has Qa_max / 128 rows of 4 ints.
 *   vdb_ivf_tc_finish: the k2 = top_k + 1 results of vdb_ivf_topk_tc (list-major positions) -> the reference's stable order
 *     (distance, enumeration sequence; ivf.py:177-193), external ids, tie[q] = 1 when equal distances straddle the cut. */
/* Probe selection of a batch from the kc = n_probe + 4 (<= 16) nearest centroids per query returned by vdb_exact_topk_tc
 * (L2, ids = NULL): the candidates are re-scored in numpy's summation order (ivf.py:170), queries whose n_probe-th candidate
 * is not provably inside the candidate set are flagged and go through the full numpy-order kernels inside the call. */
int vdb_ivf_probe_refine(const float* q /*[dev] Q*D*/, int Q, int D, const float* centroids /*[dev] K*D*/, int K,
                         const float* cand_d /*[dev] Q*kc*/, const int64_t* cand_i /*[dev] Q*kc*/, int kc, int n_probe,
                         const int32_t* list_ptr /*[dev] K+1 or NULL*/, int32_t* probes /*[dev] Q*n_probe*/,
                         int32_t* prefix /*[dev] Q*(n_probe+1) or NULL*/, float* cdist_scratch /*[dev] Q*K*/,
                         int32_t* flags /*[dev] Q*/, void* stream);
size_t vdb_ivf_group_workspace(int K);
int vdb_ivf_group_pairs(const int32_t* probes /*[dev] Q*n_probe*/, int Q, int n_probe, const int32_t* list_tile_ptr /*[dev] K+1*/,
                        int K, int64_t Qa_max, int32_t* row_query /*[dev] Qa_max*/, int32_t* row_list /*[dev] Qa_max*/,
                        int32_t* pair_rows /*[dev] Q*n_probe*/, int32_t* unit_table /*[dev] Qa_max/128*4*/, int32_t* meta /*[dev] 2*/,
                        void* ws, size_t ws_bytes, void* stream);
int vdb_ivf_tc_finish(const float* d /*[dev] Q*k2*/, const int64_t* pos /*[dev] Q*k2*/, int Q, int k2, int top_k,
                      const int32_t* probes /*[dev] Q*n_probe*/, const int32_t* prefix /*[dev] Q*(n_probe+1)*/, int n_probe,
                      const int32_t* lm_list /*[dev] N: list of every list-major row*/, const int64_t* lm_start /*[dev] K*/,
                      const int64_t* lm_ids /*[dev] N*/, float* out_d /*[dev] Q*top_k*/, int64_t* out_ids /*[dev] Q*top_k*/,
                      int32_t* tie /*[dev] Q*/, void* stream);

/* bench/debug only: bit 0 = skip the whole epilogue of the coarse pass, bit 1 = skip only its threshold
 * test and bin-list inserts (results are then invalid) */
void vdb_debug_set_tc_flags(int flags);

/* ---- a11 PQ encode -------------------------------------------------------------------------
 * codes[i][j] = argmin_c sqrt(sum((C_j[c]-x_ij)^2)), first minimum wins
 * (src/vector_db/inference/pq.py:77-89).  codebooks laid out [m][K][dsub] fp32, K <= 256,
 * codes uint8 [N][m]. */
int vdb_pq_encode(const float* x /*[dev] N*D*/, int64_t N, int D, int m, int K,
                  const float* codebooks /*[dev] m*K*dsub*/, uint8_t* codes /*[dev] N*m*/, void* stream);

/* Same result (bit-exact codes) through the tensor cores, for dsub == 8 and K <= 256: two
 * tcgen05.mma.kind::tf32 per (128-row tile, sub-space) give every |c|^2 - 2 x.c score, the epilogue
 * keeps the centroids within a rigorous rounding margin of the minimum and resolves them with the
 * reference's exact arithmetic.  ws: vdb_pq_encode_tc_workspace(m) bytes, 256-byte aligned. */
int vdb_pq_encode_tc_supported(int64_t N, int D, int m, int K);
size_t vdb_pq_encode_tc_workspace(int m);
int vdb_pq_encode_tc(const float* x, int64_t N, int D, int m, int K, const float* codebooks, uint8_t* codes,
                     void* ws, size_t ws_bytes, void* stream);
/* bench/debug only: 1 (default) = the f16-accumulator kernel, 0 = the fp32-accumulator kernel; counters of the last
 * f16-accumulator call on workspace `ws`: out[0] = (row, sub-space) pairs the resolver warp decided, out[1] = of which
 * with all-ones candidate masks, out[2] = pairs an epilogue lane resolved itself because the queue was full */
void vdb_debug_set_pq_tc_mode(int mode);
int vdb_debug_pq_tc_stats(const void* ws, int m, unsigned long long* out /*[host] 3*/, void* stream);

/* ---- a12 PQ training (Lloyd iterations of scipy kmeans2 as called at pq.py:11) ------------------
 * One assignment + mean-update round for all m sub-spaces at once; an empty cluster keeps its
 * centroid.  `labels` int32 [m][N] receives the assignment made BEFORE the update (kmeans2's
 * return convention).  ws: m*K*(dsub+1) floats. */
size_t vdb_pq_kmeans_workspace(int m, int K, int dsub);
int vdb_pq_kmeans_step(const float* x, int64_t N, int D, int m, int K, float* codebooks /*in/out*/,
                       int32_t* labels /*[dev] m*N or NULL*/, void* ws, size_t ws_bytes, void* stream);
/* The same round with the assignment supplied as uint8 codes [N][m] (vdb_pq_encode / vdb_pq_encode_tc of the
 * CURRENT codebooks: the nearest-centroid rule of pq.py:77-80) -- mean update only.  Same workspace. */
int vdb_pq_kmeans_step_coded(const float* x, int64_t N, int D, int m, int K, const uint8_t* codes /*[dev] N*m*/,
                             float* codebooks /*in/out*/, void* ws, size_t ws_bytes, void* stream);

/* ---- a13 ADC scan (no reference symbol; sub-space layout pq.py:48-62, distance pq.py:77-80) ----
 * lut[q][j][c] = sum_t (q_jt - C_j[c][t])^2, fp32 [Q][m][K]. */
int vdb_pq_lut(const float* q /*[dev] Q*D*/, int Q, int D, int m, int K, const float* codebooks,
               float* lut /*[dev] Q*m*K*/, void* stream);
/* d(q,i) = sum_j lut[q][j][codes[i][j]] on a per-query fixed-point table (lutq = rint(lut * 2^e),
 * e = 23 - exponent(max entry of the query's table); exact integer sums, so the result does not depend on
 * the summation order and is bit-reproducible); the R smallest by (d_int, row), reported as
 * float(d_int) * 2^-e.  K must be 256 (nbits=8), m <= 255.  out_i are row indices (int64), or ids[row]
 * when ids != NULL.  row_mask (nullable): bitmap over rows, a cleared bit skips the row -- tombstones and
 * the metadata pre-filter (filter_ids of src/vector_db/indexing/hnsw.py:96,127) -- the result is exact over
 * the remaining rows. */
size_t vdb_pq_adc_topk_workspace(int Q, int64_t N, int m, int R);
int vdb_pq_adc_topk(const uint8_t* codes /*[dev] N*m*/, int64_t N, int m, const float* lut, int Q, int R,
                    const int64_t* ids, const uint32_t* row_mask /*[dev] nullable*/, float* out_d /*[dev] Q*R*/,
                    int64_t* out_i /*[dev] Q*R*/, void* ws, size_t ws_bytes, void* stream);
/* Debug/A-B switch: 1 (default) = the threshold-collect pass runs on 16-bit lower-bound tables, eight queries per
 * shared-memory lookup, and the collected rows are re-evaluated on the 32-bit tables (same result, bit for bit);
 * 0 = the collect pass itself uses the 32-bit tables, four queries per lookup. */
void vdb_debug_set_adc_coarse(int on);
/* Test hook: the 15-bit tables of that collect pass for given fixed-point tables lutq [Q][m][256] and thresholds tau [Q]
 * (all device pointers).  lut16: [ceil(Q/G)][ceil(m/8) * 2048 slots][G] u16 with slot = ((j/8)*256 + c)*8 + j%8;
 * thr16[q] = 0x8000 + threshold.  G = queries per lookup group (2, 4 or 8). */
int vdb_debug_adc_coarse_tables(const uint32_t* lutq, int m, int Q, const uint32_t* tau, uint16_t* lut16, uint32_t* thr16,
                                int G, void* stream);

/* ---- a4/a5/a6 batched HNSW search -------------------------------------------------------------
 * One CTA per query: greedy descent on levels max_level..1 (hnsw.py:289-328, strict '<'), then the
 * best-first beam of hnsw.py:79-134 on level 0 with the same termination rule
 * (len(found) >= ef and d > worst), result order (dist,row) and result-only filtering.
 * Graph layout: adj0 int32 [N][deg0] (-1 padded, deg0 = 2M); upper levels: node r with top level
 * L_r owns rows up_ptr[r] .. up_ptr[r]+L_r-1 of adj_up int32 [*][deg_up] (level l at up_ptr[r]+l-1).
 * `alive` (optional bitmap): cleared rows are tombstones -- marked visited, never evaluated
 * (hnsw.py:116-121).  `filter` (optional bitmap): only set rows may enter the result list.
 * out_stats (optional) uint32 [Q][2] = {distance evaluations, expansions}.  ef <= 1024.
 * The visited set is an open-addressing hash: hash_slots = 0 sizes it automatically (shared memory
 * when it fits, else global memory inside ws); a power of two forces that many slots in global
 * memory.  After the stream has drained, the first int32 of ws is non-zero if any query filled its
 * visited table (its result may be truncated): retry with a larger hash_slots. */
size_t vdb_hnsw_search_workspace(int Q, int ef, int deg0);
int vdb_hnsw_search(const float* q, int Q, const float* db, int64_t N, int D,
                    const int32_t* adj0, int deg0, const int32_t* adj_up, int deg_up,
                    const int32_t* up_ptr, int32_t entry_row, int max_level, int ef, int k,
                    const uint32_t* alive, const uint32_t* filter,
                    const int32_t* levels /*[dev] N or NULL: top level of every row; a list above it reads as empty
                       (the reference's entry point can lack its top lists, hnsw.py:179-191)*/,
                    int hash_slots,
                    float* out_d /*[dev] Q*k*/, int32_t* out_rows /*[dev] Q*k*/, uint32_t* out_stats,
                    int32_t* out_status /*[dev] Q or NULL: per query, bit 0 = visited table filled up, bit 1 = the
                       bounded frontier dropped candidates (filtered search only: the reference's heap is unbounded,
                       hnsw.py:102-131) -- the caller finishes such queries with the exact masked scan*/,
                    void* ws, size_t ws_bytes, void* stream);

/* bench/debug only: CTA shared-memory footprint above which the visited hash moves to global memory */
void vdb_debug_set_hnsw_smem_hash_limit(int bytes);
/* bench/debug only: resident CTAs per SM the search grid is sized for (default 12, at most 16) */
void vdb_debug_set_hnsw_ctas_per_sm(int n);
void vdb_debug_set_hnsw_insert_threads(int n);   /* 128 | 256 (default) | 512 threads of the sequential insert CTA */

/* ---- a7 HNSW construction ----------------------------------------------------------------------
 * Sequential-semantics insert of rows first..first+count-1 (their vectors already in `db`, their
 * levels drawn by the host from random.Random, hnsw.py:66-68,160) by ONE CTA, reproducing
 * hnsw.py:143-225: greedy descent, ef_construction beam per level, M closest as links, reverse
 * links with closest-cap truncation.  state int32[4] = {entry_row, max_level, evals_lo, evals_hi} is
 * read and updated (evals = running count of distance evaluations, for parity checks against the
 * reference's call count).  First int32 of ws = visited-table overflow flag, STICKY over the calls of a build: the
 * caller zeroes it before the first chunk and reads it after the last. */
size_t vdb_hnsw_insert_workspace(int efc, int deg0);
int vdb_hnsw_insert(const float* db, int64_t N, int D, int32_t* adj0, int deg0, int32_t* adj_up,
                    int deg_up, const int32_t* up_ptr, const int32_t* levels /*[dev] N*/,
                    int32_t first, int32_t count, int M, int efc, int32_t* state /*[dev] 4*/,
                    const uint32_t* alive, void* ws, size_t ws_bytes, void* stream);

/* ---- a7, wave-parallel construction -----------------------------------------
 * The reference builds node by node (hnsw.py:143-225, ~50 inserts/s); these two entry points build
 * level by level, top-down, in waves: vdb_hnsw_build_search runs the reference's insert-time search
 * (greedy descent hnsw.py:174-176, ef_construction beam hnsw.py:182) for every node of a wave against
 * the graph built so far and returns its M closest (hnsw.py:188); vdb_hnsw_reverse_merge links the
 * wave into its neighbours' lists and cuts them to the closest `deg` (hnsw.py:199-213).  Lists are
 * kept sorted by distance, with the distances in a parallel float array.  The result is an
 * approximation of the sequential build (wave members do not see each other), not the reference's
 * graph; vdb_hnsw_insert remains the parity builder.  First int32 of ws = visited-table overflow flag,
 * STICKY over the calls of a build: the caller zeroes it before the first wave and reads it after the last. */
size_t vdb_hnsw_build_search_workspace(int count, int efc, int deg0);
int vdb_hnsw_build_search(const float* db, int64_t N, int D, const int32_t* adj0, int deg0,
                          const int32_t* adj_up, int deg_up, const int32_t* up_ptr,
                          const int32_t* nodes /*[dev] count rows of this wave*/, int count, int level,
                          int32_t entry_row, int max_level, int32_t seed_entry, int efc, int M,
                          int32_t* out_rows /*[dev] count*M*/, float* out_d /*[dev] count*M*/,
                          void* ws, size_t ws_bytes, void* stream);
int vdb_hnsw_reverse_merge(int32_t* adj /*[dev] lists*deg*/, float* adjd /*[dev] lists*deg*/, int deg,
                           const int32_t* list_idx /*[dev] segments*/, const int32_t* seg_start /*[dev] segments+1*/,
                           int segments, const int32_t* in_rows /*[dev]*/, const float* in_d /*[dev]*/, void* stream);

/* ---- f3: IVF (src/vector_db/indexing/ivf.py) -------------------------------------------------
 * The index is K centroids [K,D] and K inverted lists in CSR form: list_ptr[K+1], list_rows[] = row
 * indices into the fp32 matrix `db`, each list in the reference's append order (ivf.py:78-81,113).
 *
 * vdb_ivf_centroid_dist: out[q][c] = np.linalg.norm(centroids[c] - query[q]) (ivf.py:109,170) in numpy's
 *   operation order (fp32 subtract and square, pairwise_sum along the row, sqrt) -- bit-identical values,
 *   so the probe set and the nearest-centroid assignment of add() are the reference's.
 * vdb_ivf_probe: probes[q][:] = np.argsort(out[q])[:n_probe] (ivf.py:171; ties by centroid index) and,
 *   when `prefix` is given, prefix[q][p] = number of candidates enumerated before probe p
 *   (prefix[q][n_probe] = total).  n_probe = 1 is the argmin of ivf.py:110.
 * vdb_ivf_scan: euclidean_vector_distance (util/distance.py:3-4) to every member of the probed lists,
 *   candidates ordered by distance with enumeration order (probe rank, list position) as tie-break --
 *   the stable sort of ivf.py:177-193 -- first k returned.  Rows whose bit in `alive` is cleared are
 *   skipped (ivf.py:188-190).  out_scanned[q] (nullable) = candidates enumerated.  The distance
 *   arithmetic is vdb_exact_topk_f32's, so n_probe == K returns that scan's values.
 * vdb_ivf_centroid_mean: Lloyd mean update over the lists (scipy kmeans2 as called at ivf.py:75: an
 *   empty cluster keeps its centroid). */
int vdb_ivf_centroid_dist(const float* q /*[dev] Q*D*/, int Q, int D, const float* centroids /*[dev] K*D*/, int K,
                          float* out /*[dev] Q*K*/, void* stream);
int vdb_ivf_probe(const float* cdist /*[dev] Q*K*/, int Q, int K, int n_probe, const int32_t* list_ptr /*[dev] K+1, nullable*/,
                  int32_t* probes /*[dev] Q*n_probe*/, int32_t* prefix /*[dev] Q*(n_probe+1), nullable*/, void* stream);
size_t vdb_ivf_scan_workspace(int Q, int k);
int vdb_ivf_scan(const float* q /*[dev]*/, int Q, const float* db /*[dev] rows*D*/, int D,
                 const int32_t* list_ptr /*[dev]*/, const int32_t* list_rows /*[dev]*/, const int32_t* probes /*[dev]*/,
                 const int32_t* prefix /*[dev]*/, int n_probe, const uint32_t* alive /*[dev] nullable*/,
                 const int64_t* ids /*[dev] nullable*/, int k, float* out_d /*[dev] Q*k*/, int64_t* out_i /*[dev] Q*k*/,
                 int64_t* out_scanned /*[dev] Q, nullable*/, void* ws, size_t ws_bytes, void* stream);
int vdb_ivf_centroid_mean(const float* x /*[dev] rows*D*/, int D, const int32_t* list_ptr /*[dev] K+1*/,
                          const int32_t* list_rows /*[dev]*/, int K, float* centroids /*[dev] K*D, in place*/, void* stream);

/* IVF-PQ: ADC (vdb_pq_adc_topk's fixed-point arithmetic) over the members of the probed lists only.
 * codes are LIST-MAJOR: the code of CSR position p (row list_rows[p]) is codes[p*m .. p*m+m).  Returns the
 * R best by (d_int, candidate enumeration order) as squared-distance estimates and ROW indices (-1 padded),
 * ready for vdb_rerank_f32.  No reference symbol (the reference has IVF, ivf.py, and PQ, pq.py, separately). */
size_t vdb_ivf_adc_workspace(int Q, int m, int R);
int vdb_ivf_adc_scan(const uint8_t* codes /*[dev] total*m*/, int m, const float* lut /*[dev] Q*m*256*/, int Q,
                     const int32_t* list_ptr /*[dev]*/, const int32_t* list_rows /*[dev]*/, const int32_t* probes /*[dev]*/,
                     const int32_t* prefix /*[dev]*/, int n_probe, int R, float* out_d /*[dev] Q*R*/,
                     int32_t* out_rows /*[dev] Q*R*/, void* ws, size_t ws_bytes, void* stream);

/* ---- f4: metadata pre-filter --------------------------------------------------------------------
 * Replaces StorageService.filter_by_metadata (src/vector_db/services/storage_service.py:106-128: a Python
 * scan that JSON-decodes every node) for the search path.  column = dictionary codes of ONE metadata key,
 * int32 [n] in row order (0 = key missing / value None); ANDs `column[r] == want` (and the alive bitmap,
 * if given) into bitmap [ceil(n/32)] -- first != 0 overwrites instead.  The bitmap is the row_mask /
 * filter_bitmap argument of the search entry points (what hnsw.py:96,127 does with filter_ids). */
int vdb_filter_eq(const int32_t* column /*[dev] n*/, int64_t n, int32_t want, const uint32_t* alive /*[dev] nullable*/,
                  uint32_t* bitmap /*[dev] ceil(n/32)*/, int first, void* stream);

/* ---- e: multi-GPU merge -----------------------------------------------------------------------
 * lists: L sorted lists per query, d [L][Q][k] / i [L][Q][k] (the all-gathered per-shard results);
 * selects the k best by (d, list, position) -- equal to (d,id) order when shard l owns smaller ids
 * than shard l+1.  descending != 0 selects the largest (cosine similarities). */
int vdb_merge_topk(const float* d, const int64_t* i, int L, int Q, int k, int descending,
                   float* out_d, int64_t* out_i, void* stream);

/* Exchange + merge in one kernel over NVLink peer memory (csrc/exchange.cu).  Every rank owns an exchange
 * buffer of vdb_exchange_bytes(world, cap) bytes, zero-filled once, mapped into every peer (torch symmetric
 * memory; cap >= Q*k elements per list plane); peer_bufs is the device array of the `world` base addresses in
 * rank order.  Call number c (0, 1, ... in step on all ranks, one stream per rank) passes parity = c & 1 and
 * target = the running sum of vdb_exchange_grid(Q) over the calls made so far with that parity, this one
 * included.  Each rank stores its lists d/i [Q][k] into every rank's buffer, waits for all ranks' stores and
 * merges: out = k best by (value, shard, position), exactly vdb_merge_topk of the all-gathered lists.
 * status (device int32) becomes 1 if a wait ran into its clock limit (a peer never arrived).
 * Replaces, for this path, torch.distributed.all_gather + vdb_merge_topk (no reference symbol: the reference
 * scales by replicas only, README.md:116-126). */
size_t vdb_exchange_bytes(int world, int64_t cap);
int vdb_exchange_grid(int Q);
int vdb_exchange_merge_topk(const float* d /*[dev] Q*k*/, const int64_t* i /*[dev] Q*k*/, int Q, int k, int descending,
                            const uint64_t* peer_bufs /*[dev] world*/, int rank, int world, int64_t cap, int parity,
                            uint32_t target, float* out_d /*[dev] Q*k*/, int64_t* out_i /*[dev] Q*k*/,
                            int32_t* status /*[dev]*/, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VDB_B200_H */
