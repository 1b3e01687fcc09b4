/*
 * muninn_cuda.h — C ABI of libmuninn_cuda: the B200 (sm_100a) implementation of
 * sqlite-muninn's vector-search hot path.
 *
 * Every entry point below replaces, or is the batched form of, one reference interface
 * (paths relative to /root/reference/):
 *
 *   muninn_cuda_index_create / _destroy   hnsw_create / hnsw_destroy         src/hnsw_algo.h:56-59
 *   muninn_cuda_index_upload              load_index_from_shadow             src/hnsw_vtab.c:286-341
 *   muninn_cuda_index_patch               the effect of hnsw_insert / hnsw_delete on the stored graph
 *                                                                            src/hnsw_algo.c:520-666, :717-805
 *   muninn_cuda_search_batch              hnsw_search, called at             src/hnsw_vtab.c:604
 *                                         (signature src/hnsw_algo.h:73), for nq queries at once
 *   muninn_cuda_search_batch_device       same, device-resident queries/results (benchmark `value`)
 *   muninn_cuda_exact_knn                 brute_force_knn test helper        test/test_hnsw_algo.c:10-34
 *   muninn_cuda_set_exact_mode / _exact_stats / _exact_filter_probe   evaluation strategy of the same
 *   muninn_cuda_distance_batch            vec_get_distance_func(metric)(a,b) src/vec_math.c:180-190
 *   muninn_cuda_parse_metric              vec_parse_metric                   src/vec_math.c:192-204
 *   muninn_cuda_build_device              the searches inside a loop of hnsw_insert calls (bulk build)
 *                                                                            src/hnsw_algo.c:553-579
 *
 * Plain C types only: pointers, sizes and ints.  No CPU fallback exists anywhere in the
 * library: without a CUDA device every compute entry point returns MUNINN_CUDA_E_NODEVICE.
 *
 * Threading: entry points are not re-entrant on one handle (the reference has no locks
 * either; SQLite serialises calls on a connection).  Separate handles are
 * independent.  Multi-GPU = one process (or at least one handle) per device.
 */
#ifndef MUNINN_CUDA_H
#define MUNINN_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* VecMetric, src/vec_math.h:13-16 — the integer stored in the _config shadow table */
enum { MUNINN_METRIC_L2 = 0, MUNINN_METRIC_COSINE = 1, MUNINN_METRIC_INNER_PRODUCT = 2 };

enum {
    MUNINN_CUDA_OK = 0,
    MUNINN_CUDA_E_ARG = -1,      /* bad argument (message in muninn_cuda_last_error) */
    MUNINN_CUDA_E_NODEVICE = -2, /* no usable CUDA device */
    MUNINN_CUDA_E_CUDA = -3,     /* a CUDA runtime call failed */
    MUNINN_CUDA_E_NOMEM = -4,    /* host or device allocation failed */
    MUNINN_CUDA_E_STATE = -5     /* e.g. search before upload */
};

typedef struct MuninnCudaIndex MuninnCudaIndex;

/* Per-query work counters. n_dist / n_expand are deterministic functions of (graph, query,
 * ef) and must equal the CPU oracle's; they also define the algorithmic
 * bytes of the roofline. Layout matches oracle/hnsw_oracle.c OrcStats. */
typedef struct {
    int64_t n_dist;         /* distance evaluations the reference would make (dist_func calls) */
    int64_t n_dist_upper;   /* of which in the greedy descent, src/hnsw_algo.c:257-282 */
    int32_t n_expand;       /* layer-0 candidates whose neighbour lists were scanned */
    int32_t n_greedy_steps; /* neighbour lists (re)scanned in the descent */
    int32_t term;           /* 0 candidates exhausted, 1 distance bound, 2 patience (src/hnsw_algo.c:382-392) */
    int32_t entry0;         /* slot at which the descent handed over to the layer-0 beam */
} MuninnSearchStats;

int muninn_cuda_device_count(void);
const char *muninn_cuda_version(void);
const char *muninn_cuda_strerror(int code);
/* "l2" | "cosine" | "inner_product" → metric, else -1 (vec_parse_metric). Host-only. */
int muninn_cuda_parse_metric(const char *name);

/* hnsw_create: dim, metric, M as in `CREATE VIRTUAL TABLE … USING hnsw_index(dimensions, metric, m, …)`.
 * M_max0 = 2*M (src/hnsw_algo.c:188). `device` is the CUDA ordinal. */
int muninn_cuda_index_create(int device, int dim, int metric, int M, MuninnCudaIndex **out);
void muninn_cuda_index_destroy(MuninnCudaIndex *idx);
/* Message of the last failure on this handle ("" if none). NULL handle → last create failure. */
const char *muninn_cuda_last_error(const MuninnCudaIndex *idx);

/*
 * Upload the graph exactly as load_index_from_shadow leaves it in host memory:
 *   n             stored rows of <name>_nodes, deleted ones included
 *   ids[n]        rowids, strictly ascending; the dense "slot" of a node is its rank here
 *   levels[n]     _nodes.level
 *   deleted[n]    _nodes.deleted (may be NULL = none)
 *   vectors       [n × dim] fp32, row-major, packed
 *   list_offsets  [L+1], L = Σ(levels[s]+1): lists are node-major, level-minor — list
 *                 (s, l) has index first(s)+l with first(s) = Σ_{t<s}(levels[t]+1)
 *   targets       slot per edge, in shadow-table order (ascending target_id within a list,
 *                 src/hnsw_vtab.c:163-169,322); -1 for an edge whose target row is missing
 *   entry_slot    slot of _config.entry_point, or -1 for an empty index
 *   max_level     _config.max_level
 * Device layout built from it: packed fp32 slab (row stride padded to 4 floats), layer-0
 * neighbour table [n × 2M] int32 (-1 padded), upper-layer lists [·× M], per-row squared
 * norms for cosine. A second upload replaces the first (the vtab's dirty-flag re-upload).
 */
int muninn_cuda_index_upload(MuninnCudaIndex *idx, int64_t n, const int64_t *ids, const int32_t *levels,
                             const uint8_t *deleted, const float *vectors, const int64_t *list_offsets,
                             const int32_t *targets, int32_t entry_slot, int32_t max_level);

/* Rows only — an index without a graph, for exact k-NN over a shard of base rows (BASELINE.json config 5; the
 * reference's brute_force_knn helper, test/test_hnsw_algo.c:10-34, needs nothing but the vectors). `rows` [n × dim] fp32
 * may be host OR device memory (a shard already resident on the device is adopted with one copy); row indices are
 * 0..n-1 and nothing is deleted. Search and patch refuse such a handle (MUNINN_CUDA_E_STATE). */
int muninn_cuda_index_upload_rows(MuninnCudaIndex *idx, int64_t n, const float *rows);

/*
 * Incremental form of muninn_cuda_index_upload for the vtab's xUpdate path, so that an INSERT or DELETE does not
 * re-send the whole index. The reference mutates little per call: hnsw_insert (src/hnsw_algo.c:520-666) creates
 * one node and rewrites only the lists of the nodes it links it to; hnsw_delete (:717-805) flags one node and
 * rewrites only the lists of its former neighbours (the node stays stored). The caller sends the records of
 * exactly those nodes, in the upload's format:
 *   n_total        stored rows after the change; new rows must carry ids above every stored id, so that they take
 *                  slots n_old..n_total-1 (otherwise slots shift: re-upload instead)
 *   n_patch        nodes whose record follows; every new node must be among them
 *   slots[n_patch] their slots; ids / levels / deleted / vectors [n_patch × dim]: their records (levels of stored
 *                  nodes cannot change, a deleted node cannot come back). A stored vector never changes in the
 *                  reference (no vector update in src/hnsw_algo.c; UPDATE is rejected, src/hnsw_vtab.c:782-783), so
 *                  only the vectors of NEW nodes are read: vectors[i*dim ..] for a record i of an already stored
 *                  node need not be addressable (the device-assisted builder puts the new nodes first and passes
 *                  a pointer into its base array)
 *   list_offsets, targets   their neighbour lists, node-major level-minor, as in muninn_cuda_index_upload
 *   entry_slot, max_level   the index's after the change
 * References to a newly deleted slot become holes in every list of the device copy, as upload stores them.
 * Synchronous. A search after the patch returns exactly what it would after a full upload of the new state.
 * Every record is validated before anything is changed: a patch the library refuses (MUNINN_CUDA_E_ARG) leaves the
 * handle answering as before. A CUDA failure half-way leaves it unusable until the next muninn_cuda_index_upload.
 */
int muninn_cuda_index_patch(MuninnCudaIndex *idx, int64_t n_total, int32_t n_patch, const int32_t *slots, const int64_t *ids,
                            const int32_t *levels, const uint8_t *deleted, const float *vectors, const int64_t *list_offsets,
                            const int32_t *targets, int32_t entry_slot, int32_t max_level);

/*
 * Batched hnsw_search. Host buffers; the call copies queries to the device, runs the
 * search kernel, copies results back and returns when they are in place.
 *   queries   [nq × dim] fp32
 *   k, ef     as hnsw_search; ef = max(ef, k) is applied here (src/hnsw_algo.c:673). The
 *             vtab's default ef = 2k (src/hnsw_vtab.c:591) stays in the vtab.
 *   out_ids   [nq × k] rowids ascending by distance; unused tail = -1
 *   out_dist  [nq × k] fp32 distances (L2 is squared, as in the reference); unused tail = 0
 *   out_count [nq] number of results per query (may be < k; 0 for an empty index)
 *   out_stats [nq] or NULL
 * Batches of 16 384 queries and more stream in behind ONE launch: the kernel starts on the first 8 192 queries while the
 * copy engine delivers the rest and raises a device-side arrival mark that the query teams check (results identical;
 * end-to-end throughput of an 80 000-query call rose from 0.92 to 0.96-0.99 of the device-resident rate).
 */
int muninn_cuda_search_batch(MuninnCudaIndex *idx, const float *queries, int nq, int k, int ef, int64_t *out_ids,
                             float *out_dist, int32_t *out_count, MuninnSearchStats *out_stats);

/* Same, all buffers device-resident on the handle's device; asynchronous on the handle's
 * stream (muninn_cuda_set_stream), no host synchronisation. d_stats may be NULL. */
int muninn_cuda_search_batch_device(MuninnCudaIndex *idx, const float *d_queries, int nq, int k, int ef,
                                    int64_t *d_ids, float *d_dist, int32_t *d_count, MuninnSearchStats *d_stats);

/*
 * muninn_cuda_search_batch_device without any allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it). Same arguments, same results (ids, distances and counts are bit-identical to
 * muninn_cuda_search_batch_device). The kernels' scratch — work counters, retry lists, per-team visited tables — is a
 * file-scope device array of 17 301 760 bytes (256 + 512 KiB + 16 MiB) that every GPU which loads the library holds until
 * the process exits; the call itself issues one 32-byte memset and the search launches on the handle's stream and calls
 * no cudaMalloc, cudaFree or allocating library routine. Limits, rejected with MUNINN_CUDA_E_ARG and a message in
 * muninn_cuda_last_error: nq <= 65 536 queries per call; indexes of at most 2 097 119 nodes at M = 16, in general
 * n + 2M + 1 <= 2 097 152 (the last-resort visited table of 2(n + 2M + 1) entries, rounded up to a power of two, must
 * fit the 4 Mi-entry arena; muninn_cuda_static_scratch_limits reports 2 096 896, safe for every M). Fewer query teams are
 * resident than with a grown arena (1 024 at ef = 64, M = 16), so large batches run slower than through
 * muninn_cuda_search_batch_device. ALL calls on one device share the array and it keeps the previous call's contents
 * (the call clears the counters; tables and lists are written before they are read): calls that overlap on the GPU —
 * two streams, two handles, two threads — race; serialise them on one stream.
 */
int muninn_cuda_search_batch_device_static(MuninnCudaIndex *idx, const float *d_queries, int nq, int k, int ef,
                                           int64_t *d_ids, float *d_dist, int32_t *d_count, MuninnSearchStats *d_stats);
/* The limits above, for callers that size their batches against them (any pointer may be NULL). No counterpart in the
 * reference; added with muninn_cuda_search_batch_device_static. */
int muninn_cuda_static_scratch_limits(int *max_nq, int64_t *max_nodes, int64_t *bytes);

/* Stream on which the handle launches (a cudaStream_t passed as void*; NULL = the handle's own). */
int muninn_cuda_set_stream(MuninnCudaIndex *idx, void *stream);
/* Device time of the handle's most recent search kernel, from CUDA events recorded around
 * the launch on its stream (synchronises on the end event). < 0 if none. */
float muninn_cuda_last_kernel_ms(MuninnCudaIndex *idx);
/* Number of kernels this handle has launched since creation (bench.py's gpu_launches). */
int64_t muninn_cuda_launch_count(const MuninnCudaIndex *idx);
/* Launch geometry knobs for tuning: warps cooperating on one query (1, 2 or 4; 0 = automatic: the widest
 * team that still keeps every query of the batch resident at once, so a lone MATCH gets 4 warps and a full
 * batch one warp per query) and query teams resident per SM (0 = automatic). Results do not depend on them. */
int muninn_cuda_set_tuning(MuninnCudaIndex *idx, int warps_per_query, int teams_per_sm);
/* Team width and kernel generation (7 = hnsw_search.cuh, 8 = hnsw_beam.cuh) the handle's most recent search launched with. */
int muninn_cuda_last_team_width(const MuninnCudaIndex *idx);
int muninn_cuda_last_kernel_gen(const MuninnCudaIndex *idx);
/* Diagnostic options (no reference counterpart; results do not depend on them): "kernel_gen" = 0 (default: by
 * shape), 7 (hnsw_search.cuh: one warp per query, 13-row tile) or 8 (hnsw_beam.cuh: teams of 1/2/4 warps, 16-row
 * batches); "vis_factor" = size of a team's
 * first-level visited table in units of ef·2M entries (default 2); "exact_qs" (1/0), "exact_qs_min_nq",
 * "exact_superblock_mb", "exact_auto_3x_min_nq", "exact_seed_dense" (1/0: the first chunk of rows evaluated densely
 * instead of filtered and gathered): shape of the exact k-NN path; "vis_in_smem" (1/0): teams of several warps keep
 * their first-level visited table in shared memory when it fits; "tiny_zero_copy" (1/0): host batches of at most
 * 64 KB are read and written by the kernel straight from pinned host memory, without copy operations. */
int muninn_cuda_set_option(MuninnCudaIndex *idx, const char *name, double value);

/*
 * Exact k-NN over the uploaded rows of this handle (deleted rows skipped), reference
 * distance arithmetic, output ascending by (distance, row index) — the tie rule of
 * test/test_hnsw_algo.c:10-34. out_idx holds ROW INDICES (slots) + index_base, not rowids,
 * so a base sharded over several handles merges by plain comparison.
 */
int muninn_cuda_exact_knn(MuninnCudaIndex *idx, const float *queries, int nq, int k, int64_t index_base,
                          int64_t *out_idx, float *out_dist, int32_t *out_count);
int muninn_cuda_exact_knn_device(MuninnCudaIndex *idx, const float *d_queries, int nq, int k, int64_t index_base,
                                 int64_t *d_idx, float *d_dist, int32_t *d_count);
/*
 * How exact k-NN is evaluated. The OUTPUT is identical in every mode (bit-exact ids and distances,
 * tests/test_gpu_exact_tc.py); the modes differ in how the n·nq distance evaluations are pruned.
 *   MUNINN_EXACT_CUDA_CORE   every pair in the reference's fp32 chain order on the CUDA cores
 *   MUNINN_EXACT_TC_3XTF32   tiled Q·Xᵀ on the tensor cores as 3xTF32 (tcgen05 + TMA + TMEM)
 *                            with a fused per-tile selection against each query's running k-th
 *                            distance; only the survivors are evaluated in the reference's order
 *   MUNINN_EXACT_TC_1XTF32   same with a single TF32 product and a correspondingly wider margin; the
 *                            MMA reads the fp32 rows as stored (no split pass: the base is read once)
 *   MUNINN_EXACT_TC_AUTO     (default) 1xTF32 for every batch size: below 1024 queries the rows stream once past all
 *                            queries (HBM-bound), above it the query-stationary filter keeps a 128-query tile in shared
 *                            memory and sweeps L2-sized super-blocks of rows; muninn_cuda_exact_stats reports the mode
 * tile_rows = base rows per CTA tile (0 = default 256, or 128). In the tensor-core modes the
 * device entry point synchronises once with the host before it returns.
 */
enum { MUNINN_EXACT_CUDA_CORE = 0, MUNINN_EXACT_TC_3XTF32 = 1, MUNINN_EXACT_TC_1XTF32 = 2, MUNINN_EXACT_TC_AUTO = 3 };
int muninn_cuda_set_exact_mode(MuninnCudaIndex *idx, int mode, int tile_rows);

/* Counters of the handle's most recent exact k-NN call. */
typedef struct {
    int32_t mode;             /* MUNINN_EXACT_* that ran */
    int32_t chunks;           /* row chunks (split → filter → re-rank rounds) */
    int64_t candidates;       /* (query, row) pairs that survived the filter and were re-ranked exactly */
    int32_t overflow_queries; /* queries whose survivor list overflowed; recomputed on the CUDA cores */
    float filter_ms;          /* Σ device time of the tensor-core filter launches (CUDA events) */
    float c_eps;              /* error constant of the filter: |d~ - d_ref| <= c_eps·(|q|² + |x|²), or c_eps for cosine;
                                 in terms of the contraction: |sum~ - sum(q·x)| <= c_eps/2 · (|q|² + |x|²) */
} MuninnExactStats;
int muninn_cuda_exact_stats(MuninnCudaIndex *idx, MuninnExactStats *out);

/* Error probe for the tests: the filter's tensor-core contraction sum(q·x) of every (query, row)
 * pair of rows [0, rows) → out_approx [nq × rows] (host buffers). Needs a tensor-core exact mode. */
int muninn_cuda_exact_filter_probe(MuninnCudaIndex *idx, const float *queries, int nq, int64_t rows, float *out_approx);

/* Merge `parts` per-shard result lists ([parts × nq × k], as gathered by ncclAllGather) into
 * the global top-k under the same (distance, index) order. Device buffers. */
int muninn_cuda_topk_merge_device(MuninnCudaIndex *idx, int parts, int nq, int k, const int64_t *d_idx_parts,
                                  const float *d_dist_parts, const int32_t *d_count_parts, int64_t *d_idx,
                                  float *d_dist, int32_t *d_count);

/*
 * Sharded exact k-NN: the base rows are split over `world` handles (one per GPU, one process per GPU),
 * every rank calls this with the SAME queries and the first global row index of its shard. Each rank scans its shard,
 * packs its lists into 64-bit keys (ordered distance << 32 | global row index), the lists are exchanged with ONE
 * ncclAllGather of nq·k·8 bytes per rank on the handle's stream, and every rank merges them under the (distance,
 * index) rule of test/test_hnsw_algo.c:10-34 — all ranks end with the identical global answer in d_idx / d_dist /
 * d_count (device buffers). Global row indices must fit 32 bits. `nccl_comm` is an ncclComm_t passed as void*: the
 * host's own (e.g. torch.distributed's) or one made by muninn_cuda_comm_create. NCCL is bound at run time from the
 * libnccl.so.2 already loaded in the process, else the system's.
 */
int muninn_cuda_exact_knn_sharded_device(MuninnCudaIndex *idx, void *nccl_comm, const float *d_queries, int nq, int k,
                                         int64_t index_base, int64_t *d_idx, float *d_dist, int32_t *d_count);
/* A communicator for hosts that have none: rank 0 calls _unique_id (128 bytes, ncclUniqueId) and ships the bytes to the
 * other ranks by any means; every rank then calls _create (collective). */
int muninn_cuda_comm_unique_id(void *out_id128);
int muninn_cuda_comm_create(int device, int world, int rank, const void *id128, void **out_comm);
void muninn_cuda_comm_destroy(void *comm);
/* The two halves of the exchange for callers that move the keys themselves: result lists → keys, and the merge of
 * `parts` gathered key lists [parts × nq × k]. Device buffers. */
int muninn_cuda_topk_pack_keys_device(MuninnCudaIndex *idx, int nq, int k, const int64_t *d_idx, const float *d_dist,
                                      const int32_t *d_count, uint64_t *d_keys);
int muninn_cuda_topk_merge_keys_device(MuninnCudaIndex *idx, int parts, int nq, int k, const uint64_t *d_keys,
                                       int64_t *d_idx, float *d_dist, int32_t *d_count);

/* n distances metric(a[i], b[i]) with the reference's summation order (src/vec_math.c:78-143).
 * Host buffers [n × dim]; used by the known-answer tests of test/test_vec_math.c. */
int muninn_cuda_distance_batch(int device, int metric, int dim, int64_t n, const float *a, const float *b, float *out);

/* The handle as the searcher of the host builder's group loop (include/muninn_build.h: muninn_build_run): fills
 * `out` with this handle and thin wrappers over muninn_cuda_index_upload / _index_patch / _search_batch, so that a bulk
 * build — the reference's `for each row: hnsw_insert(...)` (src/hnsw_algo.c:520-666) — runs its layer-0 candidate
 * searches on the device. Host-only call. */
struct MuninnBuildDevice;
int muninn_cuda_build_device(MuninnCudaIndex *idx, struct MuninnBuildDevice *out);

/* Diagnostic for the roofline report (no reference counterpart): GB/s this device delivers for
 * independent random gathers of row_bytes-byte rows out of a slab_bytes slab (slab >> L2) — the
 * access pattern of hnsw_search's neighbour-vector reads with nothing dependent in the way. */
int muninn_cuda_gather_ceiling(int device, int row_bytes, int64_t slab_bytes, float *out_gbps);

#ifdef __cplusplus
}
#endif
#endif /* MUNINN_CUDA_H */
