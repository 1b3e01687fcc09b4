/*
 * muninn_cuda.cu — libmuninn_cuda: handle management, graph upload and the C ABI declared in
 * include/muninn_cuda.h. Kernels live in hnsw_search.cuh and exact_knn.cuh.
 *
 * Upload mirrors what load_index_from_shadow (/root/reference/src/hnsw_vtab.c:286-341) leaves
 * in host memory, re-laid out for the device:
 *   vec        [n × ld] fp32, ld = dim rounded up to 4 (16-byte rows for 128-bit gathers)
 *   nbr0       [n × 2M] int32 layer-0 neighbour slots, -1 padded; a deleted or missing target
 *              is stored as -1 in place (the reference skips both, src/hnsw_algo.c:270,408)
 *   upper_base [n] first upper-layer list of a node (or -1); nbrU [lists × M], same convention
 *   norm2      [n] Σx² in the reference's summation order (cosine only)
 * There is no CPU path: without a device every entry point fails with MUNINN_CUDA_E_NODEVICE.
 */
#include "../../include/muninn_cuda.h"
#include "../../include/muninn_build.h" /* MuninnBuildDevice: the searcher interface of the host builder */
#include "exact_knn.cuh"
#include "exact_tc.cuh"
#include "hnsw_search.cuh"
#include "hnsw_beam.cuh"

#ifndef MUNINN_D128
#define MUNINN_D128 1 /* constant-trip-count instances of the search kernel for dim 96 / 128 / 768 */
#endif
#ifndef MUNINN_VIS_FACTOR
#define MUNINN_VIS_FACTOR 2
#endif

#include <algorithm>
#include <climits>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

using namespace muninn;

struct MuninnCudaIndex {
    int device = 0, dim = 0, ld = 0, metric = 0, M = 0, M0 = 0;
    int W0 = 0, WU = 0; /* device row widths of the layer-0 / upper-layer tables: 2M / M, or the longest stored list when the
                           reference's delete-time reconnection grew one past its nominal capacity (src/hnsw_algo.c:780-781) */
    long long n = 0, n_live = 0;
    int entry_slot = -1, max_level = -1;
    bool uploaded = false, rows_only = false; /* rows_only: muninn_cuda_index_upload_rows, exact k-NN only */
    /* host mirrors muninn_cuda_index_patch needs, and the device capacities it grows */
    std::vector<int32_t> h_levels, h_upper_base;
    std::vector<uint8_t> h_deleted;
    long long n_upper = 0, cap_n = 0, cap_upper = 0;
    int64_t last_id = 0;
    /* device graph */
    float *d_vec = nullptr, *d_norm2 = nullptr;
    int32_t *d_nbr0 = nullptr, *d_upper_base = nullptr, *d_nbrU = nullptr;
    uint8_t *d_deleted = nullptr;
    int64_t *d_ids = nullptr;
    /* launch state */
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaStream_t copy_in = nullptr; cudaEvent_t ev_first = nullptr; unsigned *d_arrived = nullptr, *h_arrived = nullptr; /* large host batches stream in behind the launch */
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool have_timing = false;
    int num_sms = 0, teams_per_sm = 0, warps_per_query = 0;
    int64_t launches = 0;
    unsigned *d_vis = nullptr; size_t vis_entries = 0;  /* visited arena: teams × cap 32-bit slots */
    int *d_ctl = nullptr;        /* [8]: work counters ×3, retry counts ×2 */
    int *d_retry = nullptr; size_t retry_cap = 0; /* 2 × nq */
    /* staging for the host-buffer entry points */
    void *d_stage = nullptr; size_t stage_bytes = 0;
    void *h_pinned = nullptr; size_t pinned_bytes = 0;
    size_t occ_smem[6] = {0, 0, 0, 0, 0, 0}; int occ_val[6] = {0, 0, 0, 0, 0, 0}; /* cached occupancy per (generation, team width[, table in shared memory]) */
    int vis_in_smem = 1; /* small-batch teams keep their first-level visited table in shared memory when it fits */
    int kernel_gen = 0, last_W = 1, last_gen = 7; double vis_factor = MUNINN_VIS_FACTOR; /* kernel_gen 0 = by shape */
    int tiny_zero_copy = 1; /* host-buffer calls of ≤ 64 KB: the kernel reads and writes pinned host memory directly */
    /* exact k-NN scratch */
    float *d_panel = nullptr; size_t panel_floats = 0;
    unsigned long long *d_topk = nullptr; int *d_topk_size = nullptr; size_t topk_cap = 0, topk_q = 0;
    /* exact k-NN, tensor-core filter (exact_tc.cuh) */
    int exact_mode = MUNINN_EXACT_TC_AUTO, tc_tile_rows = 256;
    int tc_qs = 1, tc_qs_min_nq = 1024, tc_superblock_mb = 32, tc_x_stride = 0; /* query-stationary 1xTF32 filter (exact_tc_filter_qs) */
    int tc_growth = 0; /* large batches: chunk growth factor − 1 (0 = default 2: the rows done grow ×3 per round) */
    int tc_seed_dense = 1; /* large batches: the first chunk (every row survives) is evaluated densely, not filtered + gathered */
    int tc_pair = 0; CUtensorMap tc_map_xh; /* CTA-pair form of the query-stationary filter: row-tile map with 128-row boxes */
    int tc_auto_1x_min_nq = 1 << 30; /* AUTO: batches of at least this many queries still take 3xTF32 (none by default) */
    int exact_mode_run = MUNINN_EXACT_TC_3XTF32; /* what AUTO resolved to for the current call */
    float *d_qsplit = nullptr; size_t qsplit_floats = 0;   /* q_hi | q_lo, [nq × kp] each */
    float *d_xsplit = nullptr; size_t xsplit_floats = 0;   /* x_hi | x_lo of one chunk */
    float *d_qaux = nullptr, *d_tau = nullptr; int *d_cand_count = nullptr, *d_overflow = nullptr, *d_qmap = nullptr;
    size_t tcq_cap = 0;
    int *d_cand = nullptr; size_t cand_entries = 0;
    float *d_qgather = nullptr; size_t qgather_floats = 0;
    float *d_dbg = nullptr; size_t dbg_floats = 0;
    unsigned long long *d_tc_ctr = nullptr;                /* [0] survivors re-ranked, [1] flagged queries (int) */
    std::vector<cudaEvent_t> tc_ev;
    MuninnExactStats exact_stats = {};
    unsigned long long *d_shard_keys = nullptr; size_t shard_keys_cap = 0; /* sharded exact k-NN: own keys | gathered keys */
    std::string err;
};

/* the last-resort visited table holds 2·(n + row width + 1) 32-bit entries rounded up to a power of two: beyond
 * 2^30 nodes that size no longer fits the 32-bit capacity the kernels index with */
static const long long MAX_NODES = (1LL << 30) - 1024;

static thread_local std::string g_create_err;

static int fail(MuninnCudaIndex *h, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf; else g_create_err = buf;
    return code;
}

#define CK(h, call)                                                                                       \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess)                                                                            \
            return fail(h, e_ == cudaErrorMemoryAllocation ? MUNINN_CUDA_E_NOMEM : MUNINN_CUDA_E_CUDA,    \
                        "%s: %s", #call, cudaGetErrorString(e_));                                        \
    } while (0)

extern "C" int muninn_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" const char *muninn_cuda_version(void) { return "muninn-b200 0.1 (sm_100a)"; }

extern "C" const char *muninn_cuda_strerror(int code) {
    switch (code) {
    case MUNINN_CUDA_OK: return "ok";
    case MUNINN_CUDA_E_ARG: return "invalid argument";
    case MUNINN_CUDA_E_NODEVICE: return "no CUDA device";
    case MUNINN_CUDA_E_CUDA: return "CUDA runtime error";
    case MUNINN_CUDA_E_NOMEM: return "out of memory";
    case MUNINN_CUDA_E_STATE: return "index not uploaded";
    }
    return "unknown error";
}

/* vec_parse_metric, /root/reference/src/vec_math.c:192-204 */
extern "C" int muninn_cuda_parse_metric(const char *name) {
    if (!name) return -1;
    if (!strcmp(name, "l2")) return MUNINN_METRIC_L2;
    if (!strcmp(name, "cosine")) return MUNINN_METRIC_COSINE;
    if (!strcmp(name, "inner_product")) return MUNINN_METRIC_INNER_PRODUCT;
    return -1;
}

extern "C" const char *muninn_cuda_last_error(const MuninnCudaIndex *h) {
    return h ? h->err.c_str() : g_create_err.c_str();
}

static void free_graph(MuninnCudaIndex *h) {
    cudaFree(h->d_vec); cudaFree(h->d_norm2); cudaFree(h->d_nbr0); cudaFree(h->d_upper_base);
    cudaFree(h->d_nbrU); cudaFree(h->d_deleted); cudaFree(h->d_ids);
    h->d_vec = h->d_norm2 = nullptr; h->d_nbr0 = h->d_upper_base = h->d_nbrU = nullptr;
    h->d_deleted = nullptr; h->d_ids = nullptr;
    h->uploaded = false; h->rows_only = false;
}

extern "C" int muninn_cuda_index_create(int device, int dim, int metric, int M, MuninnCudaIndex **out) {
    if (!out) return fail(nullptr, MUNINN_CUDA_E_ARG, "out is NULL");
    *out = nullptr;
    if (dim < 1) return fail(nullptr, MUNINN_CUDA_E_ARG, "dimensions must be >= 1 (got %d)", dim);
    if (metric < 0 || metric > 2) return fail(nullptr, MUNINN_CUDA_E_ARG, "unknown metric %d", metric);
    if (M < 2 || 2 * M > MAX_M0) return fail(nullptr, MUNINN_CUDA_E_ARG, "m must be in [2, %d] (got %d)", MAX_M0 / 2, M);
    int ndev = muninn_cuda_device_count();
    if (ndev <= 0) return fail(nullptr, MUNINN_CUDA_E_NODEVICE, "no CUDA device visible; libmuninn_cuda has no CPU path");
    if (device < 0 || device >= ndev) return fail(nullptr, MUNINN_CUDA_E_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    MuninnCudaIndex *h = new (std::nothrow) MuninnCudaIndex();
    if (!h) return fail(nullptr, MUNINN_CUDA_E_NOMEM, "host allocation failed");
    h->device = device; h->dim = dim; h->ld = (dim + 3) & ~3; h->metric = metric; h->M = M; h->M0 = 2 * M; h->W0 = h->M0; h->WU = M;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev1);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_ctl, 8 * sizeof(int));
    if (e != cudaSuccess) {
        int rc = fail(nullptr, MUNINN_CUDA_E_CUDA, "device %d setup: %s", device, cudaGetErrorString(e));
        delete h;
        return rc;
    }
    h->stream = h->own_stream;
    /* experiment knobs (A/B runs of one build): same meaning as muninn_cuda_set_option / _set_tuning */
    if (const char *e = getenv("MUNINN_VIS_SMEM")) h->vis_in_smem = atoi(e) != 0;
    if (const char *e = getenv("MUNINN_EXACT_GROWTH")) h->tc_growth = std::max(0, std::min(31, atoi(e)));
    if (const char *e = getenv("MUNINN_EXACT_SEED_DENSE")) h->tc_seed_dense = atoi(e) != 0;
    if (const char *e = getenv("MUNINN_TINY_ZERO_COPY")) h->tiny_zero_copy = atoi(e) != 0;
    if (const char *e = getenv("MUNINN_KERNEL")) { const int g = atoi(e); if (g == 0 || g == 7 || g == 8) h->kernel_gen = g; }
    if (const char *e = getenv("MUNINN_VIS_FACTOR")) { const double f = atof(e); if (f >= 0.5 && f <= 64.0) h->vis_factor = f; }
    if (const char *e = getenv("MUNINN_WARPS")) { const int w = atoi(e); if (w == 0 || w == 1 || w == 2 || w == 4) h->warps_per_query = w; }
    if (const char *e = getenv("MUNINN_EXACT_QS")) h->tc_qs = atoi(e) != 0;
    if (const char *e = getenv("MUNINN_EXACT_PAIR")) h->tc_pair = atoi(e) != 0;
    if (const char *e = getenv("MUNINN_EXACT_SB_MB")) { const int v = atoi(e); if (v >= 1 && v <= 4096) h->tc_superblock_mb = v; }
    if (const char *e = getenv("MUNINN_TEAMS_PER_SM")) { const int t = atoi(e); if (t >= 0 && t <= 32) h->teams_per_sm = t; }
    *out = h;
    return MUNINN_CUDA_OK;
}

extern "C" void muninn_cuda_index_destroy(MuninnCudaIndex *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    free_graph(h);
    cudaFree(h->d_vis); cudaFree(h->d_ctl); cudaFree(h->d_retry); cudaFree(h->d_stage);
    cudaFree(h->d_panel); cudaFree(h->d_topk); cudaFree(h->d_topk_size);
    cudaFree(h->d_qsplit); cudaFree(h->d_xsplit); cudaFree(h->d_qaux); cudaFree(h->d_tau); cudaFree(h->d_cand_count);
    cudaFree(h->d_overflow); cudaFree(h->d_qmap); cudaFree(h->d_cand); cudaFree(h->d_qgather); cudaFree(h->d_dbg);
    cudaFree(h->d_tc_ctr); cudaFree(h->d_shard_keys);
    for (cudaEvent_t e : h->tc_ev) cudaEventDestroy(e);
    if (h->h_pinned) cudaFreeHost(h->h_pinned);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->copy_in) { cudaStreamDestroy(h->copy_in); cudaEventDestroy(h->ev_first); cudaFree(h->d_arrived); cudaFreeHost(h->h_arrived); }
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
}

extern "C" int muninn_cuda_set_stream(MuninnCudaIndex *h, void *stream) {
    if (!h) return MUNINN_CUDA_E_ARG;
    h->stream = stream ? (cudaStream_t)stream : h->own_stream;
    return MUNINN_CUDA_OK;
}

extern "C" int64_t muninn_cuda_launch_count(const MuninnCudaIndex *h) { return h ? h->launches : 0; }

extern "C" int muninn_cuda_set_tuning(MuninnCudaIndex *h, int warps_per_query, int teams_per_sm) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (warps_per_query != 0 && warps_per_query != 1 && warps_per_query != 2 && warps_per_query != 4)
        return fail(h, MUNINN_CUDA_E_ARG, "warps_per_query must be 0 (automatic), 1, 2 or 4 (got %d)", warps_per_query);
    if (teams_per_sm < 0 || teams_per_sm > 32) return fail(h, MUNINN_CUDA_E_ARG, "teams_per_sm must be in [0, 32]");
    h->warps_per_query = warps_per_query;
    h->teams_per_sm = teams_per_sm;
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_set_option(MuninnCudaIndex *h, const char *name, double value) {
    if (!h || !name) return MUNINN_CUDA_E_ARG;
    if (!strcmp(name, "kernel_gen")) {
        if (value != 0 && value != 7 && value != 8) return fail(h, MUNINN_CUDA_E_ARG, "kernel_gen must be 0 (by shape), 7 or 8");
        h->kernel_gen = (int)value;
    } else if (!strcmp(name, "vis_factor")) {
        if (!(value >= 0.5 && value <= 64.0)) return fail(h, MUNINN_CUDA_E_ARG, "vis_factor must be in [0.5, 64]");
        h->vis_factor = value;
    } else if (!strcmp(name, "vis_in_smem")) { /* 0 = the teams' visited tables stay in the global arena */
        h->vis_in_smem = value != 0;
    } else if (!strcmp(name, "tiny_zero_copy")) { /* 0 = stage tiny host batches through copies, as larger ones */
        h->tiny_zero_copy = value != 0;
    } else if (!strcmp(name, "exact_qs")) { /* 0 = the streaming 1xTF32 filter for every batch size */
        h->tc_qs = value != 0;
    } else if (!strcmp(name, "exact_seed_dense")) { /* 0 = the first chunk goes through the filter and the re-rank like the others */
        h->tc_seed_dense = value != 0;
    } else if (!strcmp(name, "exact_pair")) { /* 1 = CTA pairs (tcgen05 cta_group::2) in the query-stationary filter */
        h->tc_pair = value != 0;
    } else if (!strcmp(name, "exact_qs_min_nq")) {
        h->tc_qs_min_nq = (int)value;
    } else if (!strcmp(name, "exact_superblock_mb")) {
        if (!(value >= 1 && value <= 4096)) return fail(h, MUNINN_CUDA_E_ARG, "exact_superblock_mb must be in [1, 4096]");
        h->tc_superblock_mb = (int)value;
    } else if (!strcmp(name, "exact_auto_3x_min_nq")) {
        h->tc_auto_1x_min_nq = value <= 0 ? (1 << 30) : (int)value;
    } else return fail(h, MUNINN_CUDA_E_ARG, "unknown option %s", name);
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_last_team_width(const MuninnCudaIndex *h) { return h ? h->last_W : 0; }
extern "C" int muninn_cuda_last_kernel_gen(const MuninnCudaIndex *h) { return h ? h->last_gen : 0; }

extern "C" float muninn_cuda_last_kernel_ms(MuninnCudaIndex *h) {
    if (!h || !h->have_timing) return -1.0f;
    cudaSetDevice(h->device);
    if (cudaEventSynchronize(h->ev1) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) return -1.0f;
    return ms;
}

/* ───────────────────────── upload ───────────────────────── */

template <typename T>
static cudaError_t to_device(T **dptr, const T *src, size_t count, cudaStream_t s) {
    cudaError_t e = cudaMalloc((void **)dptr, std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    if (count) e = cudaMemcpyAsync(*dptr, src, count * sizeof(T), cudaMemcpyHostToDevice, s);
    return e;
}

extern "C" int muninn_cuda_index_upload(MuninnCudaIndex *h, int64_t n, const int64_t *ids, const int32_t *levels,
                                        const uint8_t *deleted, const float *vectors, const int64_t *list_offsets,
                                        const int32_t *targets, int32_t entry_slot, int32_t max_level) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (n < 0 || n > MAX_NODES) return fail(h, MUNINN_CUDA_E_ARG, "n = %lld out of range (at most %lld nodes)", (long long)n, (long long)MAX_NODES);
    if (n > 0 && (!ids || !levels || !vectors || !list_offsets)) return fail(h, MUNINN_CUDA_E_ARG, "NULL array");
    if (entry_slot < -1 || entry_slot >= n) return fail(h, MUNINN_CUDA_E_ARG, "entry_slot %d out of range", entry_slot);
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaStreamSynchronize(h->stream));
    free_graph(h);

    /* row widths: the nominal capacities, or the longest stored list. hnsw_delete's reconnection step appends to a
     * neighbour's list with no capacity check (src/hnsw_algo.c:780-781), so after heavy deletion the reference holds —
     * and keeps answering from — lists longer than 2M / M; the device tables are sized to hold what it holds */
    int W0 = h->M0, WU = h->M;
    {
        int64_t li0 = 0;
        for (int64_t s = 0; s < n; s++) {
            if (levels[s] < 0 || levels[s] >= 32) /* HNSW_MAX_LEVELS, src/hnsw_algo.h:14 */
                return fail(h, MUNINN_CUDA_E_ARG, "level %d of slot %lld out of range", levels[s], (long long)s);
            for (int l = 0; l <= levels[s]; l++, li0++) {
                const int64_t len = list_offsets[li0 + 1] - list_offsets[li0];
                if (len < 0) return fail(h, MUNINN_CUDA_E_ARG, "list offsets must not decrease");
                if (len > MAX_M0) return fail(h, MUNINN_CUDA_E_ARG, "list (slot %lld, level %d) has %lld entries; at most %d are supported", (long long)s, l, (long long)len, MAX_M0);
                if (l == 0 && len > W0) W0 = (int)((len + 3) & ~3);
                if (l > 0 && len > WU) WU = (int)((len + 3) & ~3);
            }
        }
    }
    const int M = WU, M0 = W0; /* from here on: device row widths */
    /* host re-layout */
    std::vector<int32_t> nbr0((size_t)n * M0, -1), upper_base((size_t)n, -1);
    std::vector<uint8_t> del((size_t)n, 0);
    long long n_upper = 0, n_live = 0;
    for (int64_t s = 0; s < n; s++) {
        if (s > 0 && ids[s] <= ids[s - 1]) return fail(h, MUNINN_CUDA_E_ARG, "ids must be strictly ascending");
        if (levels[s] > 0) { upper_base[s] = (int32_t)n_upper; n_upper += levels[s]; }
        del[s] = deleted ? (deleted[s] != 0) : 0;
        if (!del[s]) n_live++;
        if (n_upper > 0x7ffffff0LL) return fail(h, MUNINN_CUDA_E_ARG, "too many upper-layer lists");
    }
    if (entry_slot >= 0 && levels[entry_slot] < max_level)
        return fail(h, MUNINN_CUDA_E_ARG, "entry node level %d < max_level %d", levels[entry_slot], max_level);
    if (entry_slot >= 0 && (max_level < 0 || max_level >= 32)) return fail(h, MUNINN_CUDA_E_ARG, "bad max_level");
    std::vector<int32_t> nbrU((size_t)n_upper * M, -1);
    int64_t li = 0;
    for (int64_t s = 0; s < n; s++) {
        for (int l = 0; l <= levels[s]; l++, li++) {
            const int64_t b = list_offsets[li], e = list_offsets[li + 1];
            int32_t *dst = l == 0 ? &nbr0[(size_t)s * M0] : &nbrU[((size_t)upper_base[s] + (l - 1)) * M];
            for (int64_t j = b; j < e; j++) {
                int32_t t = targets[j];
                if (t < -1 || t >= n) return fail(h, MUNINN_CUDA_E_ARG, "target %d out of range", t);
                if (t >= 0 && levels[t] < l) /* the reference would index neighbors[l] out of bounds */
                    return fail(h, MUNINN_CUDA_E_ARG, "edge to slot %d at level %d above its level %d", t, l, levels[t]);
                dst[j - b] = (t >= 0 && !del[t]) ? t : -1;
            }
        }
    }

    cudaStream_t st = h->stream;
    CK(h, cudaMalloc((void **)&h->d_vec, std::max<size_t>((size_t)n * h->ld, 4) * sizeof(float)));
    if (n) {
        if (h->ld != h->dim) CK(h, cudaMemsetAsync(h->d_vec, 0, (size_t)n * h->ld * sizeof(float), st));
        CK(h, cudaMemcpy2DAsync(h->d_vec, (size_t)h->ld * 4, vectors, (size_t)h->dim * 4, (size_t)h->dim * 4, (size_t)n,
                                cudaMemcpyHostToDevice, st));
    }
    CK(h, to_device(&h->d_nbr0, nbr0.data(), nbr0.size(), st));
    CK(h, to_device(&h->d_upper_base, upper_base.data(), upper_base.size(), st));
    CK(h, to_device(&h->d_nbrU, nbrU.data(), nbrU.size(), st));
    CK(h, to_device(&h->d_deleted, del.data(), del.size(), st));
    CK(h, to_device(&h->d_ids, ids, (size_t)n, st));
    CK(h, cudaMalloc((void **)&h->d_norm2, std::max<size_t>((size_t)n, 1) * sizeof(float)));
    if (n) {
        row_norm2_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(h->d_vec, h->ld, h->dim, n, h->d_norm2);
        h->launches++;
        CK(h, cudaGetLastError());
    }
    CK(h, cudaStreamSynchronize(st)); /* host vectors above go out of scope */
    h->n = n; h->n_live = n_live; h->entry_slot = entry_slot; h->max_level = max_level;
    h->W0 = W0; h->WU = WU;
    h->h_levels.assign(levels, levels + n); h->h_upper_base = upper_base; h->h_deleted = del;
    h->n_upper = n_upper; h->cap_n = std::max<long long>(n, 1); h->cap_upper = std::max<long long>(n_upper, 1);
    h->last_id = n ? ids[n - 1] : 0;
    h->uploaded = true;
    return MUNINN_CUDA_OK;
}

/* Rows only: an index without a graph, for exact k-NN over a shard of base rows (BASELINE config 5). `rows` may be
 * host or device memory (unified addressing): a shard generated or already resident on the device is adopted with
 * one device-to-device copy. Row indices are 0..n-1, nothing is deleted. */
extern "C" int muninn_cuda_index_upload_rows(MuninnCudaIndex *h, int64_t n, const float *rows) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (n < 0 || n > 0x7ffffff0LL) return fail(h, MUNINN_CUDA_E_ARG, "n = %lld out of range", (long long)n);
    if (n > 0 && !rows) return fail(h, MUNINN_CUDA_E_ARG, "NULL array");
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaStreamSynchronize(h->stream));
    free_graph(h);
    cudaStream_t st = h->stream;
    CK(h, cudaMalloc((void **)&h->d_vec, std::max<size_t>((size_t)n * h->ld, 4) * sizeof(float)));
    CK(h, cudaMalloc((void **)&h->d_norm2, std::max<size_t>((size_t)n, 1) * sizeof(float)));
    if (n) {
        if (h->ld != h->dim) CK(h, cudaMemsetAsync(h->d_vec, 0, (size_t)n * h->ld * sizeof(float), st));
        CK(h, cudaMemcpy2DAsync(h->d_vec, (size_t)h->ld * 4, rows, (size_t)h->dim * 4, (size_t)h->dim * 4, (size_t)n, cudaMemcpyDefault, st));
        row_norm2_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(h->d_vec, h->ld, h->dim, n, h->d_norm2);
        h->launches++;
        CK(h, cudaGetLastError());
    }
    CK(h, cudaStreamSynchronize(st));
    h->n = n; h->n_live = n; h->entry_slot = -1; h->max_level = -1;
    h->h_levels.clear(); h->h_upper_base.clear(); h->h_deleted.clear();
    h->n_upper = 0; h->cap_n = std::max<long long>(n, 1); h->cap_upper = 1; h->last_id = n - 1;
    h->W0 = h->M0; h->WU = h->M;
    h->uploaded = true; h->rows_only = true;
    return MUNINN_CUDA_OK;
}

/* ───────────────────────── incremental update ───────────────────────── */

/* every reference to a slot flagged deleted becomes a hole (-1), as upload stores it: one pass over the lists,
 * one L2-resident flag lookup per live entry, independent of how many slots were deleted */
__global__ void purge_slots_kernel(int32_t *__restrict__ lists, long long count, const uint8_t *__restrict__ deleted) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        const int32_t t = lists[i];
        if (t >= 0 && deleted[t]) lists[i] = -1;
    }
}

/* staged patch records → their places: `width` ints of record r go to lists[at[r] .. at[r] + width) */
__global__ void patch_scatter_rows_kernel(int n, const long long *__restrict__ at, const int32_t *__restrict__ rows, int width,
                                          int32_t *__restrict__ lists) {
    const long long total = (long long)n * width;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        lists[at[i / width] + i % width] = rows[i];
}
__global__ void patch_scatter_flags_kernel(int n, const int32_t *__restrict__ slots, const uint8_t *__restrict__ flags,
                                           uint8_t *__restrict__ deleted) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) deleted[slots[i]] = flags[i];
}

template <typename T>
static cudaError_t regrow(T **ptr, size_t keep, size_t newcount, cudaStream_t st) {
    T *np = nullptr;
    cudaError_t e = cudaMalloc((void **)&np, std::max<size_t>(newcount, 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    if (keep) e = cudaMemcpyAsync(np, *ptr, keep * sizeof(T), cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(*ptr);
    *ptr = np;
    return e;
}

static int ensure_stage(MuninnCudaIndex *h, size_t bytes);
static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" int muninn_cuda_index_patch(MuninnCudaIndex *h, int64_t n_total, int32_t n_patch, const int32_t *slots,
                                       const int64_t *ids, const int32_t *levels, const uint8_t *deleted, const float *vectors,
                                       const int64_t *list_offsets, const int32_t *targets, int32_t entry_slot,
                                       int32_t max_level) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!h->uploaded || h->rows_only) return fail(h, MUNINN_CUDA_E_STATE, "patch before muninn_cuda_index_upload");
    if (n_total < h->n || n_total > MAX_NODES || n_patch < 0) return fail(h, MUNINN_CUDA_E_ARG, "bad n_total / n_patch");
    if (n_patch > 0 && (!slots || !ids || !levels || !vectors || !list_offsets)) return fail(h, MUNINN_CUDA_E_ARG, "NULL array");
    if (entry_slot < -1 || entry_slot >= n_total) return fail(h, MUNINN_CUDA_E_ARG, "entry_slot %d out of range", entry_slot);
    CK(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    const int M = h->WU, M0 = h->W0; /* device row widths (muninn_cuda_index_upload) */
    const long long n_old = h->n, n_new = n_total - n_old;

    /* ── phase 1: validate everything; the handle is not touched until the whole patch is known to be sound ── */
    std::vector<int32_t> order(n_patch);
    for (int i = 0; i < n_patch; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return slots[a] < slots[b]; });
    std::vector<int32_t> new_levels((size_t)n_new);
    std::vector<uint8_t> new_deleted((size_t)n_new);
    std::vector<int32_t> new_rec((size_t)n_new); /* record index of new slot n_old + j */
    long long expect_new = n_old, add_upper = 0, n_upper_rows = 0;
    int64_t prev_id = h->last_id;
    for (int oi = 0; oi < n_patch; oi++) {
        const int i = order[oi];
        const long long sl = slots[i];
        if (sl < 0 || sl >= n_total || (oi > 0 && slots[order[oi - 1]] == sl)) return fail(h, MUNINN_CUDA_E_ARG, "bad or duplicate slot %lld", sl);
        if (levels[i] < 0 || levels[i] >= 32) return fail(h, MUNINN_CUDA_E_ARG, "level out of range");
        const uint8_t d = deleted ? (deleted[i] != 0) : 0;
        if (sl >= n_old) { /* new slots must be exactly n_old..n_total-1, ids ascending past the stored maximum */
            if (sl != expect_new++) return fail(h, MUNINN_CUDA_E_ARG, "new nodes must fill slots %lld..%lld", n_old, (long long)n_total - 1);
            if (sl > 0 && ids[i] <= prev_id) return fail(h, MUNINN_CUDA_E_ARG, "new ids must ascend past the stored ones");
            prev_id = ids[i];
            add_upper += levels[i];
            new_levels[sl - n_old] = levels[i]; new_deleted[sl - n_old] = d; new_rec[sl - n_old] = i;
        } else {
            if (levels[i] != h->h_levels[sl]) return fail(h, MUNINN_CUDA_E_ARG, "level of slot %lld changed", sl);
            if (!d && h->h_deleted[sl]) return fail(h, MUNINN_CUDA_E_ARG, "slot %lld cannot be undeleted", sl);
        }
        n_upper_rows += levels[i];
    }
    if (expect_new != n_total) return fail(h, MUNINN_CUDA_E_ARG, "patch does not contain every new node");
    auto level_of = [&](int32_t t) { return t < n_old ? h->h_levels[t] : new_levels[t - n_old]; };
    if (entry_slot >= 0 && level_of(entry_slot) < max_level) return fail(h, MUNINN_CUDA_E_ARG, "entry node level < max_level");
    /* deleted flags as they will stand after the patch, for the hole rule below */
    std::vector<std::pair<int32_t, uint8_t>> newly_dead; /* (slot, 1) of old slots that turn deleted */
    for (int i = 0; i < n_patch; i++)
        if (slots[i] < n_old && deleted && deleted[i] && !h->h_deleted[slots[i]]) newly_dead.push_back({slots[i], 1});
    std::sort(newly_dead.begin(), newly_dead.end());
    auto dead_after = [&](int32_t t) -> bool {
        if (t >= n_old) return new_deleted[t - n_old] != 0;
        if (h->h_deleted[t]) return true;
        return std::binary_search(newly_dead.begin(), newly_dead.end(), std::make_pair(t, (uint8_t)1));
    };
    /* staging: level-0 rows [n_patch × M0], upper rows [n_upper_rows × M] with their destinations */
    std::vector<int32_t> rows0((size_t)n_patch * M0, -1), rowsU((size_t)n_upper_rows * M, -1);
    std::vector<long long> at0((size_t)n_patch), atU((size_t)n_upper_rows);
    std::vector<int32_t> ub_new((size_t)n_new, -1);
    {
        long long nu = h->n_upper;
        for (long long j = 0; j < n_new; j++) { if (new_levels[j] > 0) { ub_new[j] = (int32_t)nu; nu += new_levels[j]; } }
    }
    int64_t li = 0;
    long long ur = 0;
    for (int i = 0; i < n_patch; i++) {
        const int32_t sl = slots[i];
        at0[i] = (long long)sl * M0;
        const long long ub = sl < n_old ? h->h_upper_base[sl] : ub_new[sl - n_old];
        for (int l = 0; l <= levels[i]; l++, li++) {
            const int64_t b = list_offsets[li], e = list_offsets[li + 1];
            const int capl = l == 0 ? M0 : M;
            if (e < b || e - b > capl) /* a list that outgrew the device rows (delete-time reconnection): the caller re-uploads */
                return fail(h, MUNINN_CUDA_E_ARG, "list (slot %d, level %d) has %lld entries, device rows hold %d: upload again", sl, l, (long long)(e - b), capl);
            if (e > b && !targets) return fail(h, MUNINN_CUDA_E_ARG, "NULL array");
            int32_t *dst = l == 0 ? &rows0[(size_t)i * M0] : &rowsU[(size_t)ur * M];
            for (int64_t j = b; j < e; j++) {
                const int32_t t = targets[j];
                if (t < -1 || t >= n_total) return fail(h, MUNINN_CUDA_E_ARG, "target %d out of range", t);
                if (t >= 0 && level_of(t) < l) return fail(h, MUNINN_CUDA_E_ARG, "edge to slot %d at level %d above its level", t, l);
                dst[j - b] = (t >= 0 && !dead_after(t)) ? t : -1;
            }
            if (l > 0) atU[ur++] = (ub + (l - 1)) * M;
        }
    }

    /* ── phase 2: grow, then move the records (a handful of bulk copies + scatter kernels) ── */
    if (n_total > h->cap_n) {
        const long long nc = n_total + n_total / 2 + 1024;
        CK(h, regrow(&h->d_vec, (size_t)n_old * h->ld, (size_t)nc * h->ld, st));
        CK(h, regrow(&h->d_nbr0, (size_t)n_old * M0, (size_t)nc * M0, st));
        CK(h, regrow(&h->d_upper_base, (size_t)n_old, (size_t)nc, st));
        CK(h, regrow(&h->d_norm2, (size_t)n_old, (size_t)nc, st));
        CK(h, regrow(&h->d_deleted, (size_t)n_old, (size_t)nc, st));
        CK(h, regrow(&h->d_ids, (size_t)n_old, (size_t)nc, st));
        h->cap_n = nc;
    }
    if (h->n_upper + add_upper > h->cap_upper) {
        const long long nc = (h->n_upper + add_upper) * 3 / 2 + 1024;
        CK(h, regrow(&h->d_nbrU, (size_t)h->n_upper * M, (size_t)nc * M, st));
        h->cap_upper = nc;
    }
    std::vector<uint8_t> flags((size_t)n_patch);
    for (int i = 0; i < n_patch; i++) flags[i] = deleted ? (deleted[i] != 0) : 0;
    const size_t b_at0 = align256((size_t)n_patch * 8), b_rows0 = align256((size_t)n_patch * M0 * 4),
                 b_atU = align256((size_t)n_upper_rows * 8), b_rowsU = align256((size_t)n_upper_rows * M * 4),
                 b_slots = align256((size_t)n_patch * 4), b_flags = align256((size_t)n_patch);
    int rc = ensure_stage(h, b_at0 + b_rows0 + b_atU + b_rowsU + b_slots + b_flags + 256);
    if (rc) return rc;
    char *sb = (char *)h->d_stage;
    long long *d_at0 = (long long *)sb; sb += b_at0;
    int32_t *d_rows0 = (int32_t *)sb; sb += b_rows0;
    long long *d_atU = (long long *)sb; sb += b_atU;
    int32_t *d_rowsU = (int32_t *)sb; sb += b_rowsU;
    int32_t *d_slots = (int32_t *)sb; sb += b_slots;
    uint8_t *d_flags = (uint8_t *)sb;
    std::vector<float> vec_new;
    std::vector<int64_t> ids_new((size_t)n_new);
    if (n_patch > 0) {
        CK(h, cudaMemcpyAsync(d_at0, at0.data(), (size_t)n_patch * 8, cudaMemcpyHostToDevice, st));
        CK(h, cudaMemcpyAsync(d_rows0, rows0.data(), (size_t)n_patch * M0 * 4, cudaMemcpyHostToDevice, st));
        CK(h, cudaMemcpyAsync(d_slots, slots, (size_t)n_patch * 4, cudaMemcpyHostToDevice, st));
        CK(h, cudaMemcpyAsync(d_flags, flags.data(), (size_t)n_patch, cudaMemcpyHostToDevice, st));
        const long long tot = (long long)n_patch * M0;
        patch_scatter_rows_kernel<<<(unsigned)std::min<long long>((tot + 255) / 256, 4096), 256, 0, st>>>(n_patch, d_at0, d_rows0, M0, h->d_nbr0);
        patch_scatter_flags_kernel<<<(unsigned)((n_patch + 255) / 256), 256, 0, st>>>(n_patch, d_slots, d_flags, h->d_deleted);
        h->launches += 2;
    }
    if (n_upper_rows > 0) {
        CK(h, cudaMemcpyAsync(d_atU, atU.data(), (size_t)n_upper_rows * 8, cudaMemcpyHostToDevice, st));
        CK(h, cudaMemcpyAsync(d_rowsU, rowsU.data(), (size_t)n_upper_rows * M * 4, cudaMemcpyHostToDevice, st));
        const long long tot = n_upper_rows * M;
        patch_scatter_rows_kernel<<<(unsigned)std::min<long long>((tot + 255) / 256, 4096), 256, 0, st>>>((int)n_upper_rows, d_atU, d_rowsU, M, h->d_nbrU);
        h->launches++;
    }
    if (n_new > 0) { /* new slots are contiguous: rows (zero-padded to ld), ids and upper_base go out as three plain copies */
        vec_new.assign((size_t)n_new * h->ld, 0.0f);
        for (long long j = 0; j < n_new; j++) {
            memcpy(&vec_new[(size_t)j * h->ld], vectors + (size_t)new_rec[j] * h->dim, (size_t)h->dim * 4);
            ids_new[j] = ids[new_rec[j]];
        }
        CK(h, cudaMemcpyAsync(h->d_vec + (size_t)n_old * h->ld, vec_new.data(), vec_new.size() * 4, cudaMemcpyHostToDevice, st));
        CK(h, cudaMemcpyAsync(h->d_ids + n_old, ids_new.data(), (size_t)n_new * 8, cudaMemcpyHostToDevice, st));
        CK(h, cudaMemcpyAsync(h->d_upper_base + n_old, ub_new.data(), (size_t)n_new * 4, cudaMemcpyHostToDevice, st));
        row_norm2_kernel<<<(unsigned)((n_new + 255) / 256), 256, 0, st>>>(h->d_vec + (size_t)n_old * h->ld, h->ld, h->dim, n_new,
                                                                        h->d_norm2 + n_old);
        h->launches++;
    }
    if (!newly_dead.empty()) { /* after the flags above, in stream order */
        purge_slots_kernel<<<h->num_sms * 8, 256, 0, st>>>(h->d_nbr0, (long long)n_total * M0, h->d_deleted);
        const long long nu = h->n_upper + add_upper;
        if (nu) purge_slots_kernel<<<h->num_sms * 2, 256, 0, st>>>(h->d_nbrU, nu * M, h->d_deleted);
        h->launches += nu ? 2 : 1;
    }
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(st); /* the staging vectors above go out of scope */
    if (e != cudaSuccess) { /* the device copy is in an unknown state: only a fresh upload makes the handle usable again */
        h->uploaded = false;
        CK(h, e);
    }
    /* ── phase 3: commit the host mirrors ── */
    h->h_levels.resize(n_total); h->h_upper_base.resize(n_total, -1); h->h_deleted.resize(n_total, 0);
    for (long long j = 0; j < n_new; j++) {
        h->h_levels[n_old + j] = new_levels[j];
        h->h_upper_base[n_old + j] = ub_new[j];
        h->h_deleted[n_old + j] = new_deleted[j];
        if (!new_deleted[j]) h->n_live++;
    }
    for (auto &nd : newly_dead) { h->h_deleted[nd.first] = 1; h->n_live--; }
    h->n_upper += add_upper;
    h->n = n_total; h->entry_slot = entry_slot; h->max_level = max_level;
    if (n_new > 0) h->last_id = prev_id;
    return MUNINN_CUDA_OK;
}

/* ───────────────────────── searcher for the host builder's group loop ───────────────────────── */

static int bd_upload(void *ctx, int64_t n, const int64_t *ids, const int32_t *levels, const float *vectors,
                     const int64_t *list_offsets, const int32_t *targets, int32_t entry_slot, int32_t max_level) {
    return muninn_cuda_index_upload((MuninnCudaIndex *)ctx, n, ids, levels, nullptr, vectors, list_offsets, targets, entry_slot, max_level);
}
static int bd_patch(void *ctx, int64_t n_total, int32_t n_patch, const int32_t *slots, const int64_t *ids, const int32_t *levels,
                    const float *vectors, const int64_t *list_offsets, const int32_t *targets, int32_t entry_slot, int32_t max_level) {
    return muninn_cuda_index_patch((MuninnCudaIndex *)ctx, n_total, n_patch, slots, ids, levels, nullptr, vectors, list_offsets, targets,
                                   entry_slot, max_level);
}
static int bd_search(void *ctx, const float *queries, int nq, int k, int ef, int64_t *out_ids, float *out_dist, int32_t *out_count) {
    return muninn_cuda_search_batch((MuninnCudaIndex *)ctx, queries, nq, k, ef, out_ids, out_dist, out_count, nullptr);
}

extern "C" int muninn_cuda_build_device(MuninnCudaIndex *h, MuninnBuildDevice *out) {
    if (!h || !out) return MUNINN_CUDA_E_ARG;
    out->ctx = h; out->upload = bd_upload; out->patch = bd_patch; out->search = bd_search;
    return MUNINN_CUDA_OK;
}

/* ───────────────────────── batched HNSW search ───────────────────────── */

static unsigned pow2_at_least(unsigned long long v) {
    unsigned long long p = 1024;
    while (p < v) p <<= 1;
    return (unsigned)p;
}
static unsigned log2u(unsigned v) { unsigned l = 0; while ((1u << l) < v) l++; return l; }

/* ── kernel selection ──
 * Two kernels answer the same contract and are tested for identical output in every shape: hnsw_search.cuh (one warp
 * per query, 13-row whole-row tile: 24 query teams per SM, the best throughput for full batches of narrow rows) and
 * hnsw_beam.cuh (query teams of W warps, 16-row batches through a row ring). kernel_gen 0 picks by shape from the
 * measurements in profiles/r2_bench/: teams of 4 or 2 warps — the widest at which every query of the batch is resident at once
 * (a lone MATCH: 4 warps; the 1 250-query slices of a batch split over 8 GPUs: 2 warps, 8-11 % faster than one), the 16-row kernel for
 * ef >= 192 at dim 128 (5 % faster) and for batches of up to two waves that need no more waves of its teams than of the 13-row kernel's (4-15 % faster), the 13-row kernel otherwise. MUNINN_KERNEL / muninn_cuda_set_option("kernel_gen")
 * force one for A/B runs. A kernel is picked as a plain function pointer so that occupancy query, attribute and launch
 * share one code path. Ring geometry (U floats per unit, NST stages) per constant-dimension instance: */
#ifndef MUNINN_U_D128
#define MUNINN_U_D128 128 /* whole row in one unit: the best of the A/B sweeps (profiles/r2_bench/ab_b_geometry_nospec.jsonl) */
#endif
#ifndef MUNINN_NST_D128
#define MUNINN_NST_D128 1
#endif
#ifndef MUNINN_U_D96
#define MUNINN_U_D96 96
#endif
#ifndef MUNINN_NST_D96
#define MUNINN_NST_D96 1
#endif
#ifndef MUNINN_U_D768
#define MUNINN_U_D768 128
#endif
#ifndef MUNINN_NST_D768
#define MUNINN_NST_D768 2
#endif
#ifndef MUNINN_U_GEN
#define MUNINN_U_GEN 64
#endif
#ifndef MUNINN_NST_GEN
#define MUNINN_NST_GEN 2
#endif

struct KernelSel {
    const void *fn = nullptr;
    int gen = 8, W = 1, U = 0, NST = 0;
    size_t smem = 0;
    size_t vis_smem = 0; /* part of smem that holds the team's first-level visited table (SearchParams::vis_smem) */
};

template <int METRIC, int W>
static const void *beam_fn(int dim, int *U, int *NST) {
    if (MUNINN_D128 && dim == 128) { *U = MUNINN_U_D128; *NST = MUNINN_NST_D128; return (const void *)beam::hnsw_beam_kernel<METRIC, 128, MUNINN_U_D128, MUNINN_NST_D128, W>; }
    if (MUNINN_D128 && dim == 96) { *U = MUNINN_U_D96; *NST = MUNINN_NST_D96; return (const void *)beam::hnsw_beam_kernel<METRIC, 96, MUNINN_U_D96, MUNINN_NST_D96, W>; }
    if (MUNINN_D128 && dim == 768) { *U = MUNINN_U_D768; *NST = MUNINN_NST_D768; return (const void *)beam::hnsw_beam_kernel<METRIC, 768, MUNINN_U_D768, MUNINN_NST_D768, W>; }
    *U = MUNINN_U_GEN; *NST = MUNINN_NST_GEN;
    return (const void *)beam::hnsw_beam_kernel<METRIC, 0, MUNINN_U_GEN, MUNINN_NST_GEN, W>;
}
template <int METRIC>
static const void *beam_fn_w(int W, int dim, int *U, int *NST) {
    return W == 4 ? beam_fn<METRIC, 4>(dim, U, NST) : W == 2 ? beam_fn<METRIC, 2>(dim, U, NST) : beam_fn<METRIC, 1>(dim, U, NST);
}
template <int METRIC>
static const void *v7_fn(int dim, int ld) {
    if (MUNINN_D128 && dim == 128) return (const void *)hnsw_search_kernel<METRIC, RB_NARROW, 128>;
    if (MUNINN_D128 && dim == 96) return (const void *)hnsw_search_kernel<METRIC, RB_D96, 96>;
    if (MUNINN_D128 && dim == 768) return (const void *)hnsw_search_kernel<METRIC, RB_WIDE, 768>;
    return beam_rows(ld) == RB_NARROW ? (const void *)hnsw_search_kernel<METRIC, RB_NARROW, 0> : (const void *)hnsw_search_kernel<METRIC, RB_WIDE, 0>;
}

static KernelSel select_kernel(const MuninnCudaIndex *h, int gen, int W, int ef) {
    KernelSel k;
    k.gen = gen;
    if (gen == 7) {
        k.W = 1;
        k.fn = h->metric == MUNINN_METRIC_L2 ? v7_fn<METRIC_L2>(h->dim, h->ld)
               : h->metric == MUNINN_METRIC_COSINE ? v7_fn<METRIC_COSINE>(h->dim, h->ld) : v7_fn<METRIC_IP>(h->dim, h->ld);
        k.smem = search_smem_bytes(ef, h->ld, h->W0, MUNINN_D128 ? h->dim : -1);
        return k;
    }
    k.W = W;
    k.fn = h->metric == MUNINN_METRIC_L2 ? beam_fn_w<METRIC_L2>(W, h->dim, &k.U, &k.NST)
           : h->metric == MUNINN_METRIC_COSINE ? beam_fn_w<METRIC_COSINE>(W, h->dim, &k.U, &k.NST)
                                               : beam_fn_w<METRIC_IP>(W, h->dim, &k.U, &k.NST);
    k.smem = beam::smem_bytes(ef, h->ld, h->W0, k.U, k.NST, W);
    return k;
}

/* The dynamic shared memory a kernel function may be launched with is a property of the FUNCTION on a device, shared by
 * every handle of the process, and one function serves launches of different sizes (other ef, with or without the
 * visited table in shared memory): the limit is only ever raised, under a lock, and remembered per (device, function). */
static cudaError_t allow_dynamic_smem(int device, const void *fn, size_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<int, const void *>, size_t> allowed;
    std::lock_guard<std::mutex> lock(mu);
    size_t &cur = allowed[std::make_pair(device, fn)];
    if (bytes <= cur) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) cur = bytes;
    return e;
}

/* resident teams per SM of a selection, cached per (generation, W, shared-memory size) */
static cudaError_t occupancy_cached(MuninnCudaIndex *h, const KernelSel &k, int *occ) {
    const int fi = k.gen == 7 ? 3 : (k.W == 4 ? 2 : k.W == 2 ? 1 : 0);
    const int wi = k.vis_smem && k.W > 1 ? (k.W == 4 ? 4 : 5) : fi; /* the cache slot */
    cudaError_t e = allow_dynamic_smem(h->device, k.fn, k.smem);
    if (e != cudaSuccess) return e;
    if (h->occ_smem[wi] == k.smem && h->occ_val[wi] > 0) { *occ = h->occ_val[wi]; return cudaSuccess; }
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k.fn, 32 * k.W, k.smem);
    if (e != cudaSuccess) return e;
    h->occ_smem[wi] = k.smem; h->occ_val[wi] = *occ;
    return cudaSuccess;
}

static const size_t SEARCH_SMEM_LIMIT = 200 * 1024;
static const size_t VIS_SMEM_MAX = 64 * 1024; /* largest first-level visited table a team keeps in shared memory (ef = 256 at M = 16) */

/* Team width, kernel and shared-memory layout of one search call: the caller's choice (muninn_cuda_set_tuning, "kernel_gen"),
 * else by shape — see "kernel selection" above. resident_cap bounds how many teams the caller's visited arena can keep
 * resident (no bound when the arena grows with the launch); allow_tbl = the first-level table may live in shared memory. */
static int choose_search_kernel(MuninnCudaIndex *h, int nq, int ef, long long resident_cap, bool allow_tbl, KernelSel *out_sel,
                                int *out_occ) {
    /* A lone query is a dependent chain of instructions: more warps take the distance evaluations of each expansion off
     * it, as long as every query of the batch can be resident at once; a full batch wants as many independent queries
     * per SM as fit. */
    KernelSel sel;
    int occ = 0;
    {
        const long long kept = resident_cap; /* teams the visited arena lets a launch keep resident */
        const int forced_w = h->warps_per_query, forced_gen = h->kernel_gen;
        int W = forced_w ? forced_w : 1;
        /* the teams of several warps can keep their first-level visited table (sized as below) in shared memory */
        const size_t tbl = (size_t)std::min(pow2_at_least((unsigned long long)(h->vis_factor * (double)ef * h->W0)),
                                            pow2_at_least(2ull * (unsigned long long)(h->n + h->W0 + 1))) * 4;
        const bool tbl_ok = allow_tbl && h->vis_in_smem && tbl <= VIS_SMEM_MAX;
        bool use_tbl = false;
        if (!forced_w && forced_gen != 7) {
            /* automatic width: the widest team at which the whole batch is still resident — with the table in shared
             * memory if the batch is still resident at the occupancy that leaves, else without. 4 warps beat 2 by 2-5 % up
             * to ≈ 600 queries and lose badly once a second wave is needed (profiles/r2_bench/ab_team_width4.jsonl) */
            for (int w = 4; w >= 2 && W == 1; w >>= 1) {
                for (int with_tbl = tbl_ok ? 1 : 0; with_tbl >= 0 && W == 1; with_tbl--) {
                    KernelSel sw = select_kernel(h, 8, w, ef);
                    if (with_tbl) { sw.vis_smem = tbl; sw.smem += tbl; }
                    int ow = 0;
                    if (sw.smem > SEARCH_SMEM_LIMIT) continue;
                    CK(h, occupancy_cached(h, sw, &ow));
                    if (ow >= 1 && (long long)nq <= std::min((long long)h->num_sms * ow, kept)) { W = w; use_tbl = with_tbl != 0; }
                }
            }
        } else if (forced_w > 1 && tbl_ok) {
            KernelSel sw = select_kernel(h, 8, forced_w, ef);
            use_tbl = sw.smem + tbl <= SEARCH_SMEM_LIMIT;
        }
        int gen = forced_gen;
        if (gen == 0) gen = W > 1 ? 8 : ((h->dim == 128 && ef >= 192) ? 8 : 7);
        if (forced_gen == 0 && gen == 7) {
            /* Up to two waves of resident teams the time of a launch is a staircase in the NUMBER of waves, and the 13-row
             * kernel's extra residency only pays where it saves one: whenever the ring kernel needs no more waves than the
             * 13-row kernel, its 16-row batches win — by 4-15 % on one wave (profiles/r2_bench/ab_one_wave.jsonl), by 10 % on
             * two (5 000 queries, a 10k batch split over 2 GPUs: 0.842 vs 0.936 ms; 6 400: 1.13 vs 1.01 the other way;
             * profiles/r2_bench/ab_two_waves.jsonl). Longer batches keep the 13-row kernel (measured at 10k and 80k). */
            KernelSel s8 = select_kernel(h, 8, 1, ef), s7 = select_kernel(h, 7, 1, ef);
            int o8 = 0, o7 = 0;
            if (s8.smem <= SEARCH_SMEM_LIMIT && s7.smem <= SEARCH_SMEM_LIMIT) {
                CK(h, occupancy_cached(h, s8, &o8));
                CK(h, occupancy_cached(h, s7, &o7));
                if (o8 >= 1 && o7 >= 1) {
                    const long long cap8 = std::min((long long)h->num_sms * o8, kept), cap7 = std::min((long long)h->num_sms * o7, kept);
                    const long long w8 = (nq + cap8 - 1) / cap8, w7 = (nq + cap7 - 1) / cap7;
                    if (w7 <= 2 && w8 <= w7) gen = 8;
                }
            }
        }
        if (gen == 7 && W > 1) gen = 8; /* the one-warp kernel has no wider shape */
        sel = select_kernel(h, gen, W, ef);
        if (sel.smem > SEARCH_SMEM_LIMIT && W > 1) { W = 1; sel = select_kernel(h, gen, W, ef); }
        if (sel.smem > SEARCH_SMEM_LIMIT)
            return fail(h, MUNINN_CUDA_E_ARG, "ef=%d with dim=%d needs %zu B of shared memory per query (limit 200 KiB)", ef, h->dim, sel.smem);
        if (use_tbl && sel.gen == 8 && sel.W > 1 && sel.smem + tbl <= SEARCH_SMEM_LIMIT) { sel.vis_smem = tbl; sel.smem += tbl; }
        CK(h, occupancy_cached(h, sel, &occ));
    }
    *out_sel = sel; *out_occ = occ;
    return MUNINN_CUDA_OK;
}

static int search_device_impl(MuninnCudaIndex *h, const float *d_queries, int nq, int k, int ef, int64_t *d_ids, float *d_dist,
                              int32_t *d_count, MuninnSearchStats *d_stats, const unsigned *d_arrived);

extern "C" int muninn_cuda_search_batch_device(MuninnCudaIndex *h, const float *d_queries, int nq, int k, int ef,
                                               int64_t *d_ids, float *d_dist, int32_t *d_count,
                                               MuninnSearchStats *d_stats) {
    return search_device_impl(h, d_queries, nq, k, ef, d_ids, d_dist, d_count, d_stats, nullptr);
}

/* d_arrived: NULL, or a device word that the copy engine raises to the number of queries already present in d_queries
 * (muninn_cuda_search_batch streams a large batch in behind the kernel's back; SearchParams::q_arrived) */
static int search_device_impl(MuninnCudaIndex *h, const float *d_queries, int nq, int k, int ef, int64_t *d_ids, float *d_dist,
                              int32_t *d_count, MuninnSearchStats *d_stats, const unsigned *d_arrived) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!h->uploaded) return fail(h, MUNINN_CUDA_E_STATE, "search before muninn_cuda_index_upload");
    if (h->rows_only) return fail(h, MUNINN_CUDA_E_STATE, "this handle holds rows only (muninn_cuda_index_upload_rows): no graph to search");
    if (nq < 0 || k < 1) return fail(h, MUNINN_CUDA_E_ARG, "nq must be >= 0 and k >= 1 (nq=%d k=%d)", nq, k);
    if (nq == 0) return MUNINN_CUDA_OK;
    if (!d_queries || !d_ids || !d_dist || !d_count) return fail(h, MUNINN_CUDA_E_ARG, "NULL buffer");
    if (ef < k) ef = k; /* src/hnsw_algo.c:673 */
    CK(h, cudaSetDevice(h->device));

    KernelSel sel;
    int occ = 0;
    {
        const int rc_sel = choose_search_kernel(h, nq, ef, LLONG_MAX, true, &sel, &occ);
        if (rc_sel) return rc_sel;
    }
    if (occ < 1) return fail(h, MUNINN_CUDA_E_CUDA, "search kernel does not fit on an SM");
    const size_t smem = sel.smem;
    int tps = h->teams_per_sm > 0 ? std::min(h->teams_per_sm, occ) : occ;
    int teams[3];
    unsigned caps[3];
    teams[0] = (int)std::min<long long>((long long)h->num_sms * tps, nq);
    /* (Sizing the grid to ⌈nq / rounds⌉ teams so that every "round" of queries is full — 3 334 instead of 3 552 teams for
     * a 10k batch — was measured and dropped: 1.543 vs 1.492 ms at ef = 64, 46.8 vs 42.6 ms at config 3. Query lengths vary
     * enough that the work queue, not a round structure, decides; every resident slot is worth having.
     * profiles/r2_bench/ab_balance_rounds.jsonl) */
    const unsigned cap_all = pow2_at_least(2ull * (unsigned long long)(h->n + h->W0 + 1));
    /* first-level table: 2·ef·2M slots. A search visits ≈ 0.8·ef·2M nodes on the bench graphs, so the table
     * runs at ≈ 40 % load at the end of a typical query and the whole arena (teams × cap × 4 B = 58 MB at
     * config 2) stays L2-resident; the rare query that would pass 3/4 load is redone at the next level. */
    caps[0] = std::min(pow2_at_least((unsigned long long)(h->vis_factor * (double)ef * h->W0)), cap_all);
    caps[1] = (unsigned)std::min<unsigned long long>(16ull * caps[0], cap_all);
    caps[2] = cap_all;
    int nlevels = caps[0] == cap_all ? 1 : (caps[1] == cap_all ? 2 : 3);
    const size_t arena = std::max<size_t>((size_t)teams[0] * caps[0], cap_all);
    for (int lv = 1; lv < 3; lv++) teams[lv] = (int)std::max<size_t>(1, std::min<size_t>(arena / caps[lv], (size_t)teams[0]));

    if (arena > h->vis_entries) { /* no initialisation needed: every query clears its team's table */
        CK(h, cudaStreamSynchronize(h->stream));
        cudaFree(h->d_vis); h->d_vis = nullptr; h->vis_entries = 0;
        CK(h, cudaMalloc((void **)&h->d_vis, arena * 4));
        h->vis_entries = arena;
    }
    if ((size_t)2 * nq > h->retry_cap) {
        CK(h, cudaStreamSynchronize(h->stream));
        cudaFree(h->d_retry); h->d_retry = nullptr;
        CK(h, cudaMalloc((void **)&h->d_retry, (size_t)2 * nq * sizeof(int)));
        h->retry_cap = (size_t)2 * nq;
    }
    CK(h, cudaMemsetAsync(h->d_ctl, 0, 8 * sizeof(int), h->stream));

    SearchParams p;
    memset(&p, 0, sizeof p);
    p.vec = h->d_vec; p.norm2 = h->d_norm2; p.nbr0 = h->d_nbr0; p.upper_base = h->d_upper_base; p.nbrU = h->d_nbrU;
    p.deleted = h->d_deleted; p.ids = h->d_ids;
    p.dim = h->dim; p.ld = h->ld; p.M = h->WU; p.M0 = h->W0; /* row widths: 2M / M unless a stored list is longer */ p.entry_slot = h->entry_slot; p.max_level = h->max_level;
    p.n_live = h->n_live;
    p.queries = d_queries; p.nq = nq; p.k = k; p.ef = ef;
    p.patience_max = std::max(ef / 4, 10); /* src/hnsw_algo.c:372-374 */
    p.out_ids = d_ids; p.out_dist = d_dist; p.out_count = d_count; p.stats = d_stats;
    p.vis = h->d_vis;
    p.q_arrived = d_arrived;
    p.vis_smem = (int)sel.vis_smem; /* caps[0] · 4 by construction: the kernel uses it at level 0 only */

    CK(h, cudaEventRecord(h->ev0, h->stream));
    for (int lv = 0; lv < nlevels; lv++) {
        p.vis_cap = caps[lv];
        p.vis_shift = 32 - log2u(caps[lv]);
        p.work_counter = h->d_ctl + lv;
        p.in_list = lv == 0 ? nullptr : h->d_retry + (size_t)(lv - 1) * nq;
        p.in_count = lv == 0 ? nullptr : h->d_ctl + 3 + (lv - 1);
        p.out_list = h->d_retry + (size_t)(lv < 2 ? lv : 0) * nq;
        p.out_count_retry = h->d_ctl + 3 + lv;
        int level = lv;
        void *args[2] = {(void *)&p, (void *)&level};
        CK(h, cudaLaunchKernel(sel.fn, dim3((unsigned)teams[lv]), dim3((unsigned)(32 * sel.W)), args, smem, h->stream));
        h->launches++;
    }
    CK(h, cudaEventRecord(h->ev1, h->stream));
    h->have_timing = true;
    h->last_W = sel.W; h->last_gen = sel.gen;
    return MUNINN_CUDA_OK;
}

/* ── search without any allocation (muninn_cuda_search_batch_device_static) ──
 * No counterpart in the reference; added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops). The search kernels' scratch — work counters, retry lists and the per-team
 * visited tables — lives in ONE file-scope device array of fixed size, which every GPU that loads the library holds until
 * the process exits:
 *     [0, 256)                          control words (work counters ×3, retry counts ×2), cleared by every call
 *     [256, 256 + 2·MAX_NQ·4)           retry lists (queries whose visited table overflowed, per level)
 *     [.., + VIS_ENTRIES·4)             visited arena: teams × capacity 32-bit entries; every query clears its own table
 * Every call on a device shares the array: calls that overlap on the GPU (two streams, two handles) race. */
#define MUNINN_STATIC_MAX_NQ 65536
#define MUNINN_STATIC_VIS_ENTRIES (4u << 20) /* 16 MiB: 1 024 teams at ef = 64, M = 16; indexes up to 2 097 119 nodes at M = 16 */
static const size_t STATIC_CTL_BYTES = 256, STATIC_RETRY_BYTES = (size_t)2 * MUNINN_STATIC_MAX_NQ * sizeof(int),
                    STATIC_VIS_BYTES = (size_t)MUNINN_STATIC_VIS_ENTRIES * sizeof(unsigned);
static __device__ __align__(256) unsigned char g_static_scratch[256 + (size_t)2 * MUNINN_STATIC_MAX_NQ * sizeof(int) +
                                                                (size_t)MUNINN_STATIC_VIS_ENTRIES * sizeof(unsigned)];

/* The array's name is only a stand-in on the host: its device address is resolved once per device, with that device current */
static const int STATIC_MAX_DEVICES = 64;
static unsigned char *g_static_ptr[STATIC_MAX_DEVICES];
static std::mutex g_static_mu;

static int static_scratch_address(MuninnCudaIndex *h, unsigned char **out) {
    if (h->device < 0 || h->device >= STATIC_MAX_DEVICES) return fail(h, MUNINN_CUDA_E_ARG, "device %d: the static scratch is tracked for devices 0..%d", h->device, STATIC_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[h->device]) {
        void *ptr = nullptr;
        CK(h, cudaSetDevice(h->device));
        CK(h, cudaGetSymbolAddress(&ptr, g_static_scratch));
        if (!ptr) return fail(h, MUNINN_CUDA_E_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_static_ptr[h->device] = (unsigned char *)ptr;
    }
    *out = g_static_ptr[h->device];
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_search_batch_device_static(MuninnCudaIndex *h, const float *d_queries, int nq, int k, int ef,
                                                      int64_t *d_ids, float *d_dist, int32_t *d_count,
                                                      MuninnSearchStats *d_stats) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!h->uploaded) return fail(h, MUNINN_CUDA_E_STATE, "search before muninn_cuda_index_upload");
    if (h->rows_only) return fail(h, MUNINN_CUDA_E_STATE, "this handle holds rows only (muninn_cuda_index_upload_rows): no graph to search");
    if (nq < 0 || k < 1) return fail(h, MUNINN_CUDA_E_ARG, "nq must be >= 0 and k >= 1 (nq=%d k=%d)", nq, k);
    if (nq == 0) return MUNINN_CUDA_OK;
    if (!d_queries || !d_ids || !d_dist || !d_count) return fail(h, MUNINN_CUDA_E_ARG, "NULL buffer");
    if (nq > MUNINN_STATIC_MAX_NQ)
        return fail(h, MUNINN_CUDA_E_ARG, "nq=%d exceeds the static scratch's limit of %d queries per call (use muninn_cuda_search_batch_device)", nq, MUNINN_STATIC_MAX_NQ);
    if (ef < k) ef = k; /* src/hnsw_algo.c:673 */
    CK(h, cudaSetDevice(h->device));

    /* visited-table capacities exactly as in search_device_impl; the arena is the fixed array instead of a grown allocation */
    const unsigned cap_all = pow2_at_least(2ull * (unsigned long long)(h->n + h->W0 + 1));
    if (cap_all > MUNINN_STATIC_VIS_ENTRIES)
        return fail(h, MUNINN_CUDA_E_ARG, "an index of %lld nodes needs a last-resort visited table of %u entries; the static scratch holds %u "
                    "(indexes up to about %u nodes; use muninn_cuda_search_batch_device)", (long long)h->n, cap_all, MUNINN_STATIC_VIS_ENTRIES, MUNINN_STATIC_VIS_ENTRIES / 2 - 256);
    unsigned caps[3];
    caps[0] = std::min(pow2_at_least((unsigned long long)(h->vis_factor * (double)ef * h->W0)), cap_all);
    caps[1] = (unsigned)std::min<unsigned long long>(16ull * caps[0], cap_all);
    caps[2] = cap_all;
    const int nlevels = caps[0] == cap_all ? 1 : (caps[1] == cap_all ? 2 : 3);
    const long long arena_teams = MUNINN_STATIC_VIS_ENTRIES / caps[0];

    /* team width and kernel by the same policy as search_device_impl, with "resident" bounded by the arena as well and
     * every visited table in the static array */
    KernelSel sel;
    int occ = 0;
    {
        const int rc_sel = choose_search_kernel(h, nq, ef, arena_teams, false, &sel, &occ);
        if (rc_sel) return rc_sel;
    }
    if (occ < 1) return fail(h, MUNINN_CUDA_E_CUDA, "search kernel does not fit on an SM");
    const int tps = h->teams_per_sm > 0 ? std::min(h->teams_per_sm, occ) : occ;
    int teams[3];
    teams[0] = (int)std::min<long long>(std::min<long long>((long long)h->num_sms * tps, nq), arena_teams);
    for (int lv = 1; lv < 3; lv++)
        teams[lv] = (int)std::max<long long>(1, std::min<long long>(MUNINN_STATIC_VIS_ENTRIES / caps[lv], teams[0]));

    unsigned char *scratch = nullptr;
    int rc = static_scratch_address(h, &scratch);
    if (rc) return rc;
    int *ctl = reinterpret_cast<int *>(scratch);
    int *retry = reinterpret_cast<int *>(scratch + STATIC_CTL_BYTES);
    unsigned *vis = reinterpret_cast<unsigned *>(scratch + STATIC_CTL_BYTES + STATIC_RETRY_BYTES);
    /* the array keeps the previous call's contents: the counters must start at zero (tables and lists are written before read) */
    CK(h, cudaMemsetAsync(ctl, 0, 8 * sizeof(int), h->stream));

    SearchParams p;
    memset(&p, 0, sizeof p);
    p.vec = h->d_vec; p.norm2 = h->d_norm2; p.nbr0 = h->d_nbr0; p.upper_base = h->d_upper_base; p.nbrU = h->d_nbrU;
    p.deleted = h->d_deleted; p.ids = h->d_ids;
    p.dim = h->dim; p.ld = h->ld; p.M = h->WU; p.M0 = h->W0; p.entry_slot = h->entry_slot; p.max_level = h->max_level;
    p.n_live = h->n_live;
    p.queries = d_queries; p.nq = nq; p.k = k; p.ef = ef;
    p.patience_max = std::max(ef / 4, 10); /* src/hnsw_algo.c:372-374 */
    p.out_ids = d_ids; p.out_dist = d_dist; p.out_count = d_count; p.stats = d_stats;
    p.vis = vis;
    p.q_arrived = nullptr;
    for (int lv = 0; lv < nlevels; lv++) {
        p.vis_cap = caps[lv];
        p.vis_shift = 32 - log2u(caps[lv]);
        p.work_counter = ctl + lv;
        p.in_list = lv == 0 ? nullptr : retry + (size_t)(lv - 1) * nq;
        p.in_count = lv == 0 ? nullptr : ctl + 3 + (lv - 1);
        p.out_list = retry + (size_t)(lv < 2 ? lv : 0) * nq;
        p.out_count_retry = ctl + 3 + lv;
        int level = lv;
        void *args[2] = {(void *)&p, (void *)&level};
        CK(h, cudaLaunchKernel(sel.fn, dim3((unsigned)teams[lv]), dim3((unsigned)(32 * sel.W)), args, sel.smem, h->stream));
        h->launches++;
    }
    h->last_W = sel.W; h->last_gen = sel.gen;
    return MUNINN_CUDA_OK;
}

/* limits of muninn_cuda_search_batch_device_static, for callers that size their batches against them */
extern "C" int muninn_cuda_static_scratch_limits(int *max_nq, int64_t *max_nodes, int64_t *bytes) {
    if (max_nq) *max_nq = MUNINN_STATIC_MAX_NQ;
    if (max_nodes) *max_nodes = (int64_t)MUNINN_STATIC_VIS_ENTRIES / 2 - 256;
    if (bytes) *bytes = (int64_t)(STATIC_CTL_BYTES + STATIC_RETRY_BYTES + STATIC_VIS_BYTES);
    return MUNINN_CUDA_OK;
}

static int ensure_stage(MuninnCudaIndex *h, size_t bytes) {
    if (bytes <= h->stage_bytes) return MUNINN_CUDA_OK;
    CK(h, cudaStreamSynchronize(h->stream));
    cudaFree(h->d_stage); h->d_stage = nullptr; h->stage_bytes = 0;
    CK(h, cudaMalloc(&h->d_stage, bytes));
    h->stage_bytes = bytes;
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_search_batch(MuninnCudaIndex *h, const float *queries, int nq, int k, int ef,
                                        int64_t *out_ids, float *out_dist, int32_t *out_count,
                                        MuninnSearchStats *out_stats) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!h->uploaded) return fail(h, MUNINN_CUDA_E_STATE, "search before muninn_cuda_index_upload");
    if (nq < 0 || k < 1) return fail(h, MUNINN_CUDA_E_ARG, "nq must be >= 0 and k >= 1 (nq=%d k=%d)", nq, k);
    if (nq == 0) return MUNINN_CUDA_OK;
    if (!queries || !out_ids || !out_dist || !out_count) return fail(h, MUNINN_CUDA_E_ARG, "NULL buffer");
    CK(h, cudaSetDevice(h->device));
    const size_t bq = align256((size_t)nq * h->dim * 4), bi = align256((size_t)nq * k * 8),
                 bd = align256((size_t)nq * k * 4), bc = align256((size_t)nq * 4),
                 bs = out_stats ? align256((size_t)nq * sizeof(MuninnSearchStats)) : 0;
    int rc = ensure_stage(h, bq + bi + bd + bc + bs);
    if (rc) return rc;
    char *b = (char *)h->d_stage;
    float *dq = (float *)b; int64_t *di = (int64_t *)(b + bq); float *dd = (float *)(b + bq + bi);
    int32_t *dc = (int32_t *)(b + bq + bi + bd);
    MuninnSearchStats *ds = out_stats ? (MuninnSearchStats *)(b + bq + bi + bd + bc) : nullptr;
    /* small batches (the vtab's one query per statement): the call is latency, not bandwidth — stage through one
     * pinned buffer so that the results come back with ONE device→host copy instead of three pageable ones */
    const size_t out_bytes = bi + bd + bc, in_bytes = (size_t)nq * h->dim * 4;
    const bool tiny = !out_stats && in_bytes + out_bytes <= 64 * 1024;
    if (tiny && h->pinned_bytes < 128 * 1024) {
        if (h->h_pinned) { cudaFreeHost(h->h_pinned); h->h_pinned = nullptr; h->pinned_bytes = 0; }
        if (cudaMallocHost(&h->h_pinned, 128 * 1024) == cudaSuccess) h->pinned_bytes = 128 * 1024;
        else cudaGetLastError();
    }
    if (tiny && h->pinned_bytes >= 128 * 1024) {
        char *pin = (char *)h->h_pinned;
        memcpy(pin, queries, in_bytes);
        if (h->tiny_zero_copy) {
            /* no copy operations at all: pinned memory is device-visible at the same address (unified addressing), so the
             * kernel reads the query and writes its k results straight over PCIe — two copy-engine operations and their
             * cross-engine hand-overs less on a call that is pure latency (profiles/r2_bench/config1_zero_copy.json) */
            char *res = pin + 64 * 1024;
            rc = muninn_cuda_search_batch_device(h, (const float *)pin, nq, k, ef, (int64_t *)res, (float *)(res + bi), (int32_t *)(res + bi + bd), nullptr);
            if (rc) return rc;
            CK(h, cudaStreamSynchronize(h->stream));
            memcpy(out_ids, res, (size_t)nq * k * 8);
            memcpy(out_dist, res + bi, (size_t)nq * k * 4);
            memcpy(out_count, res + bi + bd, (size_t)nq * 4);
            return MUNINN_CUDA_OK;
        }
        CK(h, cudaMemcpyAsync(dq, pin, in_bytes, cudaMemcpyHostToDevice, h->stream));
        rc = muninn_cuda_search_batch_device(h, dq, nq, k, ef, di, dd, dc, nullptr);
        if (rc) return rc;
        char *res = pin + 64 * 1024;
        CK(h, cudaMemcpyAsync(res, b + bq, out_bytes, cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
        memcpy(out_ids, res, (size_t)nq * k * 8);
        memcpy(out_dist, res + bi, (size_t)nq * k * 4);
        memcpy(out_count, res + bi + bd, (size_t)nq * 4);
        return MUNINN_CUDA_OK;
    }
    /* Large batches: ONE launch that starts on the first 8 192 queries while the rest of the batch is still crossing
     * PCIe. The copy engine (own stream) raises a device word behind each slice of the copy; a team that draws a query
     * beyond it waits (SearchParams::q_arrived) — in practice never: the resident teams need several milliseconds to
     * reach query 8 192, the copy of 80 000 × 512 B takes 0.8. (Cutting the batch into several LAUNCHES instead was
     * measured and dropped: every launch pays ≈ 0.25 ms of drain at its end, which cancelled what the copies gained.) */
    if (nq >= 16384) {
        if (!h->copy_in) {
            CK(h, cudaStreamCreateWithFlags(&h->copy_in, cudaStreamNonBlocking));
            CK(h, cudaEventCreateWithFlags(&h->ev_first, cudaEventDisableTiming));
            CK(h, cudaMalloc((void **)&h->d_arrived, 64));
            CK(h, cudaMallocHost((void **)&h->h_arrived, 64));
        }
        const int first = 8192;
        h->h_arrived[0] = (unsigned)first; h->h_arrived[1] = (unsigned)nq;
        const size_t row = (size_t)h->dim * 4;
        CK(h, cudaMemcpyAsync(dq, queries, (size_t)first * row, cudaMemcpyHostToDevice, h->copy_in));
        CK(h, cudaMemcpyAsync(h->d_arrived, &h->h_arrived[0], 4, cudaMemcpyHostToDevice, h->copy_in));
        CK(h, cudaEventRecord(h->ev_first, h->copy_in));
        CK(h, cudaMemcpyAsync(dq + (size_t)first * h->dim, queries + (size_t)first * h->dim, (size_t)(nq - first) * row, cudaMemcpyHostToDevice, h->copy_in));
        CK(h, cudaMemcpyAsync(h->d_arrived, &h->h_arrived[1], 4, cudaMemcpyHostToDevice, h->copy_in));
        CK(h, cudaStreamWaitEvent(h->stream, h->ev_first, 0));
        rc = search_device_impl(h, dq, nq, k, ef, di, dd, dc, ds, h->d_arrived);
        if (rc) { cudaStreamSynchronize(h->copy_in); return rc; }
        CK(h, cudaMemcpyAsync(out_ids, di, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaMemcpyAsync(out_dist, dd, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaMemcpyAsync(out_count, dc, (size_t)nq * 4, cudaMemcpyDeviceToHost, h->stream));
        if (out_stats) CK(h, cudaMemcpyAsync(out_stats, ds, (size_t)nq * sizeof(MuninnSearchStats), cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
        CK(h, cudaStreamSynchronize(h->copy_in));
        return MUNINN_CUDA_OK;
    }
    CK(h, cudaMemcpyAsync(dq, queries, (size_t)nq * h->dim * 4, cudaMemcpyHostToDevice, h->stream));
    rc = muninn_cuda_search_batch_device(h, dq, nq, k, ef, di, dd, dc, ds);
    if (rc) return rc;
    CK(h, cudaMemcpyAsync(out_ids, di, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaMemcpyAsync(out_dist, dd, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaMemcpyAsync(out_count, dc, (size_t)nq * 4, cudaMemcpyDeviceToHost, h->stream));
    if (out_stats) CK(h, cudaMemcpyAsync(out_stats, ds, (size_t)nq * sizeof(MuninnSearchStats), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return MUNINN_CUDA_OK;
}

/* ───────────────────────── exact k-NN ───────────────────────── */

template <int METRIC>
static void launch_dist_tile(MuninnCudaIndex *h, long long row0, int rows, const float *dq, int nq, int panel_ld) {
    dim3 grid((rows + XT - 1) / XT, (nq + QT - 1) / QT);
    exact_dist_tile<METRIC><<<grid, 256, 0, h->stream>>>(h->d_vec, h->d_norm2, h->d_deleted, h->ld, h->dim, row0, rows, dq,
                                                         nq, h->d_panel, panel_ld);
}

static int ensure_topk(MuninnCudaIndex *h, int nq, int k) {
    if ((size_t)nq * k > h->topk_cap || (size_t)nq > h->topk_q) {
        CK(h, cudaStreamSynchronize(h->stream));
        cudaFree(h->d_topk); cudaFree(h->d_topk_size); h->d_topk = nullptr; h->d_topk_size = nullptr;
        h->topk_cap = h->topk_q = 0;
        CK(h, cudaMalloc((void **)&h->d_topk, (size_t)nq * k * 8));
        CK(h, cudaMalloc((void **)&h->d_topk_size, (size_t)nq * 4));
        h->topk_cap = (size_t)nq * k; h->topk_q = (size_t)nq;
    }
    return MUNINN_CUDA_OK;
}

/* CUDA-core path: reference-order chains for every (query, row) pair. List q is written to
 * output row d_qmap[q] when d_qmap is given (the tensor path's overflow fallback). */
static int exact_knn_cudacore(MuninnCudaIndex *h, const float *d_queries, int nq, int k, int64_t index_base,
                              const int *d_qmap, int64_t *d_idx, float *d_dist, int32_t *d_count) {
    /* distance panel: at most 1 GiB, whole 64-row tiles */
    long long chunk = (long long)((1ull << 28) / (unsigned long long)nq);
    chunk = std::max<long long>(XT, chunk / XT * XT);
    chunk = std::min<long long>(chunk, std::max<long long>(XT, (h->n + XT - 1) / XT * XT));
    const size_t need = (size_t)nq * (size_t)chunk;
    if (need > h->panel_floats) {
        CK(h, cudaStreamSynchronize(h->stream));
        cudaFree(h->d_panel); h->d_panel = nullptr; h->panel_floats = 0;
        CK(h, cudaMalloc((void **)&h->d_panel, need * 4));
        h->panel_floats = need;
    }
    int rc = ensure_topk(h, nq, k);
    if (rc) return rc;
    CK(h, cudaMemsetAsync(h->d_topk_size, 0, (size_t)nq * 4, h->stream));
    const size_t sel_smem = (size_t)4 * (k + 32) * 8;
    if (sel_smem > 48 * 1024)
        CK(h, cudaFuncSetAttribute(exact_select, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sel_smem));
    for (long long row0 = 0; row0 < h->n; row0 += chunk) {
        const int rows = (int)std::min<long long>(chunk, h->n - row0);
        if (h->metric == MUNINN_METRIC_L2) launch_dist_tile<METRIC_L2>(h, row0, rows, d_queries, nq, (int)chunk);
        else if (h->metric == MUNINN_METRIC_COSINE) launch_dist_tile<METRIC_COSINE>(h, row0, rows, d_queries, nq, (int)chunk);
        else launch_dist_tile<METRIC_IP>(h, row0, rows, d_queries, nq, (int)chunk);
        exact_select<<<(nq + 3) / 4, 128, sel_smem, h->stream>>>(h->d_panel, (int)chunk, rows, (unsigned)row0, nq, k,
                                                                 h->d_topk, h->d_topk_size);
        h->launches += 2;
        CK(h, cudaGetLastError());
    }
    exact_emit<<<(unsigned)(((size_t)nq * k + 255) / 256), 256, 0, h->stream>>>(h->d_topk, h->d_topk_size, nq, k, index_base,
                                                                               d_qmap, d_idx, d_dist, d_count);
    h->launches++;
    CK(h, cudaGetLastError());
    return MUNINN_CUDA_OK;
}

/* ── tensor-core filter path (exact_tc.cuh) ── */

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn tensor_map_encoder() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)sym;
    }
    return fn;
}

/* [rows × kp] fp32, row-major; box = 32 floats (one 128-byte swizzle row) × box_rows */
static int make_slab_map(MuninnCudaIndex *h, CUtensorMap *map, const float *ptr, int kp, long long rows, int box_rows) {
    /* kp = floats per row = row stride; columns past kp (the last k-block of a raw slab whose width is
     * not a multiple of 32) are out of bounds for TMA and arrive as zeros */
    EncodeTiledFn enc = tensor_map_encoder();
    if (!enc) return fail(h, MUNINN_CUDA_E_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)kp, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)kp * 4};
    cuuint32_t box[2] = {(cuuint32_t)tc::BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(h, MUNINN_CUDA_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return MUNINN_CUDA_OK;
}

template <typename T>
static cudaError_t grow(T **ptr, size_t *have, size_t need, cudaStream_t st) {
    if (need <= *have) return cudaSuccess;
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    cudaFree(*ptr); *ptr = nullptr; *have = 0;
    e = cudaMalloc((void **)ptr, need * sizeof(T));
    if (e == cudaSuccess) *have = need;
    return e;
}

/* error constant of the filter (exact_tc.cuh header): c such that |d̃ − d_ref| ≤ c·(‖q‖² + ‖x‖²) */
static float tc_eps(int mode, int kp) {
    const double split = mode == MUNINN_EXACT_TC_3XTF32 ? 3.0 * ldexp(1.0, -20) : ldexp(1.0, -9);
    const double terms = mode == MUNINN_EXACT_TC_3XTF32 ? 3.0 * kp : 1.0 * kp;
    return (float)(2.0 * (split + terms * ldexp(1.0, -23)) + 4.0 * (kp / 4.0 + 8.0) * ldexp(1.0, -24));
}

template <int METRIC, int NPASS, int BN, bool DBG = false>
static cudaError_t launch_filter(MuninnCudaIndex *h, const CUtensorMap maps[4], const tc::FilterParams &fp) {
    using Cfg = tc::FilterCfg<NPASS, BN>;
    auto kern = tc::exact_tc_filter<METRIC, NPASS, BN, DBG>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    const long long tiles = (long long)((fp.nq + tc::BM - 1) / tc::BM) * ((fp.rows + BN - 1) / BN);
    const int grid = (int)std::min<long long>(tiles, h->num_sms);
    kern<<<grid, tc::THREADS, Cfg::SMEM_BYTES, h->stream>>>(maps[0], maps[1], maps[2], maps[3], fp);
    return cudaGetLastError();
}

/* query-stationary form: one TF32 product, query tile resident in shared memory, row tiles in L2-sized super-blocks.
 * pair = two CTAs of a cluster share every row tile (tcgen05 cta_group::2, M = 256): map_x then has 128-row boxes. */
template <int METRIC>
static cudaError_t launch_filter_qs(MuninnCudaIndex *h, const CUtensorMap &map_q, const CUtensorMap &map_x, const tc::FilterParams &fp,
                                    int row_stride_floats, bool pair) {
    constexpr int BN = 256;
    const int CG = pair ? 2 : 1;
    const void *kern = pair ? (const void *)tc::exact_tc_filter_qs<METRIC, BN, 2> : (const void *)tc::exact_tc_filter_qs<METRIC, BN, 1>;
    const size_t smem = tc::qs_smem_bytes(fp.kblocks, BN, CG);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const long long tiles = (long long)((fp.nq + tc::BM * CG - 1) / (tc::BM * CG)) * ((fp.rows + BN - 1) / BN);
    const int groups = (int)std::min<long long>(tiles, h->num_sms / CG);
    int SB = (int)std::max<long long>(1, ((long long)h->tc_superblock_mb << 20) / ((long long)BN * row_stride_floats * 4));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3((unsigned)(groups * CG));
    cfg.blockDim = dim3(tc::QS_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = h->stream;
    cudaLaunchAttribute attr;
    memset(&attr, 0, sizeof attr);
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = (unsigned)CG; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    void *args[4] = {(void *)&map_q, (void *)&map_x, (void *)&fp, (void *)&SB};
    return cudaLaunchKernelExC(&cfg, kern, args);
}

template <int METRIC>
static cudaError_t launch_filter_m(MuninnCudaIndex *h, const CUtensorMap maps[4], const tc::FilterParams &fp) {
    const bool three = h->exact_mode_run == MUNINN_EXACT_TC_3XTF32;
    if (!three && !fp.dbg_panel && h->tc_tile_rows == 256 && h->tc_qs && fp.nq >= h->tc_qs_min_nq && fp.kblocks <= tc::QS_MAX_KBLOCKS)
        return launch_filter_qs<METRIC>(h, maps[0], h->tc_pair ? h->tc_map_xh : maps[2], fp, h->tc_x_stride, h->tc_pair != 0);
    if (fp.dbg_panel) return three ? launch_filter<METRIC, 3, 256, true>(h, maps, fp) : launch_filter<METRIC, 1, 256, true>(h, maps, fp);
    if (h->tc_tile_rows == 128)
        return three ? launch_filter<METRIC, 3, 128>(h, maps, fp) : launch_filter<METRIC, 1, 128>(h, maps, fp);
    return three ? launch_filter<METRIC, 3, 256>(h, maps, fp) : launch_filter<METRIC, 1, 256>(h, maps, fp);
}

/* Runs split → filter → re-rank over doubling chunks of rows [0, n_rows). With d_dbg set only the
 * first chunk (all n_rows rows, τ = +inf) is filtered and its d̃ panel kept (error probe). */
static int exact_knn_tc(MuninnCudaIndex *h, const float *d_queries, int nq, int k, int64_t index_base, int64_t *d_idx,
                        float *d_dist, int32_t *d_count, float *d_dbg, long long dbg_rows) {
    const int kp = (h->dim + tc::BK - 1) / tc::BK * tc::BK, bn = h->tc_tile_rows;
    /* AUTO: one TF32 product straight off the fp32 slab for every batch size (no split pass: the base is read once).
     * Small batches are HBM-bound and stream the rows once; large ones run the query-stationary filter, which at
     * one product per pair needs a third of the tensor work of 3xTF32 and returns the same survivors (measured
     * 1 133 vs 1 124 per query at 1M×128, k=10): the wider margin of the single product costs next to nothing. */
    h->exact_mode_run = h->exact_mode == MUNINN_EXACT_TC_AUTO ? (nq >= h->tc_auto_1x_min_nq || d_dbg ? MUNINN_EXACT_TC_3XTF32 : MUNINN_EXACT_TC_1XTF32)
                                                              : h->exact_mode;
    const bool three = h->exact_mode_run == MUNINN_EXACT_TC_3XTF32;
    const bool raw_x = !three; /* 1xTF32: the MMA reads the fp32 rows as they are (it drops the low 13 bits) */
    /* small batches are HBM- and launch-bound: larger survivor lists (re-ranked by a CTA per query) allow ×17..32
     * chunk growth, i.e. 4-5 chunks instead of 9 over millions of rows */
    const bool small = !d_dbg && nq < 1024 && k <= 512 && h->ld <= 4096; /* the CTA re-rank's shared memory: 64 KB of keys + 132 KB of row tiles + the query */
    const int cap = d_dbg ? 256 : small ? 4096 : std::max(1024, (4 * k + 255) / 256 * 256);
    const int grow_c = small ? std::min(31, std::max(3, 2 * cap / (5 * k))) : (h->tc_growth > 0 ? h->tc_growth : 2); /* expected survivors per chunk ≈ grow_c·k; ×3 per round beat ×4 and ×2 */
    const long long n = d_dbg ? dbg_rows : h->n;
    /* row scratch: ≤ 1 GiB of split slabs per chunk, whole tiles (none when the slab is read raw) */
    long long chunk_max = raw_x ? (1ll << 40) : (long long)((1ull << 30) / ((size_t)kp * 4 * (three ? 2 : 1)));
    chunk_max = std::max<long long>(cap, chunk_max / 256 * 256);
    if (d_dbg) chunk_max = std::max<long long>(chunk_max, n);
    const long long xrows = raw_x ? 1 : std::min<long long>(chunk_max, std::max<long long>(n, 1));
    cudaStream_t st = h->stream;
    CK(h, grow(&h->d_qsplit, &h->qsplit_floats, (size_t)2 * nq * kp, st));
    CK(h, grow(&h->d_xsplit, &h->xsplit_floats, (size_t)2 * xrows * kp, st));
    if ((size_t)nq > h->tcq_cap) {
        CK(h, cudaStreamSynchronize(st));
        cudaFree(h->d_qaux); cudaFree(h->d_tau); cudaFree(h->d_cand_count); cudaFree(h->d_overflow); cudaFree(h->d_qmap);
        h->d_qaux = h->d_tau = nullptr; h->d_cand_count = h->d_overflow = h->d_qmap = nullptr; h->tcq_cap = 0;
        CK(h, cudaMalloc((void **)&h->d_qaux, (size_t)nq * 4)); CK(h, cudaMalloc((void **)&h->d_tau, (size_t)nq * 4));
        CK(h, cudaMalloc((void **)&h->d_cand_count, (size_t)nq * 4)); CK(h, cudaMalloc((void **)&h->d_overflow, (size_t)nq * 4));
        CK(h, cudaMalloc((void **)&h->d_qmap, (size_t)nq * 4));
        h->tcq_cap = (size_t)nq;
    }
    CK(h, grow(&h->d_cand, &h->cand_entries, (size_t)nq * cap, st));
    if (!h->d_tc_ctr) CK(h, cudaMalloc((void **)&h->d_tc_ctr, 16));
    int rc = ensure_topk(h, nq, k);
    if (rc) return rc;
    CK(h, cudaMemsetAsync(h->d_tc_ctr, 0, 16, st));

    float *q_hi = h->d_qsplit, *q_lo = h->d_qsplit + (size_t)nq * kp;
    float *x_hi = h->d_xsplit, *x_lo = h->d_xsplit + (size_t)xrows * kp;
    const int pq_grid = (nq + 127) / 128;
    if (h->metric == MUNINN_METRIC_L2) tc::tc_prep_queries<METRIC_L2><<<pq_grid, 128, 0, st>>>(d_queries, h->dim, nq, h->d_qaux, h->d_tau, h->d_cand_count, h->d_topk_size, h->d_overflow);
    else if (h->metric == MUNINN_METRIC_COSINE) tc::tc_prep_queries<METRIC_COSINE><<<pq_grid, 128, 0, st>>>(d_queries, h->dim, nq, h->d_qaux, h->d_tau, h->d_cand_count, h->d_topk_size, h->d_overflow);
    else tc::tc_prep_queries<METRIC_IP><<<pq_grid, 128, 0, st>>>(d_queries, h->dim, nq, h->d_qaux, h->d_tau, h->d_cand_count, h->d_topk_size, h->d_overflow);
    tc::tc_split_rows<<<h->num_sms * 8, 256, 0, st>>>(d_queries, h->dim, h->dim, nq, kp, q_hi, three ? q_lo : nullptr);
    h->launches += 2;
    CK(h, cudaGetLastError());

    CUtensorMap maps[4];
    rc = make_slab_map(h, &maps[0], q_hi, kp, nq, tc::BM);
    if (!rc) rc = make_slab_map(h, &maps[1], three ? q_lo : q_hi, kp, nq, tc::BM);
    if (rc) return rc;

    size_t rr_smem = (size_t)4 * (k + 32) * 8;
    int rr_P = 0;
    if (small) {
        size_t P = 64;
        while (P < (size_t)k + cap) P <<= 1;
        rr_smem = P * 8 + ((size_t)h->ld * 4 + 15) / 16 * 16 + (size_t)tc::RR_WARPS * 16 * PSTRIDE * 4 + tc::RR_WARPS * 16 * 4 + tc::RR_WARPS * 8;
        /* the compacted keys re-use the tile region: P entries must fit (8192 × 8 B ≤ 16 × 16 × 528 B) */
        static_assert((size_t)8192 * 8 <= (size_t)tc::RR_WARPS * 16 * PSTRIDE * 4, "re-rank tiles too small for the key buffer");
        rr_P = (int)P;
        CK(h, cudaFuncSetAttribute(tc::exact_tc_rerank_cta<METRIC_L2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
        CK(h, cudaFuncSetAttribute(tc::exact_tc_rerank_cta<METRIC_COSINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
        CK(h, cudaFuncSetAttribute(tc::exact_tc_rerank_cta<METRIC_IP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
    } else if (rr_smem > 48 * 1024) {
        CK(h, cudaFuncSetAttribute(tc::exact_tc_rerank<METRIC_L2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
        CK(h, cudaFuncSetAttribute(tc::exact_tc_rerank<METRIC_COSINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
        CK(h, cudaFuncSetAttribute(tc::exact_tc_rerank<METRIC_IP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
    }
    tc::FilterParams fp;
    memset(&fp, 0, sizeof fp);
    fp.xnorm2 = h->d_norm2; fp.qaux = h->d_qaux; fp.tau = h->d_tau; fp.cand = h->d_cand; fp.cand_count = h->d_cand_count;
    fp.cand_cap = cap; fp.nq = nq; fp.kblocks = kp / tc::BK; fp.c_eps = tc_eps(h->exact_mode_run, kp);
    fp.dbg_panel = d_dbg; fp.dbg_ld = (int)n;

    int chunks = 0;
    size_t ev_used = 0;
    const bool seed_dense = h->tc_seed_dense && !small && !d_dbg;
    long long seeded_rows = 0; /* rows of the dense first chunk: evaluated exactly for every query, like survivors */
    for (long long done = 0; done < n;) {
        /* first chunk: every row survives (τ = +inf), so keep it small but ≥ 2k rows; then double */
        long long chunk = done == 0 ? std::min<long long>(small ? cap / 2 : std::min(cap, std::max(256, (2 * k + 255) / 256 * 256)), n)
                                    : std::min<long long>(std::min((long long)grow_c * done, chunk_max), n - done); /* ×(1+grow_c) per round */
        if (d_dbg) chunk = n;
        if (done == 0 && seed_dense) {
            /* Every row of the first chunk survives (τ = +inf), so filtering it is wasted work and re-ranking it row by row
             * is a gather: evaluate it densely instead — 64 × 64 (query, row) tiles of reference-order chains, then one
             * selection per query — and take τ from the result. 1M×128×10k: 0.46 ms (filter 0.05 + re-rank 0.42) → see
             * DESIGN §4.2. Same keys, same order: the list after rows [0, chunk) does not depend on how it was computed. */
            const int panel_ld = (int)((chunk + XT - 1) / XT * XT);
            const size_t need = (size_t)nq * (size_t)panel_ld;
            if (need > h->panel_floats) {
                CK(h, cudaStreamSynchronize(st));
                cudaFree(h->d_panel); h->d_panel = nullptr; h->panel_floats = 0;
                CK(h, cudaMalloc((void **)&h->d_panel, need * 4));
                h->panel_floats = need;
            }
            if (rr_smem > 48 * 1024) CK(h, cudaFuncSetAttribute(exact_select, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
            if (h->metric == MUNINN_METRIC_L2) launch_dist_tile<METRIC_L2>(h, 0, (int)chunk, d_queries, nq, panel_ld);
            else if (h->metric == MUNINN_METRIC_COSINE) launch_dist_tile<METRIC_COSINE>(h, 0, (int)chunk, d_queries, nq, panel_ld);
            else launch_dist_tile<METRIC_IP>(h, 0, (int)chunk, d_queries, nq, panel_ld);
            exact_select<<<(nq + 3) / 4, 128, rr_smem, st>>>(h->d_panel, panel_ld, (int)chunk, 0u, nq, k, h->d_topk, h->d_topk_size);
            tc::tc_seed_tau<<<(nq + 255) / 256, 256, 0, st>>>(h->d_topk, h->d_topk_size, nq, k, h->d_tau);
            h->launches += 3;
            CK(h, cudaGetLastError());
            seeded_rows = chunk;
            done += chunk;
            chunks++;
            continue;
        }
        h->tc_x_stride = raw_x ? h->ld : kp;
        if (raw_x) {
            rc = make_slab_map(h, &maps[2], h->d_vec + (size_t)done * h->ld, h->ld, chunk, bn);
            maps[3] = maps[2];
            if (!rc && h->tc_pair) rc = make_slab_map(h, &h->tc_map_xh, h->d_vec + (size_t)done * h->ld, h->ld, chunk, 128);
        } else {
            tc::tc_split_rows<<<h->num_sms * 8, 256, 0, st>>>(h->d_vec + (size_t)done * h->ld, h->ld, h->dim, chunk, kp, x_hi,
                                                              three ? x_lo : nullptr);
            h->launches++;
            rc = make_slab_map(h, &maps[2], x_hi, kp, chunk, bn);
            if (!rc) rc = make_slab_map(h, &maps[3], three ? x_lo : x_hi, kp, chunk, bn);
        }
        if (rc) return rc;
        fp.row0 = done; fp.rows = (int)chunk;
        while (h->tc_ev.size() < ev_used + 2) {
            cudaEvent_t e;
            CK(h, cudaEventCreate(&e));
            h->tc_ev.push_back(e);
        }
        CK(h, cudaEventRecord(h->tc_ev[ev_used], st));
        cudaError_t e = h->metric == MUNINN_METRIC_L2       ? launch_filter_m<METRIC_L2>(h, maps, fp)
                        : h->metric == MUNINN_METRIC_COSINE ? launch_filter_m<METRIC_COSINE>(h, maps, fp)
                                                            : launch_filter_m<METRIC_IP>(h, maps, fp);
        CK(h, e);
        CK(h, cudaEventRecord(h->tc_ev[ev_used + 1], st));
        ev_used += 2;
        h->launches++;
        if (d_dbg) { chunks = 1; break; }
        const int rr_grid = (nq + 3) / 4;
        if (small) {
            if (h->metric == MUNINN_METRIC_L2)
                tc::exact_tc_rerank_cta<METRIC_L2><<<nq, tc::RR_THREADS, rr_smem, st>>>(h->d_vec, h->d_norm2, h->d_deleted, h->ld, h->dim, d_queries, nq, k, h->d_cand, h->d_cand_count, cap, h->d_topk, h->d_topk_size, h->d_tau, h->d_overflow, h->d_tc_ctr, rr_P);
            else if (h->metric == MUNINN_METRIC_COSINE)
                tc::exact_tc_rerank_cta<METRIC_COSINE><<<nq, tc::RR_THREADS, rr_smem, st>>>(h->d_vec, h->d_norm2, h->d_deleted, h->ld, h->dim, d_queries, nq, k, h->d_cand, h->d_cand_count, cap, h->d_topk, h->d_topk_size, h->d_tau, h->d_overflow, h->d_tc_ctr, rr_P);
            else
                tc::exact_tc_rerank_cta<METRIC_IP><<<nq, tc::RR_THREADS, rr_smem, st>>>(h->d_vec, h->d_norm2, h->d_deleted, h->ld, h->dim, d_queries, nq, k, h->d_cand, h->d_cand_count, cap, h->d_topk, h->d_topk_size, h->d_tau, h->d_overflow, h->d_tc_ctr, rr_P);
        } else if (h->metric == MUNINN_METRIC_L2)
            tc::exact_tc_rerank<METRIC_L2><<<rr_grid, 128, rr_smem, st>>>(h->d_vec, h->d_norm2, h->d_deleted, h->ld, h->dim, d_queries, nq, k, h->d_cand, h->d_cand_count, cap, h->d_topk, h->d_topk_size, h->d_tau, h->d_overflow, h->d_tc_ctr);
        else if (h->metric == MUNINN_METRIC_COSINE)
            tc::exact_tc_rerank<METRIC_COSINE><<<rr_grid, 128, rr_smem, st>>>(h->d_vec, h->d_norm2, h->d_deleted, h->ld, h->dim, d_queries, nq, k, h->d_cand, h->d_cand_count, cap, h->d_topk, h->d_topk_size, h->d_tau, h->d_overflow, h->d_tc_ctr);
        else
            tc::exact_tc_rerank<METRIC_IP><<<rr_grid, 128, rr_smem, st>>>(h->d_vec, h->d_norm2, h->d_deleted, h->ld, h->dim, d_queries, nq, k, h->d_cand, h->d_cand_count, cap, h->d_topk, h->d_topk_size, h->d_tau, h->d_overflow, h->d_tc_ctr);
        h->launches++;
        CK(h, cudaGetLastError());
        done += chunk;
        chunks++;
    }
    h->exact_stats.mode = h->exact_mode_run; h->exact_stats.chunks = chunks; h->exact_stats.c_eps = fp.c_eps;
    h->exact_stats.candidates = 0; h->exact_stats.overflow_queries = 0;
    if (d_dbg) {
        CK(h, cudaStreamSynchronize(st));
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, h->tc_ev[0], h->tc_ev[1]);
        h->exact_stats.filter_ms = ms;
        return MUNINN_CUDA_OK;
    }
    exact_emit<<<(unsigned)(((size_t)nq * k + 255) / 256), 256, 0, st>>>(h->d_topk, h->d_topk_size, nq, k, index_base, nullptr,
                                                                        d_idx, d_dist, d_count);
    int *d_nflag = reinterpret_cast<int *>(h->d_tc_ctr + 1);
    tc::tc_compact_overflow<<<(nq + 255) / 256, 256, 0, st>>>(h->d_overflow, nq, h->d_qmap, d_nflag);
    h->launches += 2;
    CK(h, cudaGetLastError());
    unsigned long long ctr[2] = {0, 0};
    CK(h, cudaMemcpyAsync(ctr, h->d_tc_ctr, 16, cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st)); /* the one host round trip of this path: how many queries overflowed */
    float fms = 0.0f;
    for (size_t i = 0; i + 1 < ev_used; i += 2) {
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, h->tc_ev[i], h->tc_ev[i + 1]);
        fms += ms;
    }
    const int nflag = (int)(unsigned)ctr[1];
    h->exact_stats.filter_ms = fms; h->exact_stats.candidates = (int64_t)ctr[0] + (int64_t)nq * seeded_rows; h->exact_stats.overflow_queries = nflag;
    if (nflag > 0) {
        CK(h, grow(&h->d_qgather, &h->qgather_floats, (size_t)nflag * h->dim, st));
        tc::tc_gather_queries<<<(unsigned)(((size_t)nflag * h->dim + 255) / 256), 256, 0, st>>>(d_queries, h->dim, h->d_qmap, nflag,
                                                                                             h->d_qgather);
        h->launches++;
        CK(h, cudaGetLastError());
        rc = exact_knn_cudacore(h, h->d_qgather, nflag, k, index_base, h->d_qmap, d_idx, d_dist, d_count);
        if (rc) return rc;
    }
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_exact_knn_device(MuninnCudaIndex *h, const float *d_queries, int nq, int k, int64_t index_base,
                                            int64_t *d_idx, float *d_dist, int32_t *d_count) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!h->uploaded) return fail(h, MUNINN_CUDA_E_STATE, "exact k-NN before muninn_cuda_index_upload");
    if (nq < 0 || k < 1 || k > 2048) return fail(h, MUNINN_CUDA_E_ARG, "need nq >= 0 and 1 <= k <= 2048 (nq=%d k=%d)", nq, k);
    if (nq == 0) return MUNINN_CUDA_OK;
    if (!d_queries || !d_idx || !d_dist || !d_count) return fail(h, MUNINN_CUDA_E_ARG, "NULL buffer");
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaEventRecord(h->ev0, h->stream));
    int rc;
    if (h->exact_mode == MUNINN_EXACT_CUDA_CORE || h->n == 0) {
        h->exact_stats = MuninnExactStats{};
        rc = exact_knn_cudacore(h, d_queries, nq, k, index_base, nullptr, d_idx, d_dist, d_count);
    } else {
        rc = exact_knn_tc(h, d_queries, nq, k, index_base, d_idx, d_dist, d_count, nullptr, 0);
    }
    if (rc) return rc;
    CK(h, cudaEventRecord(h->ev1, h->stream));
    h->have_timing = true;
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_set_exact_mode(MuninnCudaIndex *h, int mode, int tile_rows) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (mode < MUNINN_EXACT_CUDA_CORE || mode > MUNINN_EXACT_TC_AUTO) return fail(h, MUNINN_CUDA_E_ARG, "unknown exact mode %d", mode);
    if (tile_rows != 0 && tile_rows != 128 && tile_rows != 256) return fail(h, MUNINN_CUDA_E_ARG, "tile_rows must be 0, 128 or 256");
    h->exact_mode = mode;
    h->tc_tile_rows = tile_rows ? tile_rows : 256;
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_exact_stats(MuninnCudaIndex *h, MuninnExactStats *out) {
    if (!h || !out) return MUNINN_CUDA_E_ARG;
    *out = h->exact_stats;
    return MUNINN_CUDA_OK;
}

/* Error probe: the filter's contraction Σq·x for every (query, row) pair of rows [0, rows) → out_approx [nq × rows] (host). */
extern "C" int muninn_cuda_exact_filter_probe(MuninnCudaIndex *h, const float *queries, int nq, int64_t rows, float *out_approx) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!h->uploaded) return fail(h, MUNINN_CUDA_E_STATE, "probe before muninn_cuda_index_upload");
    if (h->exact_mode == MUNINN_EXACT_CUDA_CORE) return fail(h, MUNINN_CUDA_E_STATE, "probe needs a tensor-core exact mode");
    if (nq < 1 || rows < 1 || rows > h->n || !queries || !out_approx || (size_t)nq * (size_t)rows > (1ull << 28))
        return fail(h, MUNINN_CUDA_E_ARG, "bad probe arguments");
    CK(h, cudaSetDevice(h->device));
    CK(h, grow(&h->d_dbg, &h->dbg_floats, (size_t)nq * rows, h->stream));
    CK(h, grow(&h->d_qgather, &h->qgather_floats, (size_t)nq * h->dim, h->stream));
    CK(h, cudaMemcpyAsync(h->d_qgather, queries, (size_t)nq * h->dim * 4, cudaMemcpyHostToDevice, h->stream));
    int rc = exact_knn_tc(h, h->d_qgather, nq, 1, 0, nullptr, nullptr, nullptr, h->d_dbg, rows);
    if (rc) return rc;
    CK(h, cudaMemcpy(out_approx, h->d_dbg, (size_t)nq * rows * 4, cudaMemcpyDeviceToHost));
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_exact_knn(MuninnCudaIndex *h, const float *queries, int nq, int k, int64_t index_base,
                                     int64_t *out_idx, float *out_dist, int32_t *out_count) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!h->uploaded) return fail(h, MUNINN_CUDA_E_STATE, "exact k-NN before muninn_cuda_index_upload");
    if (nq < 0 || k < 1) return fail(h, MUNINN_CUDA_E_ARG, "nq must be >= 0 and k >= 1");
    if (nq == 0) return MUNINN_CUDA_OK;
    if (!queries || !out_idx || !out_dist || !out_count) return fail(h, MUNINN_CUDA_E_ARG, "NULL buffer");
    CK(h, cudaSetDevice(h->device));
    const size_t bq = align256((size_t)nq * h->dim * 4), bi = align256((size_t)nq * k * 8),
                 bd = align256((size_t)nq * k * 4), bc = align256((size_t)nq * 4);
    int rc = ensure_stage(h, bq + bi + bd + bc);
    if (rc) return rc;
    char *b = (char *)h->d_stage;
    float *dq = (float *)b; int64_t *di = (int64_t *)(b + bq); float *dd = (float *)(b + bq + bi);
    int32_t *dc = (int32_t *)(b + bq + bi + bd);
    CK(h, cudaMemcpyAsync(dq, queries, (size_t)nq * h->dim * 4, cudaMemcpyHostToDevice, h->stream));
    rc = muninn_cuda_exact_knn_device(h, dq, nq, k, index_base, di, dd, dc);
    if (rc) return rc;
    CK(h, cudaMemcpyAsync(out_idx, di, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaMemcpyAsync(out_dist, dd, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaMemcpyAsync(out_count, dc, (size_t)nq * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_topk_merge_device(MuninnCudaIndex *h, int parts, int nq, int k, const int64_t *d_idx_parts,
                                             const float *d_dist_parts, const int32_t *d_count_parts, int64_t *d_idx,
                                             float *d_dist, int32_t *d_count) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (parts < 1 || nq < 0 || k < 1 || k > 2048) return fail(h, MUNINN_CUDA_E_ARG, "bad parts/nq/k");
    if (nq == 0) return MUNINN_CUDA_OK;
    CK(h, cudaSetDevice(h->device));
    const size_t smem = (size_t)4 * (k + 32) * 8;
    if (smem > 48 * 1024)
        CK(h, cudaFuncSetAttribute(topk_merge_parts, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    topk_merge_parts<<<(nq + 3) / 4, 128, smem, h->stream>>>(parts, nq, k, d_idx_parts, d_dist_parts, d_count_parts, d_idx,
                                                             d_dist, d_count);
    h->launches++;
    CK(h, cudaGetLastError());
    return MUNINN_CUDA_OK;
}

/* ───────────────────────── sharded exact k-NN: one NCCL all-gather ───────────────────────── */

/* NCCL is bound at run time: the library a host process already carries (torch's, an MPI launcher's) is the one whose
 * communicators it hands us, so the entry points are looked up in that instance first; a plain C host gets the
 * system's libnccl.so.2. Only the five calls below are used; their prototypes are NCCL's public ABI (nccl.h). */
namespace {
typedef int nccl_result_t;
typedef struct { char internal[128]; } nccl_unique_id_t;
struct NcclApi {
    nccl_result_t (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    nccl_result_t (*GetUniqueId)(nccl_unique_id_t *) = nullptr;
    nccl_result_t (*CommInitRank)(void **, int, nccl_unique_id_t, int) = nullptr;
    nccl_result_t (*CommDestroy)(void *) = nullptr;
    nccl_result_t (*CommCount)(void *, int *) = nullptr;
    const char *(*GetErrorString)(nccl_result_t) = nullptr;
    bool ok = false;
    std::string why;
};
const int NCCL_UINT64 = 5; /* ncclUint64, nccl.h ncclDataType_t */
}
#include <dlfcn.h>
static NcclApi &nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD); /* the instance already in this process, if any */
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!lib) { api.why = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : ""); return api; }
    api.AllGather = (decltype(api.AllGather))dlsym(lib, "ncclAllGather");
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(lib, "ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(lib, "ncclCommDestroy");
    api.CommCount = (decltype(api.CommCount))dlsym(lib, "ncclCommCount");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(lib, "ncclGetErrorString");
    api.ok = api.AllGather && api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.CommCount && api.GetErrorString;
    if (!api.ok) api.why = "libnccl.so.2 lacks an expected symbol";
    return api;
}

extern "C" int muninn_cuda_comm_unique_id(void *out_id128) {
    if (!out_id128) return MUNINN_CUDA_E_ARG;
    NcclApi &n = nccl_api();
    if (!n.ok) return fail(nullptr, MUNINN_CUDA_E_STATE, "%s", n.why.c_str());
    nccl_unique_id_t id;
    const nccl_result_t r = n.GetUniqueId(&id);
    if (r != 0) return fail(nullptr, MUNINN_CUDA_E_CUDA, "ncclGetUniqueId: %s", n.GetErrorString(r));
    memcpy(out_id128, &id, sizeof id);
    return MUNINN_CUDA_OK;
}

extern "C" int muninn_cuda_comm_create(int device, int world, int rank, const void *id128, void **out_comm) {
    if (!id128 || !out_comm || world < 1 || rank < 0 || rank >= world) return MUNINN_CUDA_E_ARG;
    *out_comm = nullptr;
    if (muninn_cuda_device_count() <= 0) return fail(nullptr, MUNINN_CUDA_E_NODEVICE, "no CUDA device visible");
    NcclApi &n = nccl_api();
    if (!n.ok) return fail(nullptr, MUNINN_CUDA_E_STATE, "%s", n.why.c_str());
    if (cudaSetDevice(device) != cudaSuccess) return fail(nullptr, MUNINN_CUDA_E_ARG, "device %d not usable", device);
    nccl_unique_id_t id;
    memcpy(&id, id128, sizeof id);
    const nccl_result_t r = n.CommInitRank(out_comm, world, id, rank);
    if (r != 0) return fail(nullptr, MUNINN_CUDA_E_CUDA, "ncclCommInitRank: %s", n.GetErrorString(r));
    return MUNINN_CUDA_OK;
}

extern "C" void muninn_cuda_comm_destroy(void *comm) {
    NcclApi &n = nccl_api();
    if (comm && n.ok) n.CommDestroy(comm);
}

extern "C" int muninn_cuda_exact_knn_sharded_device(MuninnCudaIndex *h, void *nccl_comm, const float *d_queries, int nq, int k,
                                                    int64_t index_base, int64_t *d_idx, float *d_dist, int32_t *d_count) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (!nccl_comm) return fail(h, MUNINN_CUDA_E_ARG, "NULL communicator");
    if (index_base < 0 || index_base + h->n > 0xffffffffLL) return fail(h, MUNINN_CUDA_E_ARG, "global row indices must fit 32 bits");
    NcclApi &n = nccl_api();
    if (!n.ok) return fail(h, MUNINN_CUDA_E_STATE, "%s", n.why.c_str());
    int world = 0;
    nccl_result_t r = n.CommCount(nccl_comm, &world);
    if (r != 0 || world < 1) return fail(h, MUNINN_CUDA_E_CUDA, "ncclCommCount: %s", n.GetErrorString(r));
    /* this shard's lists land in the caller's output arrays first (global indices), are packed into keys … */
    int rc = muninn_cuda_exact_knn_device(h, d_queries, nq, k, index_base, d_idx, d_dist, d_count);
    if (rc || nq == 0) return rc;
    const size_t cnt = (size_t)nq * k;
    if ((size_t)(world + 1) * cnt > h->shard_keys_cap) {
        CK(h, cudaStreamSynchronize(h->stream));
        cudaFree(h->d_shard_keys); h->d_shard_keys = nullptr; h->shard_keys_cap = 0;
        CK(h, cudaMalloc((void **)&h->d_shard_keys, (size_t)(world + 1) * cnt * 8));
        h->shard_keys_cap = (size_t)(world + 1) * cnt;
    }
    unsigned long long *mine = h->d_shard_keys, *all = h->d_shard_keys + cnt;
    topk_pack_keys<<<(unsigned)((cnt + 255) / 256), 256, 0, h->stream>>>(nq, k, d_idx, d_dist, d_count, mine);
    h->launches++;
    CK(h, cudaGetLastError());
    /* … exchanged with ONE all-gather of nq·k·8 bytes per rank over NVLink, and merged under (distance, index) */
    r = n.AllGather(mine, all, cnt, NCCL_UINT64, nccl_comm, h->stream);
    if (r != 0) return fail(h, MUNINN_CUDA_E_CUDA, "ncclAllGather: %s", n.GetErrorString(r));
    const size_t smem = (size_t)4 * (k + 32) * 8;
    if (smem > 48 * 1024) CK(h, cudaFuncSetAttribute(topk_merge_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    topk_merge_keys<<<(nq + 3) / 4, 128, smem, h->stream>>>(world, nq, k, all, d_idx, d_dist, d_count);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaEventRecord(h->ev1, h->stream)); /* last_kernel_ms now spans scan + exchange + merge */
    return MUNINN_CUDA_OK;
}

/* the merge alone, for callers that gathered the key lists themselves (tests: gloo on CPU, shards on one GPU) */
extern "C" int muninn_cuda_topk_pack_keys_device(MuninnCudaIndex *h, int nq, int k, const int64_t *d_idx, const float *d_dist,
                                                 const int32_t *d_count, uint64_t *d_keys) {
    if (!h || nq < 0 || k < 1) return MUNINN_CUDA_E_ARG;
    if (nq == 0) return MUNINN_CUDA_OK;
    CK(h, cudaSetDevice(h->device));
    const size_t cnt = (size_t)nq * k;
    topk_pack_keys<<<(unsigned)((cnt + 255) / 256), 256, 0, h->stream>>>(nq, k, d_idx, d_dist, d_count, (unsigned long long *)d_keys);
    h->launches++;
    CK(h, cudaGetLastError());
    return MUNINN_CUDA_OK;
}
extern "C" int muninn_cuda_topk_merge_keys_device(MuninnCudaIndex *h, int parts, int nq, int k, const uint64_t *d_keys,
                                                  int64_t *d_idx, float *d_dist, int32_t *d_count) {
    if (!h) return MUNINN_CUDA_E_ARG;
    if (parts < 1 || nq < 0 || k < 1 || k > 2048) return fail(h, MUNINN_CUDA_E_ARG, "bad parts/nq/k");
    if (nq == 0) return MUNINN_CUDA_OK;
    CK(h, cudaSetDevice(h->device));
    const size_t smem = (size_t)4 * (k + 32) * 8;
    if (smem > 48 * 1024) CK(h, cudaFuncSetAttribute(topk_merge_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    topk_merge_keys<<<(nq + 3) / 4, 128, smem, h->stream>>>(parts, nq, k, (const unsigned long long *)d_keys, d_idx, d_dist, d_count);
    h->launches++;
    CK(h, cudaGetLastError());
    return MUNINN_CUDA_OK;
}

/* ───────────────────────── distances (known-answer surface) ───────────────────────── */

extern "C" int muninn_cuda_distance_batch(int device, int metric, int dim, int64_t n, const float *a, const float *b,
                                          float *out) {
    if (metric < 0 || metric > 2 || dim < 1 || n < 0 || (n > 0 && (!a || !b || !out))) return MUNINN_CUDA_E_ARG;
    if (muninn_cuda_device_count() <= 0) return MUNINN_CUDA_E_NODEVICE;
    if (n == 0) return MUNINN_CUDA_OK;
    if (cudaSetDevice(device) != cudaSuccess) return MUNINN_CUDA_E_ARG;
    float *da = nullptr, *db = nullptr, *dout = nullptr;
    const size_t bytes = (size_t)n * dim * 4;
    int rc = MUNINN_CUDA_OK;
    if (cudaMalloc((void **)&da, bytes) != cudaSuccess || cudaMalloc((void **)&db, bytes) != cudaSuccess ||
        cudaMalloc((void **)&dout, (size_t)n * 4) != cudaSuccess)
        rc = MUNINN_CUDA_E_NOMEM;
    if (!rc && (cudaMemcpy(da, a, bytes, cudaMemcpyHostToDevice) != cudaSuccess ||
                cudaMemcpy(db, b, bytes, cudaMemcpyHostToDevice) != cudaSuccess))
        rc = MUNINN_CUDA_E_CUDA;
    if (!rc) {
        distance_pairs_kernel<<<(unsigned)((n + 127) / 128), 128>>>(metric, dim, n, da, db, dout);
        if (cudaGetLastError() != cudaSuccess || cudaMemcpy(out, dout, (size_t)n * 4, cudaMemcpyDeviceToHost) != cudaSuccess)
            rc = MUNINN_CUDA_E_CUDA;
    }
    cudaFree(da); cudaFree(db); cudaFree(dout);
    return rc;
}

/* ───────────────────────── access-pattern ceiling (diagnostic) ───────────────────────── */

/* Independent random gathers of whole rows (16-byte chunk per lane, one coalesced request per 512
 * bytes), two rows in flight per warp, 64 warps per SM, no dependent work: what the memory system
 * delivers for the search kernel's access pattern when nothing else is in the way. */
__global__ void __launch_bounds__(256) gather_ceiling_kernel(const float4 *__restrict__ slab, unsigned long long row_mask,
                                                             int chunks_per_row, int iters, float *out) {
    const int lane = threadIdx.x & 31;
    unsigned long long s = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 0x9E3779B97F4A7C15ull + 12345;
    float acc = 0.f;
    for (int it = 0; it < iters; it++) {
        float4 v[2];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            s = s * 6364136223846793005ull + 1442695040888963407ull;
            const float4 *p = slab + ((s >> 24) & row_mask) * (unsigned long long)chunks_per_row;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int c = lane; c < chunks_per_row; c += 32) {
                const float4 t = ldg_stream_f4(reinterpret_cast<const float *>(p + c));
                a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
            }
            v[r] = a;
        }
        acc += v[0].x + v[0].y + v[0].z + v[0].w + v[1].x + v[1].y + v[1].z + v[1].w;
    }
    if (acc == 123.456f) out[0] = acc;
}

extern "C" int muninn_cuda_gather_ceiling(int device, int row_bytes, int64_t slab_bytes, float *out_gbps) {
    if (!out_gbps || row_bytes < 16 || (row_bytes & 15) || slab_bytes < (int64_t)row_bytes * 1024) return MUNINN_CUDA_E_ARG;
    if (muninn_cuda_device_count() <= 0) return MUNINN_CUDA_E_NODEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return MUNINN_CUDA_E_ARG;
    unsigned long long rows = 1;
    while (2 * rows * (unsigned long long)row_bytes <= (unsigned long long)slab_bytes) rows *= 2;
    float4 *slab = nullptr; float *out = nullptr;
    int sms = 0, rc = MUNINN_CUDA_OK;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (cudaMalloc((void **)&slab, rows * row_bytes) != cudaSuccess || cudaMalloc((void **)&out, 4) != cudaSuccess) rc = MUNINN_CUDA_E_NOMEM;
    if (!rc && (cudaMemset(slab, 0, rows * row_bytes) != cudaSuccess || cudaEventCreate(&e0) != cudaSuccess ||
                cudaEventCreate(&e1) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess))
        rc = MUNINN_CUDA_E_CUDA;
    if (!rc) {
        const int grid = sms * 8, iters = (int)std::max<long long>(50, (1ll << 34) / ((long long)grid * 8 * 2 * row_bytes));
        gather_ceiling_kernel<<<grid, 256>>>(slab, rows - 1, row_bytes / 16, 20, out);
        float best = 1e30f;
        for (int rep = 0; rep < 3 && !rc; rep++) {
            cudaEventRecord(e0);
            gather_ceiling_kernel<<<grid, 256>>>(slab, rows - 1, row_bytes / 16, iters, out);
            cudaEventRecord(e1);
            float ms = 0.f;
            if (cudaEventSynchronize(e1) != cudaSuccess || cudaEventElapsedTime(&ms, e0, e1) != cudaSuccess) rc = MUNINN_CUDA_E_CUDA;
            best = std::min(best, ms);
        }
        if (!rc) *out_gbps = (float)((double)grid * 8 * iters * 2 * row_bytes / best / 1e6);
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(slab); cudaFree(out);
    return rc;
}
