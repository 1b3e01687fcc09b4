/*
 * exact_tc.cuh — exact k-NN, tensor-core path (sm_100a: TMA → tcgen05.mma kind::tf32 → TMEM).
 *
 * Output contract is unchanged: the brute_force_knn helper of the reference's tests
 * (test/test_hnsw_algo.c:10-34) with the reference's distance arithmetic (src/vec_math.c:78-143),
 * ascending by (distance, row index), bit-exact against the CPU oracle. Bit-exact fp32 chains
 * cannot be produced by the tensor pipe, so the tensor pipe is used as a PROVABLE FILTER and the
 * few survivors are re-evaluated with the reference's arithmetic:
 *
 *   1. split      x = hi + lo with hi = x & 0xffffe000 (a TF32 value, so the tensor core sees it
 *                 unchanged whatever its input rounding), lo = x − hi (exact in fp32), lo cut the
 *                 same way. Rows of a chunk and the queries are split once into K-padded slabs.
 *   2. filter     tiled Q·Xᵀ as 3xTF32 (lo·hi + hi·lo + hi·hi; NPASS = 1 keeps hi·hi only):
 *                 128 queries × BN rows per CTA tile, operands staged by TMA into 128B-swizzled
 *                 shared memory, tcgen05.mma issued by one thread, fp32 accumulators in TMEM,
 *                 double-buffered so the epilogue of tile i overlaps the MMAs of tile i+1.
 *                 Fused per-tile selection in the epilogue: thread = one query (one TMEM lane);
 *                 a row survives unless its approximate distance provably exceeds the query's
 *                 current k-th exact distance τ_q:   d̃ > τ_q + ε(q, x).
 *                 Survivors are parked in a per-thread shared-memory list and appended to the
 *                 query's global list 16 at a time (one atomic per survivor stalled the warp).
 *                 1xTF32 reads the raw fp32 slab (the MMA drops the low mantissa bits itself): no
 *                 split pass, the base is read once — the mode of HBM-bound small batches.
 *   3. re-rank    reference-order fp32 distance of every survivor, k smallest kept as the query's
 *                 sorted (distance, row) list; τ_q ← its k-th entry. One warp per query for large
 *                 batches; one CTA of 16 warps per query (TMA-staged rows, radix select) below
 *                 1 024 queries.
 *
 * The base is processed in chunks that grow geometrically (×4; ×17–32 for small batches); the first
 * chunk is small because every row of it survives, τ being +inf. A chunk g times the rows seen so
 * far yields ≈ g·k survivors per query.
 *
 * Soundness: a row of the true top-k has d_ref ≤ τ_q at all times; with |d̃ − d_ref| ≤ ε it can
 * never satisfy d̃ > τ_q + ε. ε = c·(‖q‖² + ‖x‖²) (L2, inner product) or c (cosine) with
 * c = 2·(3·2⁻²⁰ + 3K·2⁻²³) + 4(K/4 + 8)·2⁻²⁴ for 3xTF32 — twice the worst-case split + fp32
 * accumulation error of the K-term contraction plus the rounding of the reference's own chain and
 * of the norms; tests/test_gpu_exact_tc.py measures the actual error against this bound. The test
 * is written `!(d̃ > thr)` so a NaN anywhere keeps the row. A query whose survivors overflow its
 * candidate list (massive ties, e.g. duplicated rows) is flagged and recomputed by the CUDA-core
 * path (exact_knn.cuh) — still exact, only slower.
 */
#pragma once
#include "muninn_dev.cuh"
#include "hnsw_search.cuh" /* warp_dist_rows: TMA-staged rows, the reference's chain order */
#include <cuda.h> /* CUtensorMap and its enums only; cuTensorMapEncodeTiled is looked up at run time */

namespace muninn {
namespace tc {

constexpr int BM = 128;          /* queries per tile = TMEM lanes */
constexpr int BK = 32;           /* fp32 per k-block: one 128-byte swizzle row */
constexpr int UK = 8;            /* K of one tcgen05.mma.kind::tf32 */
constexpr int THREADS = 192;     /* warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 epilogue */
constexpr unsigned TF32_MASK = 0xffffe000u;

/* ───────────── PTX wrappers ───────────── */

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done;
    const uint32_t a = smem_u32(bar);
    do {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(a), "r"(parity)
                     : "memory");
    } while (!done);
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap *map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
/* arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed */
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
/* D[tmem] (+)= A[smem] · B[smem]ᵀ, both operands K-major, TF32 inputs, fp32 accumulate */
__device__ __forceinline__ void mma_tf32_ss(uint32_t d_tmem, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
                 : "memory");
}
/* 32 lanes × 32 consecutive columns → 32 registers per thread (thread = lane) */
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

/* ── CTA pair (cta_group::2): two CTAs of a cluster on the two SMs of a TPC issue ONE M = 256 MMA; each keeps its own
 * 128 rows of A and one half of B in its shared memory, its 128 rows of D in its tensor memory ── */
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_id_x() { uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_count_x() { uint32_t r; asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
/* shared::cluster address of the same shared-memory variable in CTA `rank` of the cluster */
__device__ __forceinline__ uint32_t map_to_cta(const void *p, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
/* TMA load into THIS CTA's shared memory whose completion is counted on a barrier of the pair's leader */
__device__ __forceinline__ void tma_load_2d_pair(void *dst, const CUtensorMap *map, uint32_t leader_bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
/* arrives on the barrier at this shared-memory offset in BOTH CTAs once every MMA issued so far has completed */
__device__ __forceinline__ void tc_commit_pair(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mma_tf32_ss_pair(uint32_t d_tmem, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
                 : "memory");
}

/* shared-memory matrix descriptor of a K-major tile whose rows are 128 bytes, 128B-swizzled (what
 * TMA writes with CU_TENSOR_MAP_SWIZZLE_128B): 8-row groups 1024 bytes apart, sm_100 descriptor
 * version 1. The tile base must be 1024-byte aligned; a K step of UK elements inside the swizzle
 * row is an address advance of UK*4 bytes. */
__device__ __forceinline__ uint64_t make_desc_k128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffffu) >> 4);
    d |= (uint64_t)1 << 16;             /* leading byte offset: unused with a swizzled K-major tile */
    d |= (uint64_t)(1024 >> 4) << 32;   /* stride byte offset between 8-row groups */
    d |= (uint64_t)1 << 46;             /* descriptor version (Blackwell) */
    d |= (uint64_t)2 << 61;             /* SWIZZLE_128B */
    return d;
}
/* instruction descriptor: D fp32, A and B TF32, both K-major, M = 128, N = bn */
__host__ __device__ constexpr uint32_t make_idesc_tf32(int bn, int bm = BM) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(bm >> 4) << 24);
}

/* ───────────── filter kernel ───────────── */

struct FilterParams {
    const float *xnorm2; /* [n] Σx² per row of the handle, reference chain order */
    const float *qaux;   /* [nq] Σq², reference chain order */
    const float *tau;    /* [nq] current k-th exact distance (+inf while fewer than k rows were seen) */
    int *cand;           /* [nq × cand_cap] surviving row indices of this chunk */
    int *cand_count;     /* [nq] */
    int cand_cap;
    long long row0;      /* first row of the chunk (the x tensor maps start there) */
    int rows, nq, kblocks;
    float c_eps;
    float *dbg_panel;    /* DBG instantiation only, [nq × dbg_ld]: the contraction Σq·x of every pair (error probe) */
    int dbg_ld;
};

constexpr int SBUF = 16; /* survivors an epilogue thread parks in shared memory before one atomic appends them */
template <int NPASS, int BN>
struct FilterCfg {
    static constexpr int SLABS = NPASS == 3 ? 2 : 1;
    static constexpr int A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
    static constexpr int STAGE_BYTES = SLABS * (A_BYTES + B_BYTES);
    static constexpr int NSTAGES = (200 * 1024) / STAGE_BYTES > 8 ? 8 : (200 * 1024) / STAGE_BYTES;
    static constexpr int TMEM_COLS = 2 * BN;
    /* stages | xaux[2][BN] | barriers (2·NSTAGES + 4) | tmem pointer | sbuf[128][SBUF] ; + 1024 alignment slack */
    static constexpr size_t SMEM_BYTES = 1024 + (size_t)NSTAGES * STAGE_BYTES + 2 * BN * 4 + (2 * NSTAGES + 4) * 8 + 16 + 128 * SBUF * 4;
};

template <int METRIC, int NPASS, int BN, bool DBG>
__global__ void __launch_bounds__(THREADS, 1)
exact_tc_filter(const __grid_constant__ CUtensorMap tm_qhi, const __grid_constant__ CUtensorMap tm_qlo,
                const __grid_constant__ CUtensorMap tm_xhi, const __grid_constant__ CUtensorMap tm_xlo,
                const FilterParams p) {
    using Cfg = FilterCfg<NPASS, BN>;
    constexpr int NST = Cfg::NSTAGES;
    extern __shared__ unsigned char smem_dyn[];
    /* 1024-byte aligned base (128B swizzle atom); plain offset arithmetic keeps the shared address space */
    unsigned char *base = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    unsigned char *stages = base;
    float *xaux = reinterpret_cast<float *>(base + (size_t)NST * Cfg::STAGE_BYTES);
    uint64_t *full = reinterpret_cast<uint64_t *>(xaux + 2 * BN);
    uint64_t *empty = full + NST;
    uint64_t *tmem_full = empty + NST;
    uint64_t *tmem_empty = tmem_full + 2;
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(tmem_empty + 2);
    int *sbuf = reinterpret_cast<int *>(tmem_ptr + 4); /* [128 epilogue threads × SBUF] parked survivors */

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_qt = (p.nq + BM - 1) / BM, n_xt = (p.rows + BN - 1) / BN;
    const long long n_tiles = (long long)n_qt * n_xt;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_qhi); prefetch_tmap(&tm_xhi);
        if (NPASS == 3) { prefetch_tmap(&tm_qlo); prefetch_tmap(&tm_xlo); }
        for (int s = 0; s < NST; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int a = 0; a < 2; a++) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 4); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, Cfg::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        /* ── TMA producer: one thread streams the k-blocks of every tile through the stage ring ── */
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int xt = (int)(t / n_qt), qt = (int)(t % n_qt);
                for (int kb = 0; kb < p.kblocks; kb++) {
                    mbar_wait(&empty[s], ph ^ 1);
                    unsigned char *st = stages + (size_t)s * Cfg::STAGE_BYTES;
                    mbar_expect_tx(&full[s], Cfg::STAGE_BYTES);
                    tma_load_2d(st, &tm_qhi, &full[s], kb * BK, qt * BM);
                    tma_load_2d(st + Cfg::A_BYTES, &tm_xhi, &full[s], kb * BK, xt * BN);
                    if (NPASS == 3) {
                        tma_load_2d(st + Cfg::A_BYTES + Cfg::B_BYTES, &tm_qlo, &full[s], kb * BK, qt * BM);
                        tma_load_2d(st + 2 * Cfg::A_BYTES + Cfg::B_BYTES, &tm_xlo, &full[s], kb * BK, xt * BN);
                    }
                    if (++s == NST) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        /* ── MMA issuer: one thread; accumulator stage a = tile parity ── */
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(BN);
            int s = 0; uint32_t ph = 0; int a = 0; uint32_t aph = 0;
            for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                mbar_wait(&tmem_empty[a], aph ^ 1); /* the epilogue has drained this accumulator */
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(a * BN);
                for (int kb = 0; kb < p.kblocks; kb++) {
                    mbar_wait(&full[s], ph);
                    tc_fence_after();
                    const uint32_t st = smem_u32(stages + (size_t)s * Cfg::STAGE_BYTES);
                    const uint64_t a_hi = make_desc_k128(st), b_hi = make_desc_k128(st + Cfg::A_BYTES);
                    uint32_t acc = kb > 0 ? 1u : 0u;
                    if (NPASS == 3) { /* small terms first: lo·hi, hi·lo, then hi·hi */
                        const uint64_t a_lo = make_desc_k128(st + Cfg::A_BYTES + Cfg::B_BYTES),
                                       b_lo = make_desc_k128(st + 2 * Cfg::A_BYTES + Cfg::B_BYTES);
#pragma unroll
                        for (int k4 = 0; k4 < BK / UK; k4++) { mma_tf32_ss(d_tmem, a_lo + 2 * k4, b_hi + 2 * k4, idesc, acc); acc = 1u; }
#pragma unroll
                        for (int k4 = 0; k4 < BK / UK; k4++) mma_tf32_ss(d_tmem, a_hi + 2 * k4, b_lo + 2 * k4, idesc, 1u);
                    }
#pragma unroll
                    for (int k4 = 0; k4 < BK / UK; k4++) { mma_tf32_ss(d_tmem, a_hi + 2 * k4, b_hi + 2 * k4, idesc, acc); acc = 1u; }
                    tc_commit(&empty[s]); /* frees the stage once these MMAs have read it */
                    if (kb == p.kblocks - 1) tc_commit(&tmem_full[a]);
                    if (++s == NST) { s = 0; ph ^= 1; }
                }
                if (++a == 2) { a = 0; aph ^= 1; }
            }
        }
    } else {
        /* ── epilogue: 4 warps, warp w reads TMEM lanes 32·(w mod 4)…, thread = one query ──
         * With s = Σ q·x from TMEM, "d̃ > τ + ε" is rearranged so that one add (or multiply) and one
         * compare per pair remain, everything else being per query or per row:
         *   L2      (qn+xn) − 2s > τ + c(qn+xn)   ⇔  s < [h·qn − τ/2] + [h·xn],  h = (1−c)/2
         *   IP      −s > τ + c(qn+xn)             ⇔  s < [−τ − c·qn] + [−c·xn]
         *   cosine  1 − s/(√qn√xn) > τ + c        ⇔  s·(1/√xn) < (1 − τ − c)·√qn
         * (the few extra roundings are ~2⁻²³ relative, far inside the factor-2 slack of c). A pair is
         * pruned only when the compare is TRUE, so NaN/inf anywhere keeps the row. */
        const int quarter = warp & 3, et = (warp - 2) * 32 + lane;
        const int qrow = quarter * 32 + lane;
        const float c = p.c_eps, hh = 0.5f * (1.0f - c);
        int a = 0; uint32_t aph = 0;
        /* the epilogue is the critical role when the batch is small (one MMA per 128 KB of rows): nothing in it may
         * wait on global memory. The per-row terms of the NEXT tile are fetched while this one is processed, the
         * per-query term is recomputed only when the query tile changes, and each 32-column TMEM read is issued
         * before the previous one is consumed. */
        constexpr int XPT = BN / 128;
        float xn_pre[XPT];
        auto fetch_xn = [&](long long t2) {
#pragma unroll
            for (int u = 0; u < XPT; u++) {
                xn_pre[u] = 0.0f;
                if (t2 < n_tiles) {
                    const int xt2 = (int)(t2 / n_qt), j = et + 128 * u;
                    if (xt2 * BN + j < p.rows) xn_pre[u] = __ldg(p.xnorm2 + p.row0 + (long long)xt2 * BN + j);
                }
            }
        };
        fetch_xn(blockIdx.x);
        int qt_have = -1;
        float qa = 0.0f;
        bool qvalid = false;
        int q = 0;
        int *mybuf = sbuf + et * SBUF;
        int nbuf = 0;
        auto flush = [&]() { /* this thread's parked survivors → the list of query q */
            if (nbuf == 0) return;
            const int pos = atomicAdd(p.cand_count + q, nbuf);
            for (int i = 0; i < nbuf; i++)
                if (pos + i < p.cand_cap) p.cand[(size_t)q * p.cand_cap + pos + i] = mybuf[i];
            nbuf = 0;
        };
        for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
            const int xt = (int)(t / n_qt), qt = (int)(t % n_qt);
            const int ncols = p.rows - xt * BN < BN ? p.rows - xt * BN : BN;
            const long long xrow0 = p.row0 + (long long)xt * BN;
            if (qt != qt_have) {
                flush(); /* the list belongs to the query this thread served so far */
                qt_have = qt;
                q = qt * BM + qrow;
                qvalid = q < p.nq;
                qa = 0.0f;
                if (qvalid) {
                    const float qn = p.qaux[q], tq = __ldcg(p.tau + q) + 1e-30f;
                    if (METRIC == METRIC_L2) qa = fmaf(hh, qn, -0.5f * tq);
                    else if (METRIC == METRIC_IP) qa = fmaf(-c, qn, -tq);
                    else qa = qn < 1e-18f ? __int_as_float(0x7fc00000) : (1.0f - tq - c) * sqrtf(qn);
                }
            }
            float *xa = xaux + a * BN;
#pragma unroll
            for (int u = 0; u < XPT; u++) {
                const int j = et + 128 * u;
                float v = 0.0f;
                if (j < ncols) {
                    const float xn = xn_pre[u];
                    if (METRIC == METRIC_L2) v = hh * xn;
                    else if (METRIC == METRIC_IP) v = -c * xn;
                    else v = xn < 1e-18f ? __int_as_float(0x7fc00000) : rsqrtf(xn);
                }
                xa[j] = v;
            }
            fetch_xn(t + gridDim.x); /* in flight during this tile's columns */
            asm volatile("bar.sync 1, 128;" ::: "memory");
            mbar_wait(&tmem_full[a], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(a * BN);
            auto consume = [&](const uint32_t (&v)[32], int c0) {
                if (DBG) {
                    if (qvalid)
                        for (int j = 0; j < 32 && c0 + j < ncols; j++)
                            p.dbg_panel[(size_t)q * p.dbg_ld + (size_t)(xrow0 - p.row0) + c0 + j] = __uint_as_float(v[j]);
                    return;
                }
                uint32_t keep = 0;
#pragma unroll
                for (int j4 = 0; j4 < 32; j4 += 4) {
                    const float4 x4 = *reinterpret_cast<const float4 *>(xa + c0 + j4);
                    const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
                    for (int jj = 0; jj < 4; jj++) {
                        const float sdot = __uint_as_float(v[j4 + jj]);
                        const bool prune = METRIC == METRIC_COSINE ? (sdot * xs[jj] < qa) : (sdot < qa + xs[jj]);
                        keep |= prune ? 0u : (1u << (j4 + jj));
                    }
                }
                if (ncols - c0 < 32) keep &= (1u << (ncols - c0)) - 1u;
                if (!qvalid) keep = 0;
                /* Survivors are rare per pair (≈ 0.1 %) but not per warp-step (32 queries × 32 rows): appending each
                 * with its own atomic made most steps wait for an L2 round trip — that, not the compares, was the whole
                 * cost of this epilogue (0.27 of 0.76 ms at 6.25M×128, nq=100). They are parked in a per-thread
                 * shared-memory list instead and appended SBUF at a time with one atomic. */
                while (keep) {
                    const int j = __ffs(keep) - 1;
                    keep &= keep - 1;
                    mybuf[nbuf++] = (int)(xrow0 + c0 + j);
                    if (nbuf == SBUF) flush();
                }
            };
            uint32_t va[32], vb[32];
            tmem_ld32(taddr, va);
            for (int c0 = 0; c0 < ncols; c0 += 64) {
                tmem_ld_wait();
                const bool more = c0 + 32 < ncols;
                if (more) tmem_ld32(taddr + (uint32_t)(c0 + 32), vb);
                consume(va, c0);
                if (more) {
                    tmem_ld_wait();
                    if (c0 + 64 < ncols) tmem_ld32(taddr + (uint32_t)(c0 + 64), va);
                    consume(vb, c0 + 32);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[a]);
            if (++a == 2) { a = 0; aph ^= 1; }
        }
        flush();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, Cfg::TMEM_COLS); }
}

/* ───────────── filter kernel, query-stationary form (one TF32 product) ─────────────
 *
 * The streaming kernel above re-fetches BOTH operands of every tile: 192 KB of TMA traffic for 8.4 MFLOP at K = 128,
 * i.e. ≈ 19 TB/s of L2→shared-memory traffic to keep the tensor pipe busy — more than the L2 delivers (measured
 * 3.2 µs per tile against 1.47 µs of MMA). Here the QUERY tile (128 × K, ≤ 128 KB) stays in shared memory and only
 * base-row tiles stream past it, a third less traffic per tile; to keep the base rows coming out of L2 rather than
 * HBM while ⌈nq/128⌉ query tiles each walk them, the row tiles are taken in super-blocks of SB tiles (≈ 32 MB of
 * rows): inside a super-block the (query tile, row tile) pairs are dealt out query-major in contiguous runs, so a
 * CTA keeps one query tile for ≈ 100 row tiles, and all CTAs sweep the same L2-resident super-block together.
 */
struct TileWalk {
    int n_qt, n_xt, SB, sb, nx;
    long long j, end;
    __device__ __forceinline__ void start_sb(int cta, int grid) {
        nx = n_xt - sb * SB < SB ? n_xt - sb * SB : SB;
        const long long T = (long long)n_qt * nx;
        j = T * cta / grid;
        end = T * (cta + 1) / grid;
    }
    __device__ __forceinline__ TileWalk(int n_qt_, int n_xt_, int SB_, int cta, int grid) : n_qt(n_qt_), n_xt(n_xt_), SB(SB_), sb(0) {
        start_sb(cta, grid);
        skip_empty(cta, grid);
    }
    __device__ __forceinline__ void skip_empty(int cta, int grid) {
        while (j >= end && (sb + 1) * SB < n_xt) { sb++; start_sb(cta, grid); }
    }
    __device__ __forceinline__ bool done() const { return j >= end; }
    __device__ __forceinline__ int qt() const { return (int)(j / nx); }
    __device__ __forceinline__ int xt() const { return sb * SB + (int)(j % nx); }
    __device__ __forceinline__ void advance(int cta, int grid) { j++; skip_empty(cta, grid); }
};

constexpr int QS_MAX_KBLOCKS = 8; /* query tile ≤ 128 × 256 fp32 = 128 KB */
constexpr int QS_MAX_STAGES = 8;
__host__ __device__ inline int qs_stages(int kblocks, int rows_per_stage) {
    const int left = 196 * 1024 - kblocks * BM * BK * 4;
    int n = left / (rows_per_stage * BK * 4);
    return n > QS_MAX_STAGES ? QS_MAX_STAGES : n;
}
__host__ __device__ inline size_t qs_smem_bytes(int kblocks, int BN, int CG) {
    return 1024 + (size_t)kblocks * BM * BK * 4 + (size_t)qs_stages(kblocks, BN / CG) * (BN / CG) * BK * 4 + 2 * BN * 4 + (2 * QS_MAX_STAGES + 6) * 8 + 16 + 256 * SBUF * 4;
}

/* The epilogue was the bound of the one-product filter (ncu, profiles/r2_ncu/exact_*: the compare + mask-building
 * lines were 46 % of all instructions, one epilogue warp per scheduler issuing at 0.37 IPC, tensor pipe 51 % busy), so
 * here it gets EIGHT warps — two per TMEM lane quarter, each taking half of the tile's columns — and a cheaper inner
 * loop: per 32 columns one subtract and one predicate-chained compare per element decide "no survivor in this group"
 * (the common case by far); only a group with a survivor is walked again to extract it. */
constexpr int QS_EPI_WARPS = 8, QS_THREADS = 64 + 32 * QS_EPI_WARPS;

template <int METRIC, int BN, int CG>
__global__ void __launch_bounds__(QS_THREADS, 1)
exact_tc_filter_qs(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_x, const FilterParams p, const int SB) {
    /* CG = 2: this CTA and its cluster peer form a pair — 256 queries (128 each, resident) against the same 256 rows,
     * each CTA streaming only ITS half of the row tile: half the L2→shared traffic and twice the look-ahead per byte */
    constexpr int BNL = BN / CG;             /* rows of the row tile this CTA loads */
    constexpr int A_BYTES = BM * BK * 4, B_BYTES = BNL * BK * 4;
    constexpr int EPT = 32 * QS_EPI_WARPS;   /* epilogue threads */
    constexpr int HALF = BN / 2;             /* columns per epilogue thread */
    extern __shared__ unsigned char smem_dyn[];
    unsigned char *base = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    const int NST = qs_stages(p.kblocks, BNL);
    unsigned char *qsm = base;                                   /* [kblocks][128 × 32 fp32], 128B-swizzled */
    unsigned char *stages = qsm + (size_t)p.kblocks * A_BYTES;   /* [NST][BN × 32 fp32] */
    float *xaux = reinterpret_cast<float *>(stages + (size_t)NST * B_BYTES);
    uint64_t *full = reinterpret_cast<uint64_t *>(xaux + 2 * BN);
    uint64_t *empty = full + QS_MAX_STAGES;
    uint64_t *tmem_full = empty + QS_MAX_STAGES;
    uint64_t *tmem_empty = tmem_full + 2;
    uint64_t *q_full = tmem_empty + 2;
    uint64_t *q_empty = q_full + 1;
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(q_empty + 1);
    int *sbuf = reinterpret_cast<int *>(tmem_ptr + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_qt = (p.nq + BM * CG - 1) / (BM * CG), n_xt = (p.rows + BN - 1) / BN;
    const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;      /* 0 = the pair's leader: it alone issues the MMAs */
    const int cta = CG == 2 ? (int)cluster_id_x() : (int)blockIdx.x, grid = CG == 2 ? (int)cluster_count_x() : (int)gridDim.x;
    const int QROWS = BM * CG;                                  /* queries per (pair) tile */

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_q); prefetch_tmap(&tm_x);
        for (int s = 0; s < NST; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int a = 0; a < 2; a++) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], QS_EPI_WARPS * CG); }
        mbar_init(q_full, 1); mbar_init(q_empty, 1);
        fence_barrier_init();
    }
    if (warp == 1) { if (CG == 2) tmem_alloc_pair(tmem_ptr, 2 * BN); else tmem_alloc(tmem_ptr, 2 * BN); }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all(); /* the peer's barriers are initialised before anything signals them */
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) { /* ── TMA producer ── */
            int s = 0; uint32_t ph = 0, qe_ph = 0; int qt_have = -1;
            for (TileWalk w(n_qt, n_xt, SB, cta, grid); !w.done(); w.advance(cta, grid)) {
                const int qt = w.qt(), xt = w.xt();
                if (qt != qt_have) {
                    if (qt_have >= 0) { mbar_wait(q_empty, qe_ph); qe_ph ^= 1; } /* every MMA on the old query tile has retired */
                    if (rank == 0) mbar_expect_tx(q_full, (uint32_t)(p.kblocks * A_BYTES * CG)); /* the pair's bytes land on the leader's barrier */
                    for (int kb = 0; kb < p.kblocks; kb++) {
                        if (CG == 2) tma_load_2d_pair(qsm + (size_t)kb * A_BYTES, &tm_q, map_to_cta(q_full, 0), kb * BK, qt * QROWS + (int)rank * BM);
                        else tma_load_2d(qsm + (size_t)kb * A_BYTES, &tm_q, q_full, kb * BK, qt * BM);
                    }
                    qt_have = qt;
                }
                for (int kb = 0; kb < p.kblocks; kb++) {
                    mbar_wait(&empty[s], ph ^ 1);
                    if (rank == 0) mbar_expect_tx(&full[s], B_BYTES * CG);
                    if (CG == 2) tma_load_2d_pair(stages + (size_t)s * B_BYTES, &tm_x, map_to_cta(&full[s], 0), kb * BK, xt * BN + (int)rank * BNL);
                    else tma_load_2d(stages + (size_t)s * B_BYTES, &tm_x, &full[s], kb * BK, xt * BN);
                    if (++s == NST) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) { /* ── MMA issuer (the pair's leader) ── */
            constexpr uint32_t idesc = make_idesc_tf32(BN, BM * CG);
            int s = 0; uint32_t ph = 0, qf_ph = 0; int a = 0; uint32_t aph = 0; int qt_have = -1;
            TileWalk w(n_qt, n_xt, SB, cta, grid);
            while (!w.done()) {
                const int qt = w.qt();
                if (qt != qt_have) { mbar_wait(q_full, qf_ph); qf_ph ^= 1; qt_have = qt; tc_fence_after(); }
                mbar_wait(&tmem_empty[a], aph ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(a * BN);
                for (int kb = 0; kb < p.kblocks; kb++) {
                    mbar_wait(&full[s], ph);
                    tc_fence_after();
                    const uint64_t a_d = make_desc_k128(smem_u32(qsm + (size_t)kb * A_BYTES));
                    const uint64_t b_d = make_desc_k128(smem_u32(stages + (size_t)s * B_BYTES));
#pragma unroll
                    for (int k4 = 0; k4 < BK / UK; k4++) {
                        if (CG == 2) mma_tf32_ss_pair(d_tmem, a_d + 2 * k4, b_d + 2 * k4, idesc, (kb | k4) ? 1u : 0u);
                        else mma_tf32_ss(d_tmem, a_d + 2 * k4, b_d + 2 * k4, idesc, (kb | k4) ? 1u : 0u);
                    }
                    if (CG == 2) tc_commit_pair(&empty[s]); else tc_commit(&empty[s]);
                    if (++s == NST) { s = 0; ph ^= 1; }
                }
                if (CG == 2) tc_commit_pair(&tmem_full[a]); else tc_commit(&tmem_full[a]);
                w.advance(cta, grid);
                if (w.done() || w.qt() != qt_have) { if (CG == 2) tc_commit_pair(q_empty); else tc_commit(q_empty); } /* the query tile may be replaced once these MMAs retire */
                if (++a == 2) { a = 0; aph ^= 1; }
            }
        }
    } else {
        /* ── epilogue: 8 warps; warp w reads TMEM lanes 32·(w mod 4)…, thread = one query × one half of the columns ──
         * (the compare is the one of exact_tc_filter: prune iff s < [h·qn − τ/2] + [h·xn] etc., NaN keeps the row) */
        const int ew = warp - 2;                       /* 0..7 */
        const int quarter = warp & 3, half = ew >> 2;  /* TMEM lane quarter (hardware: warp id mod 4), column half */
        const int et = ew * 32 + lane;                 /* 0..255: also the column this thread stages */
        const int qrow = quarter * 32 + lane;
        const float c = p.c_eps, hh = 0.5f * (1.0f - c);
        int a = 0; uint32_t aph = 0;
        static_assert(BN == EPT, "one staged column per epilogue thread");
        float xn_pre = 0.0f;
        auto fetch_xn = [&](bool have, int xt2) {
            xn_pre = 0.0f;
            if (have && xt2 * BN + et < p.rows) xn_pre = __ldg(p.xnorm2 + p.row0 + (long long)xt2 * BN + et);
        };
        TileWalk w(n_qt, n_xt, SB, cta, grid);
        fetch_xn(!w.done(), w.done() ? 0 : w.xt());
        int qt_have = -1;
        float qa = 0.0f;
        bool qvalid = false;
        int q = 0;
        int *mybuf = sbuf + et * SBUF;
        int nbuf = 0;
        auto flush = [&]() {
            if (nbuf == 0) return;
            const int pos = atomicAdd(p.cand_count + q, nbuf);
            for (int i = 0; i < nbuf; i++)
                if (pos + i < p.cand_cap) p.cand[(size_t)q * p.cand_cap + pos + i] = mybuf[i];
            nbuf = 0;
        };
        while (!w.done()) {
            const int qt = w.qt(), xt = w.xt();
            const int ncols = p.rows - xt * BN < BN ? p.rows - xt * BN : BN;
            const long long xrow0 = p.row0 + (long long)xt * BN;
            if (qt != qt_have) {
                flush();
                qt_have = qt;
                q = qt * QROWS + (int)rank * BM + qrow;
                qvalid = q < p.nq;
                qa = 0.0f;
                if (qvalid) {
                    const float qn = p.qaux[q], tq = __ldcg(p.tau + q) + 1e-30f;
                    if (METRIC == METRIC_L2) qa = fmaf(hh, qn, -0.5f * tq);
                    else if (METRIC == METRIC_IP) qa = fmaf(-c, qn, -tq);
                    else qa = qn < 1e-18f ? __int_as_float(0x7fc00000) : (1.0f - tq - c) * sqrtf(qn);
                }
            }
            float *xa = xaux + a * BN;
            {
                float v = 0.0f;
                if (et < ncols) {
                    const float xn = xn_pre;
                    if (METRIC == METRIC_L2) v = hh * xn;
                    else if (METRIC == METRIC_IP) v = -c * xn;
                    else v = xn < 1e-18f ? __int_as_float(0x7fc00000) : rsqrtf(xn);
                }
                xa[et] = v;
            }
            w.advance(cta, grid);
            fetch_xn(!w.done(), w.done() ? 0 : w.xt()); /* the next tile's row terms, in flight during this tile's columns */
            asm volatile("bar.sync 1, %0;" ::"n"(EPT) : "memory");
            mbar_wait(&tmem_full[a], aph);
            tc_fence_after();
            const int cbeg = half * HALF, cend = ncols < cbeg + HALF ? ncols : cbeg + HALF; /* this thread's columns */
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(a * BN);
            auto consume = [&](const uint32_t (&v)[32], int c0) {
                /* fast path: "every one of these 32 columns is pruned" as one predicate chain */
                bool all_pruned = true;
#pragma unroll
                for (int j4 = 0; j4 < 32; j4 += 4) {
                    const float4 x4 = *reinterpret_cast<const float4 *>(xa + c0 + j4);
                    const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
                    for (int jj = 0; jj < 4; jj++) {
                        const float sdot = __uint_as_float(v[j4 + jj]);
                        all_pruned = all_pruned && (METRIC == METRIC_COSINE ? (sdot * xs[jj] < qa) : (sdot - xs[jj] < qa));
                    }
                }
                if (all_pruned || !qvalid) return;
                uint32_t keep = 0;
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const float sdot = __uint_as_float(v[j]), x = xa[c0 + j];
                    const bool prune = METRIC == METRIC_COSINE ? (sdot * x < qa) : (sdot - x < qa);
                    keep |= prune ? 0u : (1u << j);
                }
                if (cend - c0 < 32) keep &= (1u << (cend - c0)) - 1u;
                while (keep) {
                    const int j = __ffs(keep) - 1;
                    keep &= keep - 1;
                    mybuf[nbuf++] = (int)(xrow0 + c0 + j);
                    if (nbuf == SBUF) flush();
                }
            };
            if (cbeg < cend) {
                uint32_t va[32], vb[32];
                tmem_ld32(taddr + (uint32_t)cbeg, va);
                for (int c0 = cbeg; c0 < cend; c0 += 64) {
                    tmem_ld_wait();
                    const bool more = c0 + 32 < cend;
                    if (more) tmem_ld32(taddr + (uint32_t)(c0 + 32), vb);
                    consume(va, c0);
                    if (more) {
                        tmem_ld_wait();
                        if (c0 + 64 < cend) tmem_ld32(taddr + (uint32_t)(c0 + 64), va);
                        consume(vb, c0 + 32);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG == 2) mbar_arrive_cluster(map_to_cta(&tmem_empty[a], 0)); else mbar_arrive(&tmem_empty[a]); }
            if (++a == 2) { a = 0; aph ^= 1; }
        }
        flush();
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all(); /* the leader's MMAs read the peer's shared memory until the last tile has retired */
    if (warp == 1) { tc_fence_after(); if (CG == 2) tmem_dealloc_pair(tmem_base, 2 * BN); else tmem_dealloc(tmem_base, 2 * BN); }
}

/* ───────────── split / prepare ───────────── */

__device__ __forceinline__ void split_tf32(float x, float &hi, float &lo) {
    hi = __uint_as_float(__float_as_uint(x) & TF32_MASK);
    lo = __uint_as_float(__float_as_uint(__fsub_rn(x, hi)) & TF32_MASK); /* x − hi is exact */
}

/* rows [0, rows) of src (row stride ld_src floats, `dim` valid) → hi / lo slabs [rows × kp], zero padded */
__global__ void tc_split_rows(const float *__restrict__ src, int ld_src, int dim, long long rows, int kp,
                              float *__restrict__ hi, float *__restrict__ lo) {
    const long long total = rows * (long long)(kp >> 2);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / (kp >> 2);
        const int c = (int)(i % (kp >> 2)) * 4;
        float v[4], h[4], l[4];
        if ((ld_src & 3) == 0 && c + 4 <= ld_src) {
            const float4 t = *reinterpret_cast<const float4 *>(src + r * ld_src + c);
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
#pragma unroll
            for (int j = 0; j < 4; j++) if (c + j >= dim) v[j] = 0.0f;
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) v[j] = c + j < dim ? src[r * ld_src + c + j] : 0.0f;
        }
#pragma unroll
        for (int j = 0; j < 4; j++) split_tf32(v[j], h[j], l[j]);
        *reinterpret_cast<float4 *>(hi + r * kp + c) = make_float4(h[0], h[1], h[2], h[3]);
        if (lo) *reinterpret_cast<float4 *>(lo + r * kp + c) = make_float4(l[0], l[1], l[2], l[3]);
    }
}

/* per query: Σq² in chain order → qaux; τ = +inf; empty candidate list, empty top-k, no overflow */
template <int METRIC>
__global__ void tc_prep_queries(const float *__restrict__ queries, int dim, int nq, float *__restrict__ qaux,
                                float *__restrict__ tau, int *__restrict__ cand_count, int *__restrict__ topk_size,
                                int *__restrict__ overflow) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const float n2 = chained_sum_serial<METRIC_IP>(queries + (size_t)q * dim, queries + (size_t)q * dim, dim);
    qaux[q] = n2;
    tau[q] = __int_as_float(0x7f800000);
    cand_count[q] = 0; topk_size[q] = 0; overflow[q] = 0;
}

/* τ of every query from its current sorted top-k list (after the dense seed over the first rows, exact_knn_tc) */
__global__ void tc_seed_tau(const unsigned long long *__restrict__ topk, const int *__restrict__ topk_size, int nq, int k,
                            float *__restrict__ tau) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < nq && topk_size[q] >= k) tau[q] = okey_inv((unsigned)(topk[(size_t)q * k + k - 1] >> 32));
}

/* ───────────── re-rank ───────────── */

/* chained sum of the reference's loop for one (query, row) pair, computed by 4 adjacent lanes:
 * lane h owns SSE lane h (elements i ≡ h mod 4, ascending), then ((s0+s1)+s2)+s3 and the scalar
 * tail (src/vec_math.c:82-94). All 4 lanes return the sum. */
template <int METRIC>
__device__ __forceinline__ float quad_chain(const float *__restrict__ q, const float *__restrict__ x, int dim, int lane, bool have) {
    const int h = lane & 3, g = lane & ~3, cfull = dim >> 2;
    float s = 0.0f;
    if (have)
        for (int c = 0; c < cfull; c++) {
            const float a = q[4 * c + h], b = x[4 * c + h];
            float pr;
            if (METRIC == METRIC_L2) { const float d = __fsub_rn(a, b); pr = __fmul_rn(d, d); } else pr = __fmul_rn(a, b);
            s = __fadd_rn(s, pr);
        }
    const float c0 = __shfl_sync(FULL, s, g), c1 = __shfl_sync(FULL, s, g + 1), c2 = __shfl_sync(FULL, s, g + 2),
                c3 = __shfl_sync(FULL, s, g + 3);
    float sum = __fadd_rn(__fadd_rn(__fadd_rn(c0, c1), c2), c3);
    if (have)
        for (int i = cfull * 4; i < dim; i++) {
            const float a = q[i], b = x[i];
            float pr;
            if (METRIC == METRIC_L2) { const float d = __fsub_rn(a, b); pr = __fmul_rn(d, d); } else pr = __fmul_rn(a, b);
            sum = __fadd_rn(sum, pr);
        }
    return sum;
}

/* the same for RU rows at once per lane quad: RU independent chains per lane, RU rows' loads in flight
 * (the survivors are random rows of the slab: latency, not arithmetic, is what this kernel waits on) */
template <int METRIC, int RU>
__device__ __forceinline__ void quad_chain_multi(const float *__restrict__ q, const float *const (&x)[RU], const bool (&have)[RU],
                                                 int dim, int lane, float (&sum)[RU]) {
    const int h = lane & 3, g = lane & ~3, cfull = dim >> 2;
    float s[RU];
#pragma unroll
    for (int u = 0; u < RU; u++) {
        s[u] = 0.0f;
        /* every 128-byte line of the row towards L2 now: the chain loop below would otherwise discover them one
         * DRAM round trip after the other */
        if (have[u])
            for (int off = 32 * h; off < dim; off += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(x[u] + off));
    }
    for (int c = 0; c < cfull; c++) {
        const float a = q[4 * c + h];
#pragma unroll
        for (int u = 0; u < RU; u++) {
            const float b = have[u] ? x[u][4 * c + h] : 0.0f;
            float pr;
            if (METRIC == METRIC_L2) { const float d = __fsub_rn(a, b); pr = __fmul_rn(d, d); } else pr = __fmul_rn(a, b);
            s[u] = __fadd_rn(s[u], pr);
        }
    }
#pragma unroll
    for (int u = 0; u < RU; u++) {
        const float c0 = __shfl_sync(FULL, s[u], g), c1 = __shfl_sync(FULL, s[u], g + 1), c2 = __shfl_sync(FULL, s[u], g + 2),
                    c3 = __shfl_sync(FULL, s[u], g + 3);
        float t = __fadd_rn(__fadd_rn(__fadd_rn(c0, c1), c2), c3);
        if (have[u])
            for (int i = cfull * 4; i < dim; i++) {
                const float a = q[i], b = x[u][i];
                float pr;
                if (METRIC == METRIC_L2) { const float d = __fsub_rn(a, b); pr = __fmul_rn(d, d); } else pr = __fmul_rn(a, b);
                t = __fadd_rn(t, pr);
            }
        sum[u] = t;
    }
}

/* one warp per query: exact distances of the chunk's survivors → sorted top-k → τ */
template <int METRIC>
__global__ void __launch_bounds__(128)
exact_tc_rerank(const float *__restrict__ vec, const float *__restrict__ norm2, const uint8_t *__restrict__ deleted, int ld,
                int dim, const float *__restrict__ queries, int nq, int k, const int *__restrict__ cand,
                int *__restrict__ cand_count, int cap, unsigned long long *__restrict__ topk, int *__restrict__ topk_size,
                float *__restrict__ tau, int *__restrict__ overflow, unsigned long long *__restrict__ n_cand_total) {
    constexpr int RU = 4; /* rows per lane quad per round: 32 survivors in flight per warp */
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int q = blockIdx.x * 4 + wib;
    unsigned long long *R = reinterpret_cast<unsigned long long *>(smem_raw) + (size_t)wib * (k + 32);
    unsigned long long *S = R + k;
    if (q >= nq) return;
    int cnt = cand_count[q];
    if (cnt > cap) { if (lane == 0) overflow[q] = 1; cnt = cap; }
    int size = topk_size[q];
    if (cnt > 0) {
        for (int i = lane; i < size; i += 32) R[i] = topk[(size_t)q * k + i];
        __syncwarp();
        const float *qv = queries + (size_t)q * dim;
        float qn2 = 0.0f;
        if (METRIC == METRIC_COSINE) qn2 = quad_chain<METRIC_IP>(qv, qv, dim, lane, true);
        int ev = 0, mp;
        for (int c0 = 0; c0 < cnt; c0 += 8 * RU) {
            const float *xr[RU];
            bool have[RU];
            int row[RU];
            float sum[RU];
#pragma unroll
            for (int u = 0; u < RU; u++) {
                const int ci = c0 + u * 8 + (lane >> 2);
                have[u] = ci < cnt;
                row[u] = have[u] ? cand[(size_t)q * cap + ci] : 0;
                xr[u] = vec + (size_t)row[u] * ld;
            }
            quad_chain_multi<METRIC, RU>(qv, xr, have, dim, lane, sum);
            /* the round's ≤ 32 keys that beat the current k-th go into the list with ONE merge (one per group of 8 rows
             * cost up to four passes over the list per round: 5.6 of 36.8 ms at k = 100 on the config-5 shard) */
            const unsigned long long worst = size >= k ? R[k - 1] : KEY_NONE;
            int ns = 0;
#pragma unroll
            for (int u = 0; u < RU; u++) {
                if (c0 + u * 8 >= cnt) break; /* warp-uniform */
                float d;
                if (METRIC == METRIC_L2) d = sum[u];
                else if (METRIC == METRIC_IP) d = -sum[u];
                else d = cosine_finish(sum[u], qn2, have[u] ? norm2[row[u]] : 1.0f);
                const bool live = have[u] && (lane & 3) == 0 && !(deleted && deleted[row[u]]);
                const unsigned long long f = live ? make_key(d, (unsigned)row[u]) : KEY_NONE;
                const bool keep = f < worst;
                const unsigned km = __ballot_sync(FULL, keep);
                if (keep) S[ns + __popc(km & ((1u << lane) - 1))] = f;
                ns += __popc(km);
            }
            if (ns > 0) {
                __syncwarp();
                size = warp_merge_sorted(R, size, k, S, ns, lane, &mp, &ev);
            }
        }
        __syncwarp();
        for (int i = lane; i < size; i += 32) topk[(size_t)q * k + i] = R[i];
        if (lane == 0) {
            topk_size[q] = size;
            if (size >= k) tau[q] = okey_inv((unsigned)(R[k - 1] >> 32));
            cand_count[q] = 0;
            atomicAdd(n_cand_total, (unsigned long long)cnt);
        }
    }
}

/* Small batches (the HBM-bound regime, nq < 1024): ONE CTA of 16 warps per query instead of one warp. The
 * survivors' exact distances are computed by all warps at once (256 rows in flight per query), the keys that
 * beat the current k-th best are appended to the query's sorted list in shared memory and the union is put in
 * order: a radix select finds the k-th smallest distance, the keys up to it are sorted bitonically — the k smallest
 * keys of R ∪ survivors, the same set in the same (distance, row) order as the warp kernel's sequence of merges.
 * smem: A[P_alloc] keys | query | 16 warps × 16-row tiles (re-used for the compacted keys) ; P_alloc = pow2 ≥ k + cap. */
constexpr int RR_WARPS = 16, RR_THREADS = RR_WARPS * 32; /* warps of the CTA-per-query re-rank */
template <int METRIC>
__global__ void __launch_bounds__(RR_THREADS)
exact_tc_rerank_cta(const float *__restrict__ vec, const float *__restrict__ norm2, const uint8_t *__restrict__ deleted, int ld,
                    int dim, const float *__restrict__ queries, int nq, int k, const int *__restrict__ cand,
                    int *__restrict__ cand_count, int cap, unsigned long long *__restrict__ topk, int *__restrict__ topk_size,
                    float *__restrict__ tau, int *__restrict__ overflow, unsigned long long *__restrict__ n_cand_total, int P_alloc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned long long *A = reinterpret_cast<unsigned long long *>(smem_raw);
    __shared__ int s_fill, s_hist[256], s_sel[2];
    constexpr int NT = RR_THREADS, NW = RR_WARPS;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, q = blockIdx.x;
    if (q >= nq) return;
    int cnt = cand_count[q];
    if (cnt > cap) { if (threadIdx.x == 0) overflow[q] = 1; cnt = cap; }
    if (cnt <= 0) return; /* block-uniform */
    const int size = topk_size[q];
    for (int i = threadIdx.x; i < size; i += NT) A[i] = topk[(size_t)q * k + i];
    if (threadIdx.x == 0) s_fill = size;
    __syncthreads();
    const unsigned long long worst = size >= k ? A[k - 1] : KEY_NONE;
    /* distances: the search kernel's row path — 16 survivors per warp per round, each row one TMA bulk copy into
     * the warp's shared-memory tile, two lanes per row walking the reference's chains (hnsw_search.cuh) */
    SearchParams sp;
    sp.vec = vec; sp.norm2 = norm2; sp.ld = ld; sp.dim = dim;
    TeamSmem sm;
    unsigned char *tiles;
    {
        unsigned char *b = smem_raw + (size_t)P_alloc * 8;
        sm.q = reinterpret_cast<float *>(b); b += ((size_t)ld * 4 + 15) / 16 * 16;
        tiles = b;
        sm.P = reinterpret_cast<float *>(b) + (size_t)wib * 16 * PSTRIDE; b += (size_t)NW * 16 * PSTRIDE * 4;
        sm.Fs = reinterpret_cast<int *>(b) + wib * 16; b += NW * 16 * 4;
        sm.bar = reinterpret_cast<unsigned long long *>(b) + wib;
        sm.phase = 0u;
        sm.R = sm.Fk = sm.S = nullptr;
    }
    for (int i = threadIdx.x; i < ld; i += NT) sm.q[i] = i < dim ? queries[(size_t)q * dim + i] : 0.0f;
    if (lane == 0) bar_init(sm.bar, 1);
    __syncthreads();
    float qn2 = 0.0f;
    if (METRIC == METRIC_COSINE) qn2 = query_norm2(sm.q, dim);
    const unsigned long long pol = l2_evict_first_policy();
    for (int c0 = wib * 16; c0 < cnt; c0 += NW * 16) {
        const int nb = cnt - c0 < 16 ? cnt - c0 : 16;
        if (lane < nb) sm.Fs[lane] = cand[(size_t)q * cap + c0 + lane];
        __syncwarp();
        const float d = warp_dist_rows<METRIC, 2, 0>(sp, sm, sm.Fs, nb, qn2, lane, pol);
        const int r = lane >> 1;
        const int row = r < nb ? sm.Fs[r] : 0;
        const bool live = (lane & 1) == 0 && r < nb && !(deleted && deleted[row]);
        const unsigned long long f = live ? make_key(d, (unsigned)row) : KEY_NONE;
        const bool keep = f < worst;
        const unsigned km = __ballot_sync(FULL, keep);
        if (km) {
            int base = 0;
            if (lane == 0) base = atomicAdd(&s_fill, __popc(km));
            base = __shfl_sync(FULL, base, 0);
            if (keep) A[base + __popc(km & ((1u << lane) - 1))] = f;
        }
        __syncwarp();
    }
    __syncthreads();
    const int total = s_fill;
    unsigned long long *srt = A;
    if (total > size) { /* something entered: the k smallest of the union, in order */
        int m = total;
        if (total > 512) {
            /* radix select on the distance half of the keys: thr = the k-th smallest distance (k ≤ 512 < total), then
             * only the keys with distance ≤ thr (k of them, plus ties of the k-th) are sorted */
            unsigned prefix = 0;
            int want = k;
            for (int pass = 3; pass >= 0; pass--) {
                const int shift = 8 * pass;
                if (threadIdx.x < 256) s_hist[threadIdx.x] = 0;
                __syncthreads();
                for (int i0 = 0; i0 < total; i0 += NT) {
                    const int i = i0 + threadIdx.x;
                    unsigned hi = 0;
                    bool in = i < total;
                    if (in) {
                        hi = (unsigned)(A[i] >> 32);
                        in = pass == 3 || (hi >> (shift + 8)) == (prefix >> (shift + 8));
                    }
                    const unsigned pm = __ballot_sync(FULL, in);
                    if (in) {
                        const unsigned bin = (hi >> shift) & 255u;
                        const unsigned peers = __match_any_sync(pm, bin);
                        if ((__ffs(peers) - 1) == lane) atomicAdd(&s_hist[bin], __popc(peers));
                    }
                }
                __syncthreads();
                if (wib == 0) {
                    int part = 0;
#pragma unroll
                    for (int b8 = 0; b8 < 8; b8++) part += s_hist[8 * lane + b8];
                    int incl = part;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const int up = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += up; }
                    const unsigned hit = __ballot_sync(FULL, incl >= want);
                    const int L = __ffs(hit) - 1; /* total ≥ want: some lane hits */
                    if (lane == L) {
                        int before = incl - part, b8 = 0;
                        while (before + s_hist[8 * lane + b8] < want) { before += s_hist[8 * lane + b8]; b8++; }
                        s_sel[0] = (int)(prefix | ((unsigned)(8 * lane + b8) << shift));
                        s_sel[1] = want - before;
                    }
                }
                __syncthreads();
                prefix = (unsigned)s_sel[0];
                want = s_sel[1];
            }
            if (threadIdx.x == 0) s_fill = 0;
            __syncthreads();
            /* the row tiles are idle now (every warp has waited for its last copy): they hold the compacted keys */
            unsigned long long *B = reinterpret_cast<unsigned long long *>(tiles);
            for (int i = threadIdx.x; i < total; i += NT) {
                const unsigned long long e = A[i];
                if ((unsigned)(e >> 32) <= prefix) B[atomicAdd(&s_fill, 1)] = e;
            }
            __syncthreads();
            m = s_fill;
            srt = B;
        }
        int P = 64;
        while (P < m) P <<= 1;
        for (int i = m + threadIdx.x; i < P; i += NT) srt[i] = KEY_NONE;
        __syncthreads();
        for (int kk = 2; kk <= P; kk <<= 1)
            for (int j = kk >> 1; j > 0; j >>= 1) {
                for (int i = threadIdx.x; i < P; i += NT) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
                        const unsigned long long x = srt[i], y = srt[ixj];
                        const bool up = (i & kk) == 0;
                        if ((x > y) == up) { srt[i] = y; srt[ixj] = x; }
                    }
                }
                __syncthreads();
            }
    }
    const int nsize = total < k ? total : k;
    for (int i = threadIdx.x; i < nsize; i += NT) topk[(size_t)q * k + i] = srt[i];
    if (threadIdx.x == 0) {
        topk_size[q] = nsize;
        if (nsize >= k) tau[q] = okey_inv((unsigned)(srt[k - 1] >> 32));
        cand_count[q] = 0;
        atomicAdd(n_cand_total, (unsigned long long)cnt);
    }
}

/* flagged queries → dense list (order irrelevant) */
__global__ void tc_compact_overflow(const int *__restrict__ overflow, int nq, int *__restrict__ qmap, int *__restrict__ n_flag) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < nq && overflow[q]) qmap[atomicAdd(n_flag, 1)] = q;
}
__global__ void tc_gather_queries(const float *__restrict__ queries, int dim, const int *__restrict__ qmap, int n,
                                  float *__restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)n * dim) out[i] = queries[(size_t)qmap[i / dim] * dim + i % dim];
}

} // namespace tc
} // namespace muninn
