/*
 * hnsw_search.cuh — batched HNSW search kernel for sm_100a.
 *
 * Restates, per query, the reference's hnsw_search (src/hnsw_algo.c:670-704):
 *   greedy_search_layer  src/hnsw_algo.c:257-282   list-order hill climb with the in-loop switch
 *   beam_search_layer    src/hnsw_algo.c:347-448   ef-bounded best-first search + patience rule
 *   distances            src/vec_math.c:78-143     bit-exact summation order
 *
 * Execution model (B200): a persistent grid of query teams. A team is one warp (one CTA) that
 * pulls query indices from a global counter until the batch is drained, so long and short
 * searches balance across the 148 SMs; 24 teams share an SM at dim ≤ 128. Per team, in shared
 * memory: the query, the result list R (ef entries, ascending (distance, slot) 64-bit keys,
 * bit 0 = "expanded"), the fresh-neighbour lists and a row staging tile (13 rows × 528 B for
 * rows of ≤ 512 B, 16 rows for dim 96 and for wide rows). Per team, in HBM (L2-resident in
 * practice): a bucketised visited table of 32-bit slot entries, four per 16-byte bucket, cleared
 * by the team at the start of each query.
 *
 * Distances: every fresh neighbour row of a batch is brought into the tile by ONE TMA bulk copy
 * (cp.async.bulk, lane r issues row r, completion on a per-team mbarrier, L2 evict-first hint);
 * lane (row, h) then walks its row in shared memory and accumulates chains 2h, 2h+1 of the
 * reference's SSE loop in ascending element order as a packed f32x2 pair (sub / add packed,
 * multiplies scalar so that nothing is contracted into an FMA), followed by the fixed
 * ((s0+s1)+s2)+s3 horizontal sum and the scalar tail — bit-identical to src/vec_math.c:78-143.
 *
 * Result-set semantics ("canonical order"): R' = the ef smallest of R ∪ F under
 * (distance, slot); next candidate = smallest unexpanded entry of R. On inputs without
 * bit-equal distances this is exactly the reference's two-heap loop (SURVEY.md finding 4);
 * with ties the reference's order is heap-dependent and unspecified (test_priority_queue.c:85-99).
 */
#pragma once
#include "muninn_dev.cuh"
#include "../../include/muninn_cuda.h"

namespace muninn {

#ifndef MUNINN_SEARCH_MIN_BLOCKS
#define MUNINN_SEARCH_MIN_BLOCKS 24 /* resident query teams per SM the register budget is sized for */
#endif
#ifndef MUNINN_RB
#define MUNINN_RB 0 /* 0 = by row width: RB_NARROW rows (24 teams/SM) for one-column-block rows, else 16 */
#endif
constexpr int DB = 8;             /* rows per distance batch in the greedy descent (4 lanes per row) */
/* Rows per beam batch. Rows of ≤ 512 bytes are staged in one go: 13-row batches shrink the tile so
 * that 24 teams fit an SM (measured +5 % over 16 rows / 20 teams, +3 % over 12 rows). Wider rows take several column
 * blocks per batch, each a memory round trip: keep the batch as large as the lanes allow. */
#ifndef MUNINN_RB_NARROW
#define MUNINN_RB_NARROW 13 /* batch rows for rows of ≤ 512 bytes: the largest tile that still lets 24 teams share an SM */
#endif
constexpr int RB_NARROW = MUNINN_RB_NARROW, RB_WIDE = 16;
__host__ __device__ inline int beam_rows(int ld) { return MUNINN_RB ? (MUNINN_RB < 16 ? RB_NARROW : RB_WIDE) : (ld <= 128 ? RB_NARROW : RB_WIDE); }
constexpr unsigned VIS_EMPTY = 0xffffffffu;
constexpr int PSTRIDE = 128 + 4;  /* floats per staged row: 32 chunks + 4 pad → conflict-free per-row reads */
constexpr int MAX_M0 = 128;       /* layer-0 list capacity supported (M ≤ 64) */
/* dim 96 (config 4): rows are 384 bytes, so a 16-row tile at stride 100 floats (≡ 4 mod 32: conflict-free like 132)
 * is SMALLER than the 13-row tile at stride 132 — the full 16-row batch at the same 24 teams/SM */
#ifndef MUNINN_D96_WIDE
#define MUNINN_D96_WIDE 1
#endif
template <int CD> struct TileOf { static constexpr int stride = (MUNINN_D96_WIDE && CD == 96) ? 100 : PSTRIDE; };
constexpr int RB_D96 = MUNINN_D96_WIDE ? 16 : MUNINN_RB_NARROW;
__host__ __device__ inline bool tile_d96(int ld, int dim) { return MUNINN_D96_WIDE && ld == 96 && dim == 96; }
__host__ __device__ inline int tile_rows(int ld, int dim) { return tile_d96(ld, dim) ? RB_D96 : beam_rows(ld); }
__host__ __device__ inline int tile_stride(int ld, int dim) { return tile_d96(ld, dim) ? 100 : PSTRIDE; }

struct SearchParams {
    /* graph */
    const float *vec; const float *norm2; const int32_t *nbr0; const int32_t *upper_base; const int32_t *nbrU;
    const uint8_t *deleted; const int64_t *ids;
    int dim, ld, M, M0, entry_slot, max_level;
    long long n_live;
    /* batch */
    const float *queries; int nq, k, ef, patience_max;
    int64_t *out_ids; float *out_dist; int32_t *out_count; MuninnSearchStats *stats;
    /* work lists: level 0 processes queries 0..nq-1; level>0 processes in_list[0..*in_count) */
    const int *in_list; const int *in_count; int *out_list; int *out_count_retry; int *work_counter;
    /* visited tables: one per team, vis_cap 32-bit slots each (VIS_EMPTY = free), cleared per query */
    unsigned *vis; unsigned vis_shift; unsigned vis_cap;
    /* host-buffer entry, large batches: queries [0, *q_arrived) have landed in `queries` (written by the copy engine behind
     * each slice of the host→device copy); NULL = all of them. The kernel starts on the first slice while the rest streams in. */
    const unsigned *q_arrived;
    /* team kernels (hnsw_beam.cuh, W > 1), level 0: bytes of shared memory behind the team's other buffers that hold its
     * visited table (= vis_cap · 4), or 0 = the table lives in the global arena like everywhere else */
    int vis_smem;
};

struct TeamSmem {
    unsigned long long *R;   /* [ef] */
    unsigned long long *Fk;  /* [M0] keys of this expansion's fresh neighbours */
    unsigned long long *S;   /* [32] survivors being merged */
    float *q;                /* [ld] */
    float *P;                /* [beam_rows * PSTRIDE] row staging tile */
    int *Fs;                 /* [fs_len(M0)] fresh slots; the greedy descent keeps list positions at Fs[fs_pos(M0) + i] */
    unsigned long long *bar; /* mbarrier the bulk row copies complete on */
    unsigned phase;          /* its current parity */
};

__host__ __device__ inline int fs_pos(int M0) { return M0 > DB ? M0 : DB; }
__host__ __device__ inline int fs_len(int M0) { return fs_pos(M0) + DB; }

__host__ __device__ inline size_t search_smem_bytes(int ef, int ld, int M0, int dim) {
    size_t b = (((size_t)ef * 8 + 15) & ~(size_t)15) + (size_t)M0 * 8 + 32 * 8;
    b += ((size_t)ld * 4 + 15) / 16 * 16;
    b += (size_t)tile_rows(ld, dim) * tile_stride(ld, dim) * 4;
    b += ((size_t)fs_len(M0) * 4 + 15) & ~(size_t)15;
    b += 16; /* mbarrier */
    return b;
}

/* L2 policy for the gathered rows: streamed once and never reused by this SM */
__device__ __forceinline__ unsigned long long l2_evict_first_policy() {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

/* TMA bulk copy global → shared of `bytes` (multiple of 16), completion counted on an mbarrier */
__device__ __forceinline__ void bulk_row_copy(void *smem_dst, const void *gmem_src, unsigned bytes, unsigned long long *bar,
                                              unsigned long long policy) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst), b = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(d), "l"(gmem_src), "r"(bytes), "r"(b), "l"(policy) : "memory");
}
__device__ __forceinline__ void bar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_wait(unsigned long long *bar, unsigned parity) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    unsigned done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
    } while (!done);
}

/* packed pair of fp32 (sm_100 FADD2): both halves are IEEE round-to-nearest, i.e. bit-identical to two
 * scalar operations. The products stay scalar FMULs on purpose: ptxas contracts mul.rn.f32x2 +
 * add.rn.f32x2 into one FFMA2, which would not be the reference's two roundings. */
__device__ __forceinline__ unsigned long long f2_pack(float lo, float hi) {
    return ((unsigned long long)__float_as_uint(hi) << 32) | __float_as_uint(lo);
}
__device__ __forceinline__ float f2_lo(unsigned long long v) { return __uint_as_float((unsigned)v); }
__device__ __forceinline__ float f2_hi(unsigned long long v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ unsigned long long f2_sub(unsigned long long a, unsigned long long b) {
    unsigned long long d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ unsigned long long f2_add(unsigned long long a, unsigned long long b) {
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

template <int METRIC>
__device__ __forceinline__ float prod1(float a, float b) {
    if (METRIC == METRIC_L2) { float d = __fsub_rn(a, b); return __fmul_rn(d, d); }
    return __fmul_rn(a, b);
}

/*
 * Distances from the staged query to up to 32/L rows (L = lanes per row: 4 → ≤ 8 rows, 2 → ≤ 16).
 * bs[0..nb) are slots (shared memory). The warp copies the rows into the staging tile with
 * asynchronous 16-byte copies (lane c moves chunk c of every row: one fully coalesced row per
 * instruction, all rows in flight at once, no registers held), then lane (row, h) walks ITS row
 * in shared memory and accumulates chains h·4/L … of the reference's SSE loop in ascending
 * element order (src/vec_math.c:82-86), followed by the fixed ((s0+s1)+s2)+s3 horizontal sum and
 * the scalar tail (:88-94). Row stride PSTRIDE = 132 floats makes both mappings bank-conflict
 * free. Every lane of a row group returns that row's distance.
 */
template <int METRIC, int L, int CD>
__device__ __forceinline__ float warp_dist_rows(const SearchParams &p, TeamSmem &sm, const int *bs, int nb,
                                                float q_norm2, int lane, unsigned long long pol) {
    constexpr int CH = 4 / L;
    constexpr int PS = TileOf<CD>::stride;
    /* CD != 0: dim == ld == CD (a multiple of 4, no tail) is a compile-time constant, so every trip count below
     * is one too — the unrolled chain loop sheds its counter/branch overhead (−9 % kernel time at dim 128) */
    const int cpr = CD ? CD / 4 : p.ld >> 2, cfull = CD ? CD / 4 : p.dim >> 2, rem = CD ? 0 : p.dim & 3;
    const int row = lane / L, h = lane % L, g = lane - h;
    const bool mine = row < nb;
    float nb2 = 1.0f;
    if (METRIC == METRIC_COSINE && mine && h == 0) nb2 = p.norm2[bs[row]];
    float s0 = 0.0f, s1 = 0.0f, tl0 = 0.0f, tl1 = 0.0f, tl2 = 0.0f;
    unsigned long long s01 = 0ull; /* (s0, s1) packed, CH == 2 */
    for (int t0 = 0; t0 < cpr; t0 += 32) {
        const int tc = cpr - t0 < 32 ? cpr - t0 : 32;
        if (lane == 0) bar_expect_tx(sm.bar, (unsigned)(nb * tc * 16));
        __syncwarp();
        if (lane < nb) bulk_row_copy(sm.P + lane * PS, p.vec + (size_t)bs[lane] * (size_t)p.ld + 4 * t0, (unsigned)(tc * 16), sm.bar, pol);
        bar_wait(sm.bar, sm.phase);
        sm.phase ^= 1u;
        if (mine) {
            int tfull = cfull - t0;
            tfull = tfull > tc ? tc : tfull;
            const float *xr = sm.P + row * PS + CH * h;
            const float *qr = sm.q + 4 * t0 + CH * h;
#pragma unroll (CD ? (CD / 4 < 32 ? CD / 4 : 32) : 8)
            for (int cc = 0; cc < tfull; cc++) {
                if (CH == 2) {
                    const unsigned long long x = *reinterpret_cast<const unsigned long long *>(xr + 4 * cc);
                    const unsigned long long q = *reinterpret_cast<const unsigned long long *>(qr + 4 * cc);
                    unsigned long long pr;
                    if (METRIC == METRIC_L2) {
                        const unsigned long long d = f2_sub(q, x);
                        pr = f2_pack(__fmul_rn(f2_lo(d), f2_lo(d)), __fmul_rn(f2_hi(d), f2_hi(d)));
                    } else pr = f2_pack(__fmul_rn(f2_lo(q), f2_lo(x)), __fmul_rn(f2_hi(q), f2_hi(x)));
                    s01 = f2_add(s01, pr);
                } else {
                    s0 = __fadd_rn(s0, prod1<METRIC>(qr[4 * cc], xr[4 * cc]));
                }
            }
            if (rem && h == 0 && cfull >= t0 && cfull < t0 + 32) { /* the partial chunk: kept for after the horizontal sum */
                const float *xt = sm.P + row * PS + 4 * (cfull - t0), *qt = sm.q + 4 * cfull;
                tl0 = prod1<METRIC>(qt[0], xt[0]);
                if (rem > 1) tl1 = prod1<METRIC>(qt[1], xt[1]);
                if (rem > 2) tl2 = prod1<METRIC>(qt[2], xt[2]);
            }
        }
        __syncwarp(); /* the tile is overwritten by the next column block */
    }
    if (CH == 2) { s0 = f2_lo(s01); s1 = f2_hi(s01); }
    float sum;
    if (L == 4) {
        float c0 = __shfl_sync(FULL, s0, g), c1 = __shfl_sync(FULL, s0, g + 1), c2 = __shfl_sync(FULL, s0, g + 2),
              c3 = __shfl_sync(FULL, s0, g + 3);
        sum = __fadd_rn(__fadd_rn(__fadd_rn(c0, c1), c2), c3);
    } else {
        float c2 = __shfl_sync(FULL, s0, g + 1), c3 = __shfl_sync(FULL, s1, g + 1);
        sum = __fadd_rn(__fadd_rn(__fadd_rn(s0, s1), c2), c3); /* meaningful in lane g */
    }
    if (rem) {
        sum = __fadd_rn(sum, tl0);
        if (rem > 1) sum = __fadd_rn(sum, tl1);
        if (rem > 2) sum = __fadd_rn(sum, tl2);
    }
    float d;
    if (METRIC == METRIC_L2) d = sum;
    else if (METRIC == METRIC_IP) d = -sum;
    else d = cosine_finish(sum, q_norm2, nb2);
    return __shfl_sync(FULL, d, g);
}

/* wait until query qi has arrived (see SearchParams::q_arrived): one lane polls with a system-scope acquire load */
__device__ __forceinline__ void wait_query_arrived(const unsigned *flag, int qi) {
    unsigned v;
    for (;;) {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v > (unsigned)qi) break;
        __nanosleep(200);
    }
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

/* Pull the rows of all `nf` fresh neighbours towards L2 in one go (4 lanes per row, one
 * 128-byte line each per step) so that the 8-row distance batches that follow overlap their
 * DRAM latency instead of paying it batch after batch. */
__device__ __forceinline__ void prefetch_rows(const SearchParams &p, const int *slots, int nf, int lane, int from_byte = 0) {
    const int row_bytes = p.ld * 4;
    for (int r0 = 0; r0 < nf; r0 += 8) {
        int r = r0 + (lane >> 2);
        if (r < nf) {
            const char *base = reinterpret_cast<const char *>(p.vec + (size_t)slots[r] * (size_t)p.ld);
            for (int off = from_byte + (lane & 3) * 128; off < row_bytes; off += 512) prefetch_l2(base + off);
        }
    }
}

/* Σ q_i² in chain order: the `norm_a` accumulator of vec_cosine_distance (src/vec_math.c:104-118) */
__device__ __forceinline__ float query_norm2(const float *q, int dim) {
    return chained_sum_serial<METRIC_IP>(q, q, dim);
}

/* visited test-and-set for one neighbour per lane; returns 1 if the slot was not yet visited.
 * The table is private to the team and cleared at the start of every query: 32-bit entries, plain L2-level
 * loads/stores, intra-warp claim conflicts resolved with match.any.
 * Entries per bucket, measured at config 2 (kernel ms at ef 64 / 128 / 256): linear probing 1.625 / 3.23 / 7.07,
 * 8-entry 1.616 / 3.42 / 7.55 (fewer round trips but twice the loads and compares: a loss once the merge makes the
 * kernel issue-bound), 4-entry 1.532 / 3.06 / 6.74 — one 16-byte load, the best everywhere.
 * Bucketised table: a key hashes to a BUCKET of four 32-bit entries (half a 32-byte sector), fetched with
 * one 16-byte load in one L2 round trip. Entries fill a bucket from the front and are never removed, so "found / first
 * empty slot / full" are all decided from that one fetch; only a full bucket (rare below 3/4 load: a search ends
 * near 40 %) sends the lane on to the next one. With one entry per probe every call paid ≈ 3.5 dependent round trips
 * (32 lanes: some lane always collides); this pays ≈ 1. Lanes claiming in the same bucket take consecutive empty
 * slots in lane order (match.any), no atomics. */
__device__ __forceinline__ int visited_test_and_set(const SearchParams &p, unsigned *tab, int t, bool valid, int lane) {
    constexpr int VB = 4, VBL = 2; /* entries per bucket, log2 */
    const unsigned bmask = (p.vis_cap >> VBL) - 1;
    unsigned b = (((unsigned)t * 2654435761u) >> p.vis_shift) >> VBL;
    bool active = valid;
    int fresh = 0;
    while (__any_sync(FULL, active)) {
        bool claim = false;
        int first_empty = VB;
        if (active) {
            const uint4 *bp = reinterpret_cast<const uint4 *>(tab + ((size_t)b << VBL));
            unsigned e[VB];
            const uint4 e0 = __ldcg(bp);
            e[0] = e0.x; e[1] = e0.y; e[2] = e0.z; e[3] = e0.w;
            bool found = false;
#pragma unroll
            for (int i = VB - 1; i >= 0; i--) {
                found |= e[i] == (unsigned)t;
                if (e[i] == VIS_EMPTY) first_empty = i;
            }
            if (found) active = false;            /* already visited */
            else if (first_empty == VB) b = (b + 1) & bmask; /* full: the key, if present, lives further on */
            else claim = true;
        }
        const unsigned cm = __ballot_sync(FULL, claim);
        if (claim) {
            const unsigned peers = __match_any_sync(cm, b);
            const int slot = first_empty + __popc(peers & ((1u << lane) - 1));
            if (slot < VB) {
                __stcg(tab + ((size_t)b << VBL) + slot, (unsigned)t);
                fresh = 1;
                active = false;
            } else b = (b + 1) & bmask; /* the bucket filled up under this very call */
        }
        __syncwarp();
    }
    return fresh;
}

__device__ __forceinline__ void visited_clear(unsigned *tab, unsigned cap, int lane) {
    uint4 *t4 = reinterpret_cast<uint4 *>(tab);
    const uint4 e = make_uint4(VIS_EMPTY, VIS_EMPTY, VIS_EMPTY, VIS_EMPTY);
    for (unsigned i = lane; i < (cap >> 2); i += 32) __stcg(t4 + i, e);
    __syncwarp();
}

/* RB = beam_rows(p.ld), as a template parameter: with a run-time batch size the L2-hinted copy loop
 * compiled (nvcc 12.9) to code that faults with "illegal instruction" on sm_100a. */
template <int METRIC, int RB, int CD>
__global__ void __launch_bounds__(32, MUNINN_SEARCH_MIN_BLOCKS) hnsw_search_kernel(const SearchParams p, const int level) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x;
    TeamSmem sm;
    {
        unsigned char *b = smem_raw;
        sm.R = reinterpret_cast<unsigned long long *>(b); b += ((size_t)p.ef * 8 + 15) & ~(size_t)15;
        sm.Fk = reinterpret_cast<unsigned long long *>(b); b += (size_t)p.M0 * 8;
        sm.S = reinterpret_cast<unsigned long long *>(b); b += 32 * 8;
        sm.q = reinterpret_cast<float *>(b); b += ((size_t)p.ld * 4 + 15) / 16 * 16;
        sm.P = reinterpret_cast<float *>(b); b += (size_t)RB * TileOf<CD>::stride * 4;
        sm.Fs = reinterpret_cast<int *>(b); b += ((size_t)fs_len(p.M0) * 4 + 15) & ~(size_t)15;
        sm.bar = reinterpret_cast<unsigned long long *>(b);
        sm.phase = 0u;
    }
    if (lane == 0) bar_init(sm.bar, 1);
    __syncwarp();
    unsigned *tab = p.vis + (size_t)blockIdx.x * (size_t)p.vis_cap;
    const int nwork = level == 0 ? p.nq : *p.in_count;
    const int ef = p.ef, FP = fs_pos(p.M0);
    const unsigned long long pol = l2_evict_first_policy();

    for (;;) {
        int w = 0;
        if (lane == 0) w = atomicAdd(p.work_counter, 1);
        w = __shfl_sync(FULL, w, 0);
        if (w >= nwork) break;
        const int qi = level == 0 ? w : p.in_list[w];
        if (p.q_arrived && level == 0) {
            if (lane == 0) wait_query_arrived(p.q_arrived, qi);
            __syncwarp();
        }

        /* stage the query (pad lanes to the 4-float row stride with zeros); L2-level loads: the rows may have been
         * written by the copy engine while this kernel runs */
        const float *qg = p.queries + (size_t)qi * (size_t)p.dim;
        for (int i = lane; i < p.ld; i += 32) sm.q[i] = i < p.dim ? __ldcg(qg + i) : 0.0f;
        __syncwarp();
        float qn2 = 0.0f;
        if (METRIC == METRIC_COSINE) qn2 = query_norm2(sm.q, p.dim);

        long long n_dist = 0, n_dist_upper = 0;
        int n_expand = 0, n_greedy = 0, term = 0, entry0 = -1, size = 0;
        bool overflow = false;

        if (p.entry_slot >= 0 && p.n_live > 0) {
            /* ── greedy descent, levels max_level..1 (src/hnsw_algo.c:679-681, :257-282) ── */
            int cur = p.entry_slot;
            if (lane == 0) sm.Fs[0] = cur;
            __syncwarp();
            float cur_d = __shfl_sync(FULL, warp_dist_rows<METRIC, 4, CD>(p, sm, sm.Fs, 1, qn2, lane, pol), 0);
            __syncwarp();
            for (int l = p.max_level; l > 0; l--) {
                n_dist++; n_dist_upper++; /* the reference re-evaluates the entry at every level (:262) */
                bool changed = true;
                while (changed) {
                    changed = false;
                    n_greedy++;
                    int i0 = 0;
                    /* scan positions i0.. of cur's list in order, DB valid entries at a time; on
                     * the first improvement switch node and keep counting positions (:267-279) */
                    for (;;) {
                        const int32_t *lst = p.nbrU + ((size_t)p.upper_base[cur] + (size_t)(l - 1)) * (size_t)p.M;
                        /* gather up to DB valid entries at positions ≥ i0 */
                        int nb = 0, last_pos = p.M;
                        for (int base = i0; base < p.M && nb < DB; base += 32) {
                            int pos = base + lane;
                            int t = pos < p.M ? lst[pos] : -1;
                            unsigned vm = __ballot_sync(FULL, t >= 0);
                            int r = __popc(vm & ((1u << lane) - 1));
                            if (t >= 0 && nb + r < DB) { sm.Fs[nb + r] = t; sm.Fs[FP + nb + r] = pos; }
                            nb += __popc(vm);
                        }
                        if (nb > DB) nb = DB;
                        __syncwarp();
                        if (nb == 0) break;
                        last_pos = sm.Fs[FP + nb - 1];
                        float d = warp_dist_rows<METRIC, 4, CD>(p, sm, sm.Fs, nb, qn2, lane, pol);
                        bool better = (lane & 3) == 0 && (lane >> 2) < nb && d < cur_d; /* strict, :274 */
                        unsigned bm = __ballot_sync(FULL, better);
                        if (bm) {
                            int src = __ffs(bm) - 1, j = src >> 2;
                            n_dist += j + 1; n_dist_upper += j + 1;
                            cur_d = __shfl_sync(FULL, d, src);
                            cur = sm.Fs[j];
                            i0 = sm.Fs[FP + j] + 1;
                            changed = true;
                            n_greedy++;
                        } else {
                            n_dist += nb; n_dist_upper += nb;
                            i0 = last_pos + 1;
                        }
                        __syncwarp();
                        if (i0 >= p.M) break;
                    }
                }
            }
            entry0 = cur;

            /* ── layer-0 beam (src/hnsw_algo.c:347-448) ── */
            if (!p.deleted[cur]) { /* a deleted entry seeds nothing (:358-366) */
                n_dist++;
                if (lane == 0) sm.R[0] = make_key(cur_d, (unsigned)cur << 1);
                size = 1;
                visited_clear(tab, p.vis_cap, lane);
                visited_test_and_set(p, tab, cur, lane == 0, lane);
                __syncwarp();
                int stale = 0, lb = 0, n_ins = 1, evicted_open = 0;
                for (;;) {
                    /* closest unexpanded entry of R */
                    int pos = -1;
                    for (int base = lb & ~31; base < size; base += 32) {
                        int idx = base + lane;
                        bool open = idx < size && idx >= lb && !(sm.R[idx] & 1ull);
                        unsigned om = __ballot_sync(FULL, open);
                        if (om) {
                            pos = base + __ffs(om) - 1;
                            break;
                        }
                    }
                    if (pos < 0) { term = (size >= ef && evicted_open) ? 1 : 0; break; }
                    if (stale >= p.patience_max && size >= ef) { term = 2; break; } /* :391-392 */
                    const int c = (int)((unsigned)sm.R[pos] >> 1);
                    __syncwarp();
                    if (lane == 0) sm.R[pos] |= 1ull;
                    lb = pos + 1;
                    n_expand++;
                    if ((unsigned)(n_ins + p.M0) > p.vis_cap - (p.vis_cap >> 2)) { overflow = true; break; } /* keep the load ≤ 3/4 */

                    /* neighbour list → visited filter → fresh slots */
                    int nf = 0;
                    const int32_t *lst = p.nbr0 + (size_t)c * (size_t)p.M0;
                    /* lists of more than 32 entries (M > 16): every 32-entry chunk used to pay its own dependent chain
                     * list line → visited probe. All list lines are requested up front (they are adjacent), and the visited
                     * buckets of the NEXT chunk are pulled towards L2 while the current chunk is probed, so that only the
                     * first probe of an expansion can be a DRAM round trip. */
                    int t_next = lane < p.M0 ? lst[lane] : -1;
                    for (int base = 0; base < p.M0; base += 32) {
                        int pos0 = base + lane;
                        const int t = t_next;
                        if (base + 32 < p.M0) {
                            t_next = pos0 + 32 < p.M0 ? lst[pos0 + 32] : -1;
                            if (t_next >= 0) prefetch_l2(tab + (size_t)((((unsigned)t_next * 2654435761u) >> p.vis_shift) & ~3u));
                        }
                        int fresh = visited_test_and_set(p, tab, t, t >= 0, lane);
                        unsigned fm = __ballot_sync(FULL, fresh);
                        if (fresh) sm.Fs[nf + __popc(fm & ((1u << lane) - 1))] = t;
                        /* this chunk's fresh rows start towards L2 while the next chunk is still being probed (narrow rows only:
                         * for 3 KB rows the 24 extra prefetches per row cost more than they hide — c3 proxy 13.68 vs 13.32 ms) */
                        if (fresh && base + 32 < p.M0 && p.ld <= 128) {
                            const char *rb = reinterpret_cast<const char *>(p.vec + (size_t)t * (size_t)p.ld);
                            for (int off = 0; off < p.ld * 4; off += 128) prefetch_l2(rb + off);
                        }
                        nf += __popc(fm);
                    }
                    n_ins += nf;
                    n_dist += nf;
                    __syncwarp();
                    if (nf > RB) prefetch_rows(p, sm.Fs + RB, nf - RB, lane); /* the first batch is copied right away */
                    /* wide rows: the first batch's later column blocks would each pay a DRAM round trip */
                    if (p.ld > 128) prefetch_rows(p, sm.Fs, nf < RB ? nf : RB, lane, 512);
                    /* distances */
                    for (int b0 = 0; b0 < nf; b0 += RB) {
                        const int nb = nf - b0 < RB ? nf - b0 : RB;
                        if (nb > 8) {
                            float d = warp_dist_rows<METRIC, 2, CD>(p, sm, sm.Fs + b0, nb, qn2, lane, pol);
                            if ((lane & 1) == 0 && (lane >> 1) < nb)
                                sm.Fk[b0 + (lane >> 1)] = make_key(d, (unsigned)sm.Fs[b0 + (lane >> 1)] << 1);
                        } else {
                            float d = warp_dist_rows<METRIC, 4, CD>(p, sm, sm.Fs + b0, nb, qn2, lane, pol);
                            if ((lane & 3) == 0 && (lane >> 2) < nb)
                                sm.Fk[b0 + (lane >> 2)] = make_key(d, (unsigned)sm.Fs[b0 + (lane >> 2)] << 1);
                        }
                    }
                    __syncwarp();
                    /* R ← ef smallest of R ∪ F */
                    bool improved = false;
                    for (int base = 0; base < nf; base += 32) {
                        unsigned long long worst = size >= ef ? sm.R[ef - 1] : KEY_NONE;
                        unsigned long long f = base + lane < nf ? sm.Fk[base + lane] : KEY_NONE;
                        bool keep = (unsigned)(f >> 32) < (unsigned)(worst >> 32); /* strict on the DISTANCE, as :419 — the slot only orders */
                        unsigned km = __ballot_sync(FULL, keep);
                        int ns = __popc(km);
                        if (ns == 0) continue;
                        improved = true;
                        if (keep) sm.S[__popc(km & ((1u << lane) - 1))] = f;
                        __syncwarp();
                        int mp;
                        size = warp_merge_sorted(sm.R, size, ef, sm.S, ns, lane, &mp, &evicted_open);
                        lb = mp < lb ? mp : lb;
                    }
                    stale = improved ? 0 : stale + 1; /* :428-432 */
                }
            }
        }

        if (overflow) {
            /* visited table too small for this query: hand it to the next, larger level */
            if (lane == 0) p.out_list[atomicAdd(p.out_count_retry, 1)] = qi;
            continue;
        }
        /* results: first min(size, k) of R (src/hnsw_algo.c:695-699) */
        const int cnt = size < p.k ? size : p.k;
        for (int i = lane; i < p.k; i += 32) {
            int64_t id = -1; float d = 0.0f;
            if (i < cnt) {
                unsigned long long e = sm.R[i];
                id = p.ids[(unsigned)e >> 1];
                d = okey_inv((unsigned)(e >> 32));
            }
            p.out_ids[(size_t)qi * p.k + i] = id;
            p.out_dist[(size_t)qi * p.k + i] = d;
        }
        if (lane == 0) {
            p.out_count[qi] = cnt;
            if (p.stats) {
                MuninnSearchStats st;
                st.n_dist = n_dist; st.n_dist_upper = n_dist_upper; st.n_expand = n_expand;
                st.n_greedy_steps = n_greedy; st.term = term; st.entry0 = entry0;
                p.stats[qi] = st;
            }
        }
        __syncwarp();
    }
}

} // namespace muninn
