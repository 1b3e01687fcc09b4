/*
 * hnsw_beam.cuh — second-generation batched HNSW search kernel for sm_100a ("v8").
 *
 * Same contract as hnsw_search.cuh (the reference's hnsw_search, src/hnsw_algo.c:670-704, with
 * greedy_search_layer :257-282, beam_search_layer :347-448 and the distance arithmetic of
 * src/vec_math.c:78-143 reproduced bit for bit; canonical (distance, slot) tie order), same
 * SearchParams, same per-query counters. What changes is how a query is executed:
 *
 *  1. QUERY TEAMS OF W WARPS (W = 1, 2, 4; one CTA = one team). Warp 0 leads: it pops the next
 *     candidate, reads its neighbour list, runs the visited test-and-set and merges the survivors
 *     into R. The distance evaluations of one expansion — the bulk of the instructions — are dealt
 *     out in batches of 16 rows over all W warps, each warp staging and walking its rows through
 *     its own shared-memory ring. W = 1 is the throughput shape (24 teams per SM); W = 2 / 4 cut
 *     the dependent instruction chain of a lone query (the vtab's one MATCH per statement, the
 *     1 250-query slices of a batch split over 8 GPUs, the device-assisted builder's groups).
 *
 *  2. ROW RING. Rows no longer arrive as one 512-byte copy per row into a tile that is walked
 *     and then refilled: every row is cut into units of U floats (U = 32 / 64 / 128) and a ring of
 *     NST stages of 16 × U floats is kept full by TMA bulk copies (cp.async.bulk, one per row and
 *     unit, each stage with its own mbarrier). The unit sequence runs across batches and column
 *     blocks, so the copies of the next unit — the next column block of a wide row, or the next
 *     batch of rows — are in flight while the current one is walked. Everything the ring does not
 *     hold yet is pulled into L2 up front with one bulk prefetch per row.
 *
 * Tried here and removed (round 2, profiles/r2_bench/ab_a_ring_spec_teams.jsonl): speculation on the next
 * candidate — requesting the runner-up's neighbour list at pop time, looking its entries up read-only in
 * the visited table and pulling the unvisited rows into L2. Slower everywhere (ef=64: 2.18 vs 1.88 ms,
 * 3 KB rows: 26.5 vs 18.8 ms, a lone query 0.436 vs 0.383 ms): the runner-up is often overtaken by a fresh
 * neighbour and the kernel is bandwidth-bound enough that speculative rows cost more than they hide.
 */
#pragma once
#include "hnsw_search.cuh"

namespace muninn {
namespace beam {

constexpr int RB = 16;     /* rows per batch: two lanes per row, chains (2h, 2h+1) packed as f32x2 */
constexpr int DBB = 16;    /* valid entries of an upper-layer list evaluated per step of the greedy descent: a whole list at M ≤ 16, so a
                            * step that finds no improvement is one round trip, not two (the one-warp kernel's DB = 8 suits its 4-lane rows) */
constexpr int LIST_CH = 2; /* 32-entry chunks of a neighbour list requested up front (M0 ≤ 64; longer lists read the rest in the loop) */

__host__ __device__ inline size_t a16(size_t v) { return (v + 15) & ~(size_t)15; }

__host__ __device__ inline int fk_len(int M0) { return fs_pos(M0) > DBB ? fs_pos(M0) : DBB; } /* keys of one expansion, or of one descent step */

__host__ __device__ inline size_t smem_bytes(int ef, int ld, int M0, int U, int NST, int W) {
    size_t b = a16((size_t)ef * 8) + (size_t)fk_len(M0) * 8 + 32 * 8;
    b += a16((size_t)ld * 4);
    b += (size_t)W * NST * RB * (U + 4) * 4;
    b += a16((size_t)(fs_pos(M0) + DBB) * 4);
    b += a16((size_t)W * NST * 8);
    b += 16; /* team control words */
    return b;
}

/* Visited test-and-set on a table in SHARED memory (one key per lane, open addressing, linear probing, claim by
 * compare-and-swap): what the small-batch teams use at level 0. The global-memory table costs a lone query ≈ 1 600 of
 * its ≈ 6 700 cycles per expansion — a warp-wide gather of 32 scattered buckets from L2 (≈ 800 cycles: the slowest of 32
 * requests) and a match.any to order the claims (≈ 450) — profiles/r2_bench/lone_query_phases.jsonl; in shared memory
 * both are tens of cycles. Same membership semantics, so the results cannot differ. */
__device__ __forceinline__ int visited_test_and_set_smem(unsigned *tab, unsigned cap_mask, unsigned shift, int t, bool valid) {
    int fresh = 0;
    if (valid) {
        unsigned i = ((unsigned)t * 2654435761u) >> shift;
        for (;;) { /* ends: the caller keeps the load ≤ 3/4 */
            const unsigned old = atomicCAS(tab + i, VIS_EMPTY, (unsigned)t);
            if (old == VIS_EMPTY) { fresh = 1; break; }
            if (old == (unsigned)t) break;
            i = (i + 1) & cap_mask;
        }
    }
    __syncwarp();
    return fresh;
}

/* warp_merge_sorted (muninn_dev.cuh) for lists of at most 64 keys, as the small-batch teams use it: same contract, same
 * result. Both 32-key halves of R sit in registers, so the rank of every new key in R comes out of two ballots per key
 * instead of a binary search through shared memory, the shift of every old key out of the same comparisons, and nothing
 * is read after the first write — one pass, two barriers (the general routine: ≈ 1 400 cycles per expansion of a lone
 * query, profiles/r2_bench/lone_query_phases.jsonl). */
__device__ __forceinline__ int warp_merge_sorted64(unsigned long long *R, int size, int cap, const unsigned long long *S,
                                                   int ns, int lane, int *min_pos, int *evicted_open) {
    const unsigned long long r0 = lane < size ? R[lane] : KEY_NONE, r1 = 32 + lane < size ? R[32 + lane] : KEY_NONE;
    const unsigned long long f = lane < ns ? S[lane] : KEY_NONE;
    int c0 = 0, c1 = 0, rF = 0, rR = 0, mp = 0x7fffffff;
    for (int i = 0; i < ns; i++) {
        const unsigned long long fi = S[i];
        const int below = __popc(__ballot_sync(FULL, r0 < fi)) + __popc(__ballot_sync(FULL, r1 < fi)); /* rank of key i in R */
        c0 += fi < r0 ? 1 : 0;
        c1 += fi < r1 ? 1 : 0;
        rF += fi < f ? 1 : 0;
        if (i == lane) rR = below;
        mp = below < mp ? below : mp;
    }
    __syncwarp();
    int ev = 0;
    if (lane < size && c0 > 0) {
        if (lane + c0 < cap) R[lane + c0] = r0;
        else if (!(r0 & 1ull)) ev = 1;
    }
    if (32 + lane < size && c1 > 0) {
        if (32 + lane + c1 < cap) R[32 + lane + c1] = r1;
        else if (!(r1 & 1ull)) ev = 1;
    }
    if (lane < ns) {
        if (rR + rF < cap) R[rR + rF] = f;
        else ev = 1; /* a new key that does not fit is an open candidate the reference would have queued */
    }
    __syncwarp();
    if (__any_sync(FULL, ev)) *evicted_open = 1;
    *min_pos = mp;
    const int nsz = size + ns;
    return nsz < cap ? nsz : cap;
}

/* whole row (or its tail) towards L2 with one TMA prefetch */
__device__ __forceinline__ void l2_prefetch_bytes(const void *g, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(g), "r"(bytes) : "memory");
}

/*
 * The ring of one warp. Rows Fs[0..nf) are dealt out in batches of RB over W warps (batch b goes to warp b mod W);
 * this warp's batches × the row's units form one sequence i = 0..total) that is streamed through NST stages.
 */
template <int METRIC, int CD, int U, int NST>
struct RowRing {
    static constexpr int PS = U + 4; /* stage row stride in floats: ≡ 4 (mod 32) → conflict-free 64-bit reads */
    float *P;
    unsigned long long *bar;
    unsigned phase;
    const int *Fs;
    int nf, wid, W, upb, total, rowf;
    int rb; /* rows per batch of this call: RB, or an even share of nf when the team has a warp for every share */

    __device__ __forceinline__ void issue(const SearchParams &p, int bi, int u, int s, int lane, unsigned long long pol) {
        const int r0 = (wid + bi * W) * rb;
        const int nb = nf - r0 < rb ? nf - r0 : rb;
        const int fl = CD ? U : (rowf - u * U < U ? rowf - u * U : U);
        if (lane == 0) bar_expect_tx(bar + s, (unsigned)(nb * fl * 4));
        if (lane < nb)
            bulk_row_copy(P + (s * RB + lane) * PS, p.vec + (size_t)Fs[r0 + lane] * (size_t)p.ld + u * U, (unsigned)(fl * 4), bar + s, pol);
    }

    /* start streaming: the first NST units of this warp's share */
    __device__ __forceinline__ void begin(const SearchParams &p, const int *fs, int n, int warp_id, int n_warps, int lane,
                                          unsigned long long pol) {
        Fs = fs; nf = n; wid = warp_id; W = n_warps;
        rowf = CD ? CD : p.ld;
        upb = (rowf + U - 1) / U;
        /* a team of several warps is there to cut the latency of ONE expansion: when its fresh rows fit one batch per warp
         * they are dealt out evenly (21 rows over 4 warps: 6 + 6 + 6 + 3 instead of 16 + 5), so that no warp issues more
         * than its share of the per-row TMA copies — the issue loop, not the memory, was the longest phase of a lone
         * query's expansion (profiles/r2_bench/lone_query_phases.jsonl) */
        rb = RB;
        if (W > 1 && nf <= W * RB) { rb = (nf + W - 1) / W; rb = rb < 1 ? 1 : rb; }
        const int nbatch = (nf + rb - 1) / rb;
        const int mine = wid < nbatch ? (nbatch - wid + W - 1) / W : 0;
        total = mine * upb;
        int bi = 0, u = 0;
#pragma unroll
        for (int i = 0; i < NST; i++) {
            if (i < total) issue(p, bi, u, i, lane, pol);
            if (++u == upb) { u = 0; bi++; }
        }
    }

    /* walk the units as they land; key of row r → Fk[r] */
    __device__ __forceinline__ void run(const SearchParams &p, const float *q, unsigned long long *Fk, float qn2, int lane,
                                        unsigned long long pol) {
        const int row = lane >> 1, h = lane & 1;
        const int cfull = CD ? CD / 4 : p.dim >> 2, rem = CD ? 0 : p.dim & 3;
        unsigned long long s01 = 0ull;
        float tl0 = 0.0f, tl1 = 0.0f, tl2 = 0.0f, nb2 = 1.0f;
        int bi = 0, u = 0, s = 0;           /* unit being walked */
        int nbi = 0, nu = 0;                /* unit to issue next = i + NST */
#pragma unroll
        for (int i = 0; i < NST; i++) { if (++nu == upb) { nu = 0; nbi++; } }
        for (int i = 0; i < total; i++) {
            const int r0 = (wid + bi * W) * rb;
            const int nb = nf - r0 < rb ? nf - r0 : rb;
            const bool mine = row < nb;
            if (u == 0) {
                s01 = 0ull; tl0 = tl1 = tl2 = 0.0f;
                if (METRIC == METRIC_COSINE && mine && h == 0) nb2 = p.norm2[Fs[r0 + row]];
            }
            bar_wait(bar + s, (phase >> s) & 1u);
            phase ^= 1u << s;
            if (mine) {
                const float *xr = P + (s * RB + row) * PS + 2 * h;
                const float *qr = q + u * U + 2 * h;
                int nfull = U / 4;
                if (!CD) {
                    nfull = cfull - u * (U / 4);
                    nfull = nfull < 0 ? 0 : (nfull > U / 4 ? U / 4 : nfull);
                }
#pragma unroll (CD ? U / 4 : 8)
                for (int cc = 0; cc < nfull; cc++) {
                    const unsigned long long x = *reinterpret_cast<const unsigned long long *>(xr + 4 * cc);
                    const unsigned long long qq = *reinterpret_cast<const unsigned long long *>(qr + 4 * cc);
                    unsigned long long pr;
                    if (METRIC == METRIC_L2) {
                        const unsigned long long d = f2_sub(qq, x);
                        pr = f2_pack(__fmul_rn(f2_lo(d), f2_lo(d)), __fmul_rn(f2_hi(d), f2_hi(d)));
                    } else pr = f2_pack(__fmul_rn(f2_lo(qq), f2_lo(x)), __fmul_rn(f2_hi(qq), f2_hi(x)));
                    s01 = f2_add(s01, pr);
                }
                if (!CD && rem && h == 0 && cfull >= u * (U / 4) && cfull < (u + 1) * (U / 4)) {
                    /* the partial chunk: its products are added after the horizontal sum (src/vec_math.c:91-94) */
                    const float *xt = P + (s * RB + row) * PS + 4 * (cfull - u * (U / 4)), *qt = q + 4 * cfull;
                    tl0 = prod1<METRIC>(qt[0], xt[0]);
                    if (rem > 1) tl1 = prod1<METRIC>(qt[1], xt[1]);
                    if (rem > 2) tl2 = prod1<METRIC>(qt[2], xt[2]);
                }
            }
            __syncwarp(); /* every lane is done with stage s */
            if (i + NST < total) issue(p, nbi, nu, s, lane, pol);
            if (++nu == upb) { nu = 0; nbi++; }
            if (u == upb - 1) {
                const float s0 = f2_lo(s01), s1 = f2_hi(s01);
                const float c2 = __shfl_sync(FULL, s0, lane | 1), c3 = __shfl_sync(FULL, s1, lane | 1);
                float sum = __fadd_rn(__fadd_rn(__fadd_rn(s0, s1), c2), c3);
                if (!CD && rem) {
                    sum = __fadd_rn(sum, tl0);
                    if (rem > 1) sum = __fadd_rn(sum, tl1);
                    if (rem > 2) sum = __fadd_rn(sum, tl2);
                }
                float d;
                if (METRIC == METRIC_L2) d = sum;
                else if (METRIC == METRIC_IP) d = -sum;
                else d = cosine_finish(sum, qn2, nb2);
                if (mine && h == 0) Fk[r0 + row] = make_key(d, (unsigned)Fs[r0 + row] << 1);
            }
            if (++u == upb) { u = 0; bi++; }
            if (++s == NST) s = 0;
        }
    }
};

template <int W>
__device__ __forceinline__ void team_sync() {
    if (W > 1) __syncthreads();
    else __syncwarp();
}

template <int W> struct MinBlocks { static constexpr int v = W == 1 ? MUNINN_SEARCH_MIN_BLOCKS : (W == 2 ? 10 : 4); }; /* W = 2: 10 teams/SM keep a 1 250-query slice resident */

template <int METRIC, int CD, int U, int NST, int W>
__global__ void __launch_bounds__(32 * W, MinBlocks<W>::v) hnsw_beam_kernel(const SearchParams p, const int level) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using Ring = RowRing<METRIC, CD, U, NST>;
    constexpr int NT = 32 * W;
    const int lane = threadIdx.x & 31, warp = W == 1 ? 0 : (int)(threadIdx.x >> 5);
    const bool leader = W == 1 || warp == 0;

    unsigned long long *R, *Fk, *S;
    float *q;
    int *Fs;
    volatile int *ctl;
    unsigned *stab; /* the visited table in shared memory, when SearchParams::vis_smem gives it room */
    Ring ring;
    {
        unsigned char *b = smem_raw;
        R = reinterpret_cast<unsigned long long *>(b); b += a16((size_t)p.ef * 8);
        Fk = reinterpret_cast<unsigned long long *>(b); b += (size_t)fk_len(p.M0) * 8;
        S = reinterpret_cast<unsigned long long *>(b); b += 32 * 8;
        q = reinterpret_cast<float *>(b); b += a16((size_t)p.ld * 4);
        ring.P = reinterpret_cast<float *>(b) + (size_t)warp * NST * RB * Ring::PS; b += (size_t)W * NST * RB * Ring::PS * 4;
        Fs = reinterpret_cast<int *>(b); b += a16((size_t)(fs_pos(p.M0) + DBB) * 4); /* fresh slots; the descent keeps list positions at Fs[FP + i] */
        ring.bar = reinterpret_cast<unsigned long long *>(b) + warp * NST; b += a16((size_t)W * NST * 8);
        ctl = reinterpret_cast<volatile int *>(b); b += 16;
        stab = reinterpret_cast<unsigned *>(b);
        ring.phase = 0u;
    }
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < NST; s++) bar_init(ring.bar + s, 1);
    }
    team_sync<W>();

    const bool svis = W > 1 && level == 0 && p.vis_smem > 0;
    unsigned *tab = svis ? stab : p.vis + (size_t)blockIdx.x * (size_t)p.vis_cap;
    const unsigned cap_mask = p.vis_cap - 1;
    const int nwork = level == 0 ? p.nq : *p.in_count;
    const int ef = p.ef, FP = fs_pos(p.M0);
    const unsigned long long pol = l2_evict_first_policy();
    const unsigned row_bytes = (unsigned)p.ld * 4u;

    for (;;) {
        int w = 0;
        if (W == 1) {
            if (lane == 0) w = atomicAdd(p.work_counter, 1);
            w = __shfl_sync(FULL, w, 0);
        } else {
            if (threadIdx.x == 0) ctl[1] = atomicAdd(p.work_counter, 1);
            __syncthreads();
            w = ctl[1];
        }
        if (w >= nwork) break;
        const int qi = level == 0 ? w : p.in_list[w];
        if (p.q_arrived && level == 0) { /* SearchParams::q_arrived: the batch may still be streaming in */
            if (threadIdx.x == 0) wait_query_arrived(p.q_arrived, qi);
            team_sync<W>();
        }

        /* stage the query (zero-padded to the row stride) and clear the team's visited table */
        const float *qg = p.queries + (size_t)qi * (size_t)p.dim;
        for (int i = threadIdx.x; i < p.ld; i += NT) q[i] = i < p.dim ? __ldcg(qg + i) : 0.0f;
        {
            uint4 *t4 = reinterpret_cast<uint4 *>(tab);
            const uint4 e = make_uint4(VIS_EMPTY, VIS_EMPTY, VIS_EMPTY, VIS_EMPTY);
            if (svis) for (unsigned i = threadIdx.x; i < (p.vis_cap >> 2); i += NT) t4[i] = e;
            else for (unsigned i = threadIdx.x; i < (p.vis_cap >> 2); i += NT) __stcg(t4 + i, e);
        }
        team_sync<W>();
        float qn2 = 0.0f;
        if (METRIC == METRIC_COSINE) qn2 = query_norm2(q, p.dim);

        long long n_dist = 0, n_dist_upper = 0;
        int n_expand = 0, n_greedy = 0, term = 0, entry0 = -1, size = 0;
        bool overflow = false;
        /* leader state of the beam */
        int stale = 0, lb = 0, n_ins = 1, evicted_open = 0;
        bool run_beam = false;

        if (leader && p.entry_slot >= 0 && p.n_live > 0) {
            /* ── greedy descent, levels max_level..1 (src/hnsw_algo.c:679-681, :257-282) ── */
            int cur = p.entry_slot;
            if (lane == 0) Fs[0] = cur;
            __syncwarp();
            ring.begin(p, Fs, 1, 0, 1, lane, pol);
            ring.run(p, q, Fk, qn2, lane, pol);
            __syncwarp();
            unsigned cur_k = (unsigned)(Fk[0] >> 32);
            __syncwarp();
            for (int l = p.max_level; l > 0; l--) {
                n_dist++; n_dist_upper++; /* the reference re-evaluates the entry at every level (:262) */
                bool changed = true;
                while (changed) {
                    changed = false;
                    n_greedy++;
                    int i0 = 0;
                    /* scan positions i0.. of cur's list in order, DBB valid entries at a time; on the first
                     * improvement switch node and keep counting positions (:267-279) */
                    for (;;) {
                        const int32_t *lst = p.nbrU + ((size_t)p.upper_base[cur] + (size_t)(l - 1)) * (size_t)p.M;
                        int nb = 0, last_pos = p.M;
                        for (int base = i0; base < p.M && nb < DBB; base += 32) {
                            const int pos = base + lane;
                            const int t = pos < p.M ? lst[pos] : -1;
                            const unsigned vm = __ballot_sync(FULL, t >= 0);
                            const int r = __popc(vm & ((1u << lane) - 1));
                            if (t >= 0 && nb + r < DBB) { Fs[nb + r] = t; Fs[FP + nb + r] = pos; }
                            nb += __popc(vm);
                        }
                        if (nb > DBB) nb = DBB;
                        __syncwarp();
                        if (nb == 0) break;
                        last_pos = Fs[FP + nb - 1];
                        ring.begin(p, Fs, nb, 0, 1, lane, pol);
                        ring.run(p, q, Fk, qn2, lane, pol);
                        __syncwarp();
                        const unsigned dk = lane < nb ? (unsigned)(Fk[lane] >> 32) : 0xffffffffu;
                        const unsigned bm = __ballot_sync(FULL, lane < nb && dk < cur_k); /* strict, :274 */
                        if (bm) {
                            const int j = __ffs(bm) - 1;
                            n_dist += j + 1; n_dist_upper += j + 1;
                            cur_k = __shfl_sync(FULL, dk, j);
                            cur = Fs[j];
                            i0 = Fs[FP + j] + 1;
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
            /* ── seed of the layer-0 beam (src/hnsw_algo.c:358-366): a deleted entry seeds nothing ── */
            if (!p.deleted[cur]) {
                n_dist++;
                if (lane == 0) R[0] = ((unsigned long long)cur_k << 32) | ((unsigned)cur << 1);
                size = 1;
                if (svis) visited_test_and_set_smem(tab, cap_mask, p.vis_shift, cur, lane == 0);
                else visited_test_and_set(p, tab, cur, lane == 0, lane);
                __syncwarp();
                run_beam = true;
            }
        }
        if (W > 1) {
            if (threadIdx.x == 0) ctl[2] = run_beam ? 1 : 0;
            __syncthreads();
            run_beam = ctl[2] != 0;
        }

        /* ── layer-0 beam (src/hnsw_algo.c:347-448) ── */
        while (run_beam) {
            int nf = -1;
            if (leader) {
                /* closest unexpanded entry of R */
                int pos = -1;
                for (int base = lb & ~31; base < size; base += 32) {
                    const int idx = base + lane;
                    const bool open = idx < size && idx >= lb && !(R[idx] & 1ull);
                    const unsigned om = __ballot_sync(FULL, open);
                    if (om) { pos = base + __ffs(om) - 1; break; }
                }
                if (pos < 0) term = (size >= ef && evicted_open) ? 1 : 0;
                else if (stale >= p.patience_max && size >= ef) term = 2; /* :391-392 */
                else {
                    const int c = (int)((unsigned)R[pos] >> 1);
                    __syncwarp();
                    if (lane == 0) R[pos] |= 1ull;
                    lb = pos + 1;
                    n_expand++;
                    if ((unsigned)(n_ins + p.M0) > p.vis_cap - (p.vis_cap >> 2)) overflow = true; /* keep the load ≤ 3/4 */
                    else {
                        /* neighbour list: all of its lines are requested up front */
                        const int32_t *lst = p.nbr0 + (size_t)c * (size_t)p.M0;
                        int t[LIST_CH];
#pragma unroll
                        for (int ch = 0; ch < LIST_CH; ch++) {
                            const int pos0 = ch * 32 + lane;
                            t[ch] = pos0 < p.M0 ? lst[pos0] : -1;
                        }
                        /* lists of more than 32 entries: the second chunk's visited buckets move towards L2 while the first is probed */
                        if (!svis && LIST_CH > 1 && p.M0 > 32 && t[1] >= 0)
                            prefetch_l2(tab + (size_t)((((unsigned)t[1] * 2654435761u) >> p.vis_shift) & ~3u));
                        /* visited filter → fresh slots */
                        nf = 0;
                        for (int base = 0; base < p.M0; base += 32) {
                            const int pos0 = base + lane;
                            int tt;
                            if (base == 0) tt = t[0];
                            else if (LIST_CH > 1 && base == 32) tt = t[1];
                            else tt = pos0 < p.M0 ? lst[pos0] : -1;
                            const int fresh = svis ? visited_test_and_set_smem(tab, cap_mask, p.vis_shift, tt, tt >= 0)
                                                   : visited_test_and_set(p, tab, tt, tt >= 0, lane);
                            const unsigned fm = __ballot_sync(FULL, fresh);
                            if (fresh) Fs[nf + __popc(fm & ((1u << lane) - 1))] = tt;
                            nf += __popc(fm);
                        }
                        n_ins += nf;
                        n_dist += nf;
                        __syncwarp();
                    }
                }
                if (W > 1 && lane == 0) ctl[0] = nf;
            }
            if (W > 1) {
                __syncthreads();
                nf = ctl[0];
            }
            if (nf < 0) break;

            if (nf > 0) {
                ring.begin(p, Fs, nf, warp, W, lane, pol);
                /* what the rings do not hold yet → L2, one bulk prefetch per row */
                const unsigned in_ring = (unsigned)(NST * U * 4) < row_bytes ? (unsigned)(NST * U * 4) : row_bytes;
                for (int r = threadIdx.x; r < nf; r += NT) {
                    const unsigned off = r < W * RB ? in_ring : 0u;
                    if (off < row_bytes)
                        l2_prefetch_bytes(reinterpret_cast<const char *>(p.vec + (size_t)Fs[r] * (size_t)p.ld) + off, row_bytes - off);
                }
            }
            if (nf > 0) ring.run(p, q, Fk, qn2, lane, pol);
            team_sync<W>();

            if (leader) {
                /* R ← ef smallest of R ∪ F */
                bool improved = false;
                for (int base = 0; base < nf; base += 32) {
                    const unsigned long long worst = size >= ef ? R[ef - 1] : KEY_NONE;
                    const unsigned long long f = base + lane < nf ? Fk[base + lane] : KEY_NONE;
                    const bool keep = (unsigned)(f >> 32) < (unsigned)(worst >> 32); /* strict on the DISTANCE, as :419 — the slot only orders */
                    const unsigned km = __ballot_sync(FULL, keep);
                    const int ns = __popc(km);
                    if (ns == 0) continue;
                    improved = true;
                    if (keep) S[__popc(km & ((1u << lane) - 1))] = f;
                    __syncwarp();
                    int mp;
                    size = (W > 1 && ef <= 64) ? warp_merge_sorted64(R, size, ef, S, ns, lane, &mp, &evicted_open)
                                               : warp_merge_sorted(R, size, ef, S, ns, lane, &mp, &evicted_open);
                    lb = mp < lb ? mp : lb;
                }
                stale = improved ? 0 : stale + 1; /* :428-432 */
            }
        }

        if (leader) {
            if (overflow) {
                /* visited table too small for this query: hand it to the next, larger level */
                if (lane == 0) p.out_list[atomicAdd(p.out_count_retry, 1)] = qi;
            } else {
                /* results: first min(size, k) of R (src/hnsw_algo.c:695-699) */
                const int cnt = size < p.k ? size : p.k;
                for (int i = lane; i < p.k; i += 32) {
                    int64_t id = -1; float d = 0.0f;
                    if (i < cnt) {
                        const unsigned long long e = R[i];
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
            }
        }
        team_sync<W>(); /* q, R and the control words are rewritten by the next query */
    }
}

} // namespace beam
} // namespace muninn
