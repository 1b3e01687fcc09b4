/*
 * hnsw_build.c — host-side HNSW graph construction in the reference's format.
 *
 * The search path this repo accelerates runs "over the graph the reference builds"
 * (BASELINE.json north_star).  The reference's own hnsw_insert is quadratic in two
 * places (linear-scan VisitedSet, src/hnsw_algo.c:318-325; O(deg²) mutual-neighbour
 * count evaluated for every entry of every over-full list, :460-475 + :608-617), so it
 * cannot produce the 1M–10M node graphs of BASELINE.json configs 2–4 in practical time
 * (SURVEY.md finding 7, risk R1).  This file implements the same construction
 * algorithm from scratch on flat arrays:
 *
 *   random level        src/hnsw_algo.c:19-30, :240-248   xorshift32 seed 42, floor(-ln(u)/ln(M))
 *   greedy descent      src/hnsw_algo.c:257-282           list-order scan with in-loop switch
 *   beam per level      src/hnsw_algo.c:347-448           ef = ef_construction, patience rule included
 *   closest-M select    src/hnsw_algo.c:497-516
 *   bidirectional link  src/hnsw_algo.c:582-591
 *   MN-RU prune         src/hnsw_algo.c:601-646           distance asc, mutual-neighbour count desc on ties
 *
 * Two modes:
 *   exact  (batch = 0): one insert at a time. O(1) visited stamps and a lazily evaluated
 *          mutual-neighbour count (it only matters when two distances are bit-equal) are
 *          pure representation changes, so the graph is IDENTICAL to the reference's —
 *          entry point, levels and ordered neighbour lists (tests/test_host.py diffs it
 *          against the real reference).
 *   batch  (batch > 0): nodes are inserted in id order in groups; each group searches the
 *          graph as it stood before the group (parallel, read-only) and is then linked in
 *          id order, reverse edges grouped per target so every list has one writer.  The
 *          result is deterministic for a given batch size, independent of thread count,
 *          and is a graph of the reference's construction rule but not its exact
 *          sequential graph.  Bit-equal candidate distances are ordered by slot here (the
 *          reference's heap order among equal keys is unspecified).  Makes full-size benchmark
 *          inputs and bulk-loaded tables (tools/bulk_load.py) in minutes.
 *   group  the batch mode one group at a time (muninn_build_group_begin / _group_link), for a
 *          caller that runs the layer-0 candidate searches itself — builder.build_graph_device
 *          runs them as one muninn_cuda_search_batch per group on a device copy of the graph
 *          and writes the touched records back with muninn_cuda_index_patch; muninn_build_run is
 *          the same loop in C over a MuninnBuildDevice (three function pointers).  Same graph as
 *          the batch mode wherever no two candidate distances are bit-equal.
 *
 * Every stored edge keeps the distance it was created with (as the reference's _edges table
 * does, src/hnsw_vtab.c:163-169), so a prune re-reads 2M+1 floats instead of recomputing them.
 *
 * Distances are evaluated in the reference's SSE order (src/vec_math.c:78-143) so that
 * the exact mode reproduces its comparisons bit for bit.
 */
#define _POSIX_C_SOURCE 200809L
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <immintrin.h>

#include "muninn_build.h" /* include/: the ABI this file implements */

enum { MB_L2 = 0, MB_COSINE = 1, MB_IP = 2 };
#define MB_MAX_LEVELS 32 /* src/hnsw_algo.h:14 */

typedef struct { int32_t slot; float key; } HItem;
typedef struct { HItem *a; int size, cap; } Heap;

typedef struct {
    uint32_t *stamp; uint32_t cur;
    Heap cand, res;
    int32_t *r_slot; float *r_dist; /* beam output, ascending, capacity efc */
    int32_t *sel;                   /* per level selected: [MB_MAX_LEVELS][M0] */
    float *sel_d;                   /* their distances to the new node */
    int32_t *fresh; float *fresh_d; /* unvisited entries of the list being expanded: [M0 + 1] */
    int32_t *sel_cnt;               /* [MB_MAX_LEVELS] */
} Scratch;

typedef struct {
    int dim, metric, M, M0, efc;
    int64_t n, cap;
    const float *vecs;
    int32_t *level;
    int32_t *nbr0; int32_t *cnt0;       /* layer 0: stride M0+1 */
    float *d0, *dU;                     /* distance of every stored edge, parallel to nbr0 / nbrU (the reference
                                           persists the same number in _edges.distance, src/hnsw_vtab.c:163-169) */
    int64_t *upper_base;                /* first upper list of a node, or -1 */
    int32_t *nbrU; int32_t *cntU;       /* upper lists: stride M+1 */
    int64_t n_upper, cap_upper;
    int32_t entry; int max_level;
    uint32_t rng; double level_mult;
    int n_scratch; Scratch *scratch;
    int lazy_mn_parallel_skip;
    int log_touched; int32_t *touched; int64_t n_touched, cap_touched;
    uint32_t *touch_stamp; uint32_t touch_epoch; /* a slot is logged once per epoch (= until the log is taken) */
    int64_t pending_g0, pending_g1; int pending_raise; /* group opened by muninn_build_group_begin */
    void *req_buf; int64_t req_cap; int64_t *grp_buf; int64_t grp_cap; /* link_group scratch, kept across groups */
    double t_phase[3]; /* seconds in link_group: searches + own lists, request sort, reverse edges + prune */
} Builder;

/* ───────────── distances, reference SSE order ───────────── */

static inline float hsum(__m128 v) {
    float t[4];
    _mm_storeu_ps(t, v);
    return t[0] + t[1] + t[2] + t[3];
}

static float dist_l2(const float *a, const float *b, int dim) {
    __m128 acc = _mm_setzero_ps();
    int i = 0;
    for (; i + 4 <= dim; i += 4) {
        __m128 d = _mm_sub_ps(_mm_loadu_ps(a + i), _mm_loadu_ps(b + i));
        acc = _mm_add_ps(acc, _mm_mul_ps(d, d));
    }
    float s = hsum(acc);
    for (; i < dim; i++) { float d = a[i] - b[i]; float p = d * d; s = s + p; }
    return s;
}

static float dist_ip(const float *a, const float *b, int dim) {
    __m128 acc = _mm_setzero_ps();
    int i = 0;
    for (; i + 4 <= dim; i += 4) acc = _mm_add_ps(acc, _mm_mul_ps(_mm_loadu_ps(a + i), _mm_loadu_ps(b + i)));
    float s = hsum(acc);
    for (; i < dim; i++) { float p = a[i] * b[i]; s = s + p; }
    return -s;
}

static float dist_cos(const float *a, const float *b, int dim) {
    __m128 ad = _mm_setzero_ps(), aa = _mm_setzero_ps(), ab = _mm_setzero_ps();
    int i = 0;
    for (; i + 4 <= dim; i += 4) {
        __m128 va = _mm_loadu_ps(a + i), vb = _mm_loadu_ps(b + i);
        ad = _mm_add_ps(ad, _mm_mul_ps(va, vb));
        aa = _mm_add_ps(aa, _mm_mul_ps(va, va));
        ab = _mm_add_ps(ab, _mm_mul_ps(vb, vb));
    }
    float dot = hsum(ad), na = hsum(aa), nb = hsum(ab);
    for (; i < dim; i++) {
        float p;
        p = a[i] * b[i]; dot = dot + p;
        p = a[i] * a[i]; na = na + p;
        p = b[i] * b[i]; nb = nb + p;
    }
    float denom = sqrtf(na) * sqrtf(nb);
    if (denom < 1e-30f) return 1.0f;
    return 1.0f - (dot / denom);
}

static inline float dist(const Builder *b, const float *x, const float *y) {
    return b->metric == MB_L2 ? dist_l2(x, y, b->dim) : b->metric == MB_COSINE ? dist_cos(x, y, b->dim)
                                                                                : dist_ip(x, y, b->dim);
}

/* Two rows at once on AVX2 hosts: the low and the high 128-bit half of every register run the reference's 4-lane
 * SSE chain for row x0 and row x1 respectively (the query is broadcast to both halves), so each row's result is
 * the same sequence of IEEE operations as dist_*() above — bit-identical, checked against the unmodified reference
 * by tests/test_host.py through the graphs it produces. target("avx2") does not enable FMA: mul and add stay two
 * roundings, as in the reference's SSE path. */
#define AVX2_FN __attribute__((target("avx2")))

AVX2_FN static inline __m256 two_rows(const float *x0, const float *x1) {
    return _mm256_insertf128_ps(_mm256_castps128_ps256(_mm_loadu_ps(x0)), _mm_loadu_ps(x1), 1);
}

AVX2_FN static void dist2_l2(const float *q, const float *x0, const float *x1, int dim, float *o0, float *o1) {
    __m256 acc = _mm256_setzero_ps();
    int i = 0;
    for (; i + 4 <= dim; i += 4) {
        __m256 d = _mm256_sub_ps(_mm256_broadcast_ps((const __m128 *)(q + i)), two_rows(x0 + i, x1 + i));
        acc = _mm256_add_ps(acc, _mm256_mul_ps(d, d));
    }
    float t[8];
    _mm256_storeu_ps(t, acc);
    float s0 = t[0] + t[1] + t[2] + t[3], s1 = t[4] + t[5] + t[6] + t[7];
    for (; i < dim; i++) {
        float d0 = q[i] - x0[i], d1 = q[i] - x1[i];
        float p0 = d0 * d0, p1 = d1 * d1;
        s0 = s0 + p0; s1 = s1 + p1;
    }
    *o0 = s0; *o1 = s1;
}

AVX2_FN static void dist2_ip(const float *q, const float *x0, const float *x1, int dim, float *o0, float *o1) {
    __m256 acc = _mm256_setzero_ps();
    int i = 0;
    for (; i + 4 <= dim; i += 4)
        acc = _mm256_add_ps(acc, _mm256_mul_ps(_mm256_broadcast_ps((const __m128 *)(q + i)), two_rows(x0 + i, x1 + i)));
    float t[8];
    _mm256_storeu_ps(t, acc);
    float s0 = t[0] + t[1] + t[2] + t[3], s1 = t[4] + t[5] + t[6] + t[7];
    for (; i < dim; i++) {
        float p0 = q[i] * x0[i], p1 = q[i] * x1[i];
        s0 = s0 + p0; s1 = s1 + p1;
    }
    *o0 = -s0; *o1 = -s1;
}

AVX2_FN static void dist2_cos(const float *q, const float *x0, const float *x1, int dim, float *o0, float *o1) {
    __m256 ad = _mm256_setzero_ps(), aa = _mm256_setzero_ps(), ab = _mm256_setzero_ps();
    int i = 0;
    for (; i + 4 <= dim; i += 4) {
        __m256 va = _mm256_broadcast_ps((const __m128 *)(q + i)), vb = two_rows(x0 + i, x1 + i);
        ad = _mm256_add_ps(ad, _mm256_mul_ps(va, vb));
        aa = _mm256_add_ps(aa, _mm256_mul_ps(va, va));
        ab = _mm256_add_ps(ab, _mm256_mul_ps(vb, vb));
    }
    float td[8], ta[8], tb[8];
    _mm256_storeu_ps(td, ad); _mm256_storeu_ps(ta, aa); _mm256_storeu_ps(tb, ab);
    const float *x[2] = {x0, x1};
    float *o[2] = {o0, o1};
    for (int h = 0; h < 2; h++) {
        float dot = td[4 * h] + td[4 * h + 1] + td[4 * h + 2] + td[4 * h + 3];
        float na = ta[4 * h] + ta[4 * h + 1] + ta[4 * h + 2] + ta[4 * h + 3];
        float nb = tb[4 * h] + tb[4 * h + 1] + tb[4 * h + 2] + tb[4 * h + 3];
        for (int j = i; j < dim; j++) {
            float p;
            p = q[j] * x[h][j]; dot = dot + p;
            p = q[j] * q[j]; na = na + p;
            p = x[h][j] * x[h][j]; nb = nb + p;
        }
        float denom = sqrtf(na) * sqrtf(nb);
        *o[h] = denom < 1e-30f ? 1.0f : 1.0f - (dot / denom);
    }
}

static int g_avx2 = -1;

/* distances from q to two rows; same values as two dist() calls */
static inline void dist_pair(const Builder *b, const float *q, const float *x0, const float *x1, float *o0, float *o1) {
    if (g_avx2 > 0) {
        if (b->metric == MB_L2) dist2_l2(q, x0, x1, b->dim, o0, o1);
        else if (b->metric == MB_COSINE) dist2_cos(q, x0, x1, b->dim, o0, o1);
        else dist2_ip(q, x0, x1, b->dim, o0, o1);
    } else {
        *o0 = dist(b, q, x0);
        *o1 = dist(b, q, x1);
    }
}

/* ───────────── heap with the reference's sift rules (src/priority_queue.c:18-42) ───────────── */

static void hp_reset(Heap *h, int cap) {
    if (cap < 4) cap = 4;
    if (!h->a || h->cap < cap) { free(h->a); h->a = (HItem *)malloc((size_t)(cap + 1) * sizeof(HItem)); h->cap = cap; }
    h->size = 0;
}
static void hp_push(Heap *h, int32_t slot, float key) {
    if (h->size >= h->cap) { h->cap *= 2; h->a = (HItem *)realloc(h->a, (size_t)(h->cap + 1) * sizeof(HItem)); }
    int i = ++h->size;
    h->a[i].slot = slot; h->a[i].key = key;
    while (i > 1) {
        int p = i >> 1;
        if (h->a[p].key <= h->a[i].key) break;
        HItem t = h->a[p]; h->a[p] = h->a[i]; h->a[i] = t;
        i = p;
    }
}
static HItem hp_pop(Heap *h) {
    HItem top = h->a[1];
    h->a[1] = h->a[h->size--];
    int i = 1;
    while (h->size > 0) {
        int m = i, l = 2 * i, r = l + 1;
        if (l <= h->size && h->a[l].key < h->a[m].key) m = l;
        if (r <= h->size && h->a[r].key < h->a[m].key) m = r;
        if (m == i) break;
        HItem t = h->a[i]; h->a[i] = h->a[m]; h->a[m] = t;
        i = m;
    }
    return top;
}

/* ───────────── adjacency ───────────── */

static inline const float *vec_of(const Builder *b, int32_t s) { return b->vecs + (size_t)s * (size_t)b->dim; }

static inline int32_t *list_ptr(const Builder *b, int32_t s, int l, int32_t **cnt) {
    if (l == 0) { *cnt = &b->cnt0[s]; return b->nbr0 + (size_t)s * (size_t)(b->M0 + 1); }
    int64_t u = b->upper_base[s] + (l - 1);
    *cnt = &b->cntU[u];
    return b->nbrU + (size_t)u * (size_t)(b->M + 1);
}

static inline float *dlist_ptr(const Builder *b, int32_t s, int l) {
    if (l == 0) return b->d0 + (size_t)s * (size_t)(b->M0 + 1);
    return b->dU + (size_t)(b->upper_base[s] + (l - 1)) * (size_t)(b->M + 1);
}

static int list_add(int32_t *lst, float *dl, int32_t *cnt, int32_t t, float d) { /* src/hnsw_algo.c:142-163 */
    for (int i = 0; i < *cnt; i++) if (lst[i] == t) return 0;
    lst[*cnt] = t; dl[*cnt] = d;
    (*cnt)++;
    return 1;
}

static int mutual_count(const Builder *b, int32_t x, int32_t y, int l) { /* src/hnsw_algo.c:460-475 */
    if (l > b->level[x] || l > b->level[y]) return 0;
    int32_t *cx, *cy;
    const int32_t *lx = list_ptr(b, x, l, &cx), *ly = list_ptr(b, y, l, &cy);
    int c = 0;
    for (int i = 0; i < *cx; i++)
        for (int j = 0; j < *cy; j++)
            if (ly[j] == lx[i]) { c++; break; }
    return c;
}

/* MN-RU prune of s's over-full list at level l (src/hnsw_algo.c:601-646). The reference recomputes the distance
 * from s to every entry; here each edge keeps the distance it was created with, which is the same number: all
 * three metrics are bit-symmetric in their arguments (a−b and b−a differ in sign only, products commute, cosine's
 * two norms swap places in a commutative product). The reference
 * evaluates the mutual-neighbour count of every entry up front; it is consulted only
 * when two distances compare equal, so it is evaluated on demand here (use_mn = 0 in the
 * batch linker, where other lists are being written concurrently: ties keep the earlier
 * entry there). */
static void prune(Builder *b, int32_t s, int l, int M_max, int use_mn) {
    int32_t *cnt;
    int32_t *lst = list_ptr(b, s, l, &cnt);
    int nc = *cnt;
    float d[2 * 64 + 8];
    int mn[2 * 64 + 8];
    int32_t id[2 * 64 + 8];
    float *dp = d; int *mp = mn; int32_t *ip = id;
    if (nc > (int)(sizeof(d) / sizeof(d[0]))) {
        dp = (float *)malloc((size_t)nc * sizeof(float));
        mp = (int *)malloc((size_t)nc * sizeof(int));
        ip = (int32_t *)malloc((size_t)nc * sizeof(int32_t));
    }
    float *dl = dlist_ptr(b, s, l);
    for (int j = 0; j < nc; j++) { ip[j] = lst[j]; dp[j] = dl[j]; mp[j] = -2; }
    for (int a = 0; a < M_max && a < nc; a++) {
        int best = a;
        for (int c = a + 1; c < nc; c++) {
            if (dp[c] < dp[best]) best = c;
            else if (dp[c] == dp[best] && use_mn) {
                if (mp[c] == -2) mp[c] = mutual_count(b, s, ip[c], l);
                if (mp[best] == -2) mp[best] = mutual_count(b, s, ip[best], l);
                if (mp[c] > mp[best]) best = c;
            }
        }
        if (best != a) {
            float td = dp[a]; dp[a] = dp[best]; dp[best] = td;
            int tm = mp[a]; mp[a] = mp[best]; mp[best] = tm;
            int32_t ti = ip[a]; ip[a] = ip[best]; ip[best] = ti;
        }
    }
    memcpy(lst, ip, (size_t)M_max * sizeof(int32_t));
    memcpy(dl, dp, (size_t)M_max * sizeof(float));
    *cnt = M_max;
    if (dp != d) { free(dp); free(mp); free(ip); }
}

/* ───────────── search primitives over the growing graph ───────────── */

static int32_t greedy(const Builder *b, const float *q, int32_t entry, int l) {
    int32_t cur = entry;
    float cur_d = dist(b, q, vec_of(b, cur));
    int changed = 1;
    while (changed) {
        changed = 0;
        int32_t *cnt;
        const int32_t *lst = list_ptr(b, cur, l, &cnt);
        for (int i = 0; i < *cnt; i++) {
            int32_t t = lst[i];
            float d = dist(b, q, vec_of(b, t));
            if (d < cur_d) { cur_d = d; cur = t; changed = 1; lst = list_ptr(b, cur, l, &cnt); }
        }
    }
    return cur;
}

static int beam(const Builder *b, Scratch *sc, const float *q, int32_t entry, int l, int ef) {
    hp_reset(&sc->cand, ef * 2);
    hp_reset(&sc->res, ef * 2);
    if (++sc->cur == 0) { memset(sc->stamp, 0, (size_t)b->cap * sizeof(uint32_t)); sc->cur = 1; }
    float d0 = dist(b, q, vec_of(b, entry));
    hp_push(&sc->cand, entry, d0);
    hp_push(&sc->res, entry, -d0);
    sc->stamp[entry] = sc->cur;
    int patience_max = ef / 4 < 10 ? 10 : ef / 4, stale = 0;
    const int row_bytes = b->dim * (int)sizeof(float);
    while (sc->cand.size > 0) {
        HItem c = hp_pop(&sc->cand);
        if (sc->res.size >= ef && c.key > -sc->res.a[1].key) break;
        if (stale >= patience_max && sc->res.size >= ef) break;
        int32_t *cnt;
        const int32_t *lst = list_ptr(b, c.slot, l, &cnt);
        int n = *cnt, improved = 0, nf = 0;
        int32_t *fresh = sc->fresh;
        float *fd = sc->fresh_d;
        /* a list holds a slot at most once (list_add), so marking the unvisited entries first and evaluating them
         * afterwards — two rows per call — visits the same nodes in the same order as the reference's single loop */
        for (int i = 0; i < n; i++) {
            int32_t t = lst[i];
            if (sc->stamp[t] == sc->cur) continue;
            sc->stamp[t] = sc->cur;
            { const char *pv = (const char *)vec_of(b, t); for (int o = 0; o < row_bytes; o += 64) __builtin_prefetch(pv + o); }
            fresh[nf++] = t;
        }
        for (int i = 0; i + 1 < nf; i += 2) dist_pair(b, q, vec_of(b, fresh[i]), vec_of(b, fresh[i + 1]), &fd[i], &fd[i + 1]);
        if (nf & 1) fd[nf - 1] = dist(b, q, vec_of(b, fresh[nf - 1]));
        for (int i = 0; i < nf; i++) {
            int32_t t = fresh[i];
            float d = fd[i];
            if (sc->res.size < ef) { hp_push(&sc->cand, t, d); hp_push(&sc->res, t, -d); improved = 1; }
            else if (d < -sc->res.a[1].key) { hp_push(&sc->cand, t, d); hp_pop(&sc->res); hp_push(&sc->res, t, -d); improved = 1; }
        }
        stale = improved ? 0 : stale + 1;
    }
    int count = sc->res.size;
    for (int i = count - 1; i >= 0; i--) { HItem it = hp_pop(&sc->res); sc->r_slot[i] = it.slot; sc->r_dist[i] = -it.key; }
    return count;
}

/* ───────────── level generator (src/hnsw_algo.c:19-30, :240-248) ───────────── */

static int next_level(Builder *b) {
    uint32_t x = b->rng;
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    b->rng = x;
    double r = (double)x / (double)0xFFFFFFFFu;
    if (r == 0.0) r = 1e-10;
    int lv = (int)(-log(r) * b->level_mult);
    if (lv >= MB_MAX_LEVELS) lv = MB_MAX_LEVELS - 1;
    return lv;
}

static void alloc_node(Builder *b, int32_t s, int lv) {
    b->level[s] = lv;
    b->cnt0[s] = 0;
    b->upper_base[s] = -1;
    if (lv > 0) {
        if (b->n_upper + lv > b->cap_upper) {
            int64_t nc = b->cap_upper * 2 + lv + 1024;
            b->nbrU = (int32_t *)realloc(b->nbrU, (size_t)nc * (size_t)(b->M + 1) * sizeof(int32_t));
            b->cntU = (int32_t *)realloc(b->cntU, (size_t)nc * sizeof(int32_t));
            b->dU = (float *)realloc(b->dU, (size_t)nc * (size_t)(b->M + 1) * sizeof(float));
            b->cap_upper = nc;
        }
        b->upper_base[s] = b->n_upper;
        for (int l = 0; l < lv; l++) b->cntU[b->n_upper + l] = 0;
        b->n_upper += lv;
    }
}

/* Batch mode orders bit-equal distances by slot. The reference's order inside such a group is whatever its
 * binary heap yields (unspecified, test/test_priority_queue.c:85-99); the exact mode keeps that order, the
 * batch mode fixes it so that a searcher with the device's (distance, slot) rule selects the same neighbours. */
static void order_ties_by_slot(int32_t *slot, const float *d, int n) {
    for (int i = 0; i + 1 < n;) {
        int j = i + 1;
        while (j < n && d[j] == d[i]) j++;
        for (int a = i + 1; a < j; a++) { /* insertion sort of the run [i, j) */
            int32_t v = slot[a];
            int b = a - 1;
            while (b >= i && slot[b] > v) { slot[b + 1] = slot[b]; b--; }
            slot[b + 1] = v;
        }
        i = j;
    }
}

/* phase A of an insert: find the per-level selected neighbours of node s (read-only on
 * every other node). Levels searched: min(level, max_level) .. 0 (src/hnsw_algo.c:553-579). */
static void find_links(const Builder *b, Scratch *sc, int32_t s) {
    const float *q = vec_of(b, s);
    int lv = b->level[s];
    int32_t cur = b->entry;
    for (int l = b->max_level; l > lv; l--) cur = greedy(b, q, cur, l);
    int start = lv < b->max_level ? lv : b->max_level;
    for (int l = 0; l < MB_MAX_LEVELS; l++) sc->sel_cnt[l] = 0;
    for (int l = start; l >= 0; l--) {
        int M_max = l == 0 ? b->M0 : b->M;
        int found = beam(b, sc, q, cur, l, b->efc);
        order_ties_by_slot(sc->r_slot, sc->r_dist, found);
        int k = found < M_max ? found : M_max;
        memcpy(sc->sel + (size_t)l * (size_t)b->M0, sc->r_slot, (size_t)k * sizeof(int32_t));
        memcpy(sc->sel_d + (size_t)l * (size_t)b->M0, sc->r_dist, (size_t)k * sizeof(float));
        sc->sel_cnt[l] = k;
        if (found > 0) cur = sc->r_slot[0];
    }
}

/* exact sequential insert (src/hnsw_algo.c:520-666). The search of level l-1 must see the
 * links made at level l, so search and link are interleaved exactly as in the reference. */
static void insert_exact(Builder *b, int32_t s, int lv) {
    Scratch *sc = &b->scratch[0];
    alloc_node(b, s, lv);
    if (b->entry < 0) { b->entry = s; b->max_level = lv; return; }
    const float *q = vec_of(b, s);
    int32_t cur = b->entry;
    for (int l = b->max_level; l > lv; l--) cur = greedy(b, q, cur, l);
    int start = lv < b->max_level ? lv : b->max_level;
    for (int l = start; l >= 0; l--) {
        int M_max = l == 0 ? b->M0 : b->M;
        int found = beam(b, sc, q, cur, l, b->efc);
        int k = found < M_max ? found : M_max;
        int32_t *mycnt;
        int32_t *mine = list_ptr(b, s, l, &mycnt);
        float *mine_d = dlist_ptr(b, s, l);
        for (int i = 0; i < k; i++) {
            int32_t t = sc->r_slot[i];
            list_add(mine, mine_d, mycnt, t, sc->r_dist[i]);
            if (l <= b->level[t]) {
                int32_t *tc;
                int32_t *tl = list_ptr(b, t, l, &tc);
                list_add(tl, dlist_ptr(b, t, l), tc, s, sc->r_dist[i]);
                if (*tc > M_max) prune(b, t, l, M_max, 1);
            }
        }
        if (found > 0) cur = sc->r_slot[0];
    }
    if (lv > b->max_level) { b->entry = s; b->max_level = lv; }
}

typedef struct { int32_t target, level, src; float d; } RevReq;

static int cmp_req(const void *a, const void *b) {
    const RevReq *x = (const RevReq *)a, *y = (const RevReq *)b;
    if (x->target != y->target) return x->target < y->target ? -1 : 1;
    if (x->level != y->level) return x->level < y->level ? -1 : 1;
    return x->src < y->src ? -1 : x->src > y->src;
}

/* Order by (target, level, src) — cmp_req — as an LSD radix sort on the packed key target·2^26 | level·2^21 |
 * (src − g0) when the group is small enough for 21 bits (a serial qsort of the ≈ 33 requests per node was 8 % of a
 * 1 024-node group's time); otherwise qsort. */
static void sort_requests(RevReq *req, int64_t tot, int32_t g0) {
    typedef struct { uint64_t key; float d; } KD;
    int ok = tot > 256;
    for (int64_t i = 0; ok && i < tot; i++)
        if (req[i].src - g0 >= (1 << 21) || req[i].src < g0) ok = 0;
    KD *key = ok ? (KD *)malloc((size_t)tot * 2 * sizeof(KD)) : NULL;
    if (!key) { qsort(req, (size_t)tot, sizeof(RevReq), cmp_req); return; }
    KD *base = key, *tmp = key + tot;
    uint64_t top = 0;
    for (int64_t i = 0; i < tot; i++) {
        key[i].key = ((uint64_t)(uint32_t)req[i].target << 26) | ((uint64_t)req[i].level << 21) | (uint64_t)(req[i].src - g0);
        key[i].d = req[i].d;
        if (key[i].key > top) top = key[i].key;
    }
    enum { BITS = 11 };
    size_t *cnt = (size_t *)malloc(((size_t)1 << BITS) * sizeof(size_t));
    for (int shift = 0; shift < 64 && (top >> shift) != 0; shift += BITS) {
        memset(cnt, 0, ((size_t)1 << BITS) * sizeof(size_t));
        for (int64_t i = 0; i < tot; i++) cnt[(key[i].key >> shift) & ((1u << BITS) - 1)]++;
        size_t run = 0;
        for (size_t d = 0; d < ((size_t)1 << BITS); d++) { size_t c = cnt[d]; cnt[d] = run; run += c; }
        for (int64_t i = 0; i < tot; i++) tmp[cnt[(key[i].key >> shift) & ((1u << BITS) - 1)]++] = key[i];
        KD *sw = key; key = tmp; tmp = sw;
    }
    for (int64_t i = 0; i < tot; i++) {
        req[i].target = (int32_t)(key[i].key >> 26);
        req[i].level = (int32_t)((key[i].key >> 21) & 31);
        req[i].src = g0 + (int32_t)(key[i].key & ((1u << 21) - 1));
        req[i].d = key[i].d;
    }
    free(cnt);
    free(base);
}

/* ───────────── public C API ───────────── */

static void scratch_init(Builder *b, Scratch *sc) {
    memset(sc, 0, sizeof(*sc));
    sc->stamp = (uint32_t *)calloc((size_t)b->cap, sizeof(uint32_t));
    sc->r_slot = (int32_t *)malloc((size_t)b->efc * sizeof(int32_t));
    sc->r_dist = (float *)malloc((size_t)b->efc * sizeof(float));
    sc->sel = (int32_t *)malloc((size_t)MB_MAX_LEVELS * (size_t)b->M0 * sizeof(int32_t));
    sc->sel_d = (float *)malloc((size_t)MB_MAX_LEVELS * (size_t)b->M0 * sizeof(float));
    sc->sel_cnt = (int32_t *)calloc(MB_MAX_LEVELS, sizeof(int32_t));
    sc->fresh = (int32_t *)malloc((size_t)(b->M0 + 1) * sizeof(int32_t));
    sc->fresh_d = (float *)malloc((size_t)(b->M0 + 1) * sizeof(float));
}

void *muninn_build_create(int dim, int metric, int M, int ef_construction, int64_t capacity, const float *vectors) {
    Builder *b = (Builder *)calloc(1, sizeof(Builder));
    if (!b) return NULL;
    if (g_avx2 < 0) muninn_build_set_avx2(-1);
    b->dim = dim; b->metric = metric; b->M = M; b->M0 = 2 * M; /* src/hnsw_algo.c:188 */
    b->efc = ef_construction; b->cap = capacity; b->vecs = vectors;
    b->level = (int32_t *)malloc((size_t)capacity * sizeof(int32_t));
    b->nbr0 = (int32_t *)malloc((size_t)capacity * (size_t)(b->M0 + 1) * sizeof(int32_t));
    b->cnt0 = (int32_t *)calloc((size_t)capacity, sizeof(int32_t));
    b->d0 = (float *)malloc((size_t)capacity * (size_t)(b->M0 + 1) * sizeof(float));
    b->upper_base = (int64_t *)malloc((size_t)capacity * sizeof(int64_t));
    b->entry = -1; b->max_level = -1;
    b->pending_g0 = b->pending_g1 = -1;
    b->rng = 42; /* src/hnsw_algo.c:205 */
    b->level_mult = 1.0 / log((double)M);
    if (!b->level || !b->nbr0 || !b->cnt0 || !b->d0 || !b->upper_base) return NULL;
    return b;
}

/* Two-rows-per-call AVX2 distances on (1) / off (0) / as detected (-1); returns what is in effect. The graphs do not
 * depend on it (tests/test_host.py); MUNINN_BUILD_NO_AVX2 in the environment also turns it off. */
int muninn_build_set_avx2(int on) {
    if (on < 0) g_avx2 = getenv("MUNINN_BUILD_NO_AVX2") ? 0 : (__builtin_cpu_supports("avx2") ? 1 : 0);
    else g_avx2 = on && __builtin_cpu_supports("avx2");
    return g_avx2;
}

void muninn_build_seed(void *h, uint32_t seed) { ((Builder *)h)->rng = seed ? seed : 1; } /* src/hnsw_algo.c:222-224 */

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* ───────────── touched-node log (what a device copy of the graph must be patched with) ───────────── */

static void touch(Builder *b, int32_t s) {
    if (!b->log_touched || b->touch_stamp[s] == b->touch_epoch) return;
    b->touch_stamp[s] = b->touch_epoch;
    if (b->n_touched == b->cap_touched) {
        b->cap_touched = b->cap_touched * 2 + 4096;
        b->touched = (int32_t *)realloc(b->touched, (size_t)b->cap_touched * sizeof(int32_t));
    }
    b->touched[b->n_touched++] = s;
}

/* insert_exact changes the new node's lists and the lists of the nodes it links itself to
 * (src/hnsw_algo.c:582-646: a prune rewrites only the over-full list itself). */
static void insert_exact_logged(Builder *b, int32_t s, int lv) {
    insert_exact(b, s, lv);
    if (!b->log_touched) return;
    touch(b, s);
    for (int l = 0; l <= lv; l++) {
        int32_t *cnt;
        const int32_t *lst = list_ptr(b, s, l, &cnt);
        for (int i = 0; i < *cnt; i++) touch(b, lst[i]);
    }
}

static int ensure_scratch(Builder *b, int nthreads) {
    if (b->n_scratch < nthreads) {
        b->scratch = (Scratch *)realloc(b->scratch, (size_t)nthreads * sizeof(Scratch));
        for (int t = b->n_scratch; t < nthreads; t++) scratch_init(b, &b->scratch[t]);
        b->n_scratch = nthreads;
    }
    return 0;
}

/* Draw levels for the next group of at most `batch` nodes (slots b->n ..) and allocate them. A node
 * that would raise max_level closes the group before it and is left allocated as the pending node
 * (*raise = 1); when it is the first node drawn it is inserted alone, exactly, and the draw goes on. */
static void form_group(Builder *b, int64_t end, int batch, int64_t *g0_out, int64_t *g1_out, int *raise_out) {
    int64_t g0 = b->n, g1 = g0;
    int raise = 0;
    while (g1 < end && g1 - g0 < batch) {
        int lv = next_level(b);
        if (lv > b->max_level) {
            if (g1 == g0) { insert_exact_logged(b, (int32_t)g1, lv); b->n++; g0 = g1 = b->n; continue; }
            alloc_node(b, (int32_t)g1, lv);
            raise = 1;
            break;
        }
        alloc_node(b, (int32_t)g1, lv);
        g1++;
    }
    *g0_out = g0; *g1_out = g1; *raise_out = raise;
}

/* Link the group [g0, g1) into the graph as it stood before the group. cand == NULL: every node
 * searches on the host (find_links). cand != NULL: node g0+i with cand_cnt[i] >= 0 takes the first
 * min(cand_cnt[i], M0) slots of cand[i*cand_stride ..] as its layer-0 selection — the caller's
 * (device) search of the same graph with ef = ef_construction; only level-0 nodes may be supplied. */
static void link_group(Builder *b, int64_t g0, int64_t g1, int raise, int nthreads, const int32_t *cand,
                       const int32_t *cand_cnt, int cand_stride) {
    int64_t gn = g1 - g0;
    if (gn > 0) {
        /* phase A: parallel read-only searches; each node writes only its own lists */
        int64_t per = (int64_t)b->M0 + (int64_t)b->M * MB_MAX_LEVELS;
        if (gn * per > b->req_cap) {
            free(b->req_buf);
            b->req_buf = malloc((size_t)(gn * per) * sizeof(RevReq));
            b->req_cap = gn * per;
        }
        RevReq *req = (RevReq *)b->req_buf;
        int32_t *req_n = (int32_t *)calloc((size_t)gn, sizeof(int32_t));
        double t0 = now_s();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
        for (int64_t i = 0; i < gn; i++) {
#ifdef _OPENMP
            Scratch *sc = &b->scratch[omp_get_thread_num()];
#else
            Scratch *sc = &b->scratch[0];
#endif
            int32_t s = (int32_t)(g0 + i);
            if (cand && cand_cnt[i] >= 0 && b->level[s] == 0) {
                int k = cand_cnt[i] < b->M0 ? cand_cnt[i] : b->M0;
                for (int l = 0; l < MB_MAX_LEVELS; l++) sc->sel_cnt[l] = 0;
                memcpy(sc->sel, cand + (size_t)i * (size_t)cand_stride, (size_t)k * sizeof(int32_t));
                for (int j = 0; j < k; j++) sc->sel_d[j] = dist(b, vec_of(b, s), vec_of(b, sc->sel[j]));
                sc->sel_cnt[0] = k;
            } else {
                find_links(b, sc, s);
            }
            RevReq *r = req + i * per;
            int nr = 0;
            for (int l = b->level[s] < b->max_level ? b->level[s] : b->max_level; l >= 0; l--) {
                int32_t *mycnt;
                int32_t *mine = list_ptr(b, s, l, &mycnt);
                float *mine_d = dlist_ptr(b, s, l);
                for (int j = 0; j < sc->sel_cnt[l]; j++) {
                    int32_t t = sc->sel[(size_t)l * (size_t)b->M0 + j];
                    float d = sc->sel_d[(size_t)l * (size_t)b->M0 + j];
                    list_add(mine, mine_d, mycnt, t, d);
                    r[nr++] = (RevReq){t, l, s, d};
                }
            }
            req_n[i] = nr;
        }
        /* compact + sort requests by (target, level, src) */
        double t1 = now_s();
        int64_t tot = 0;
        for (int64_t i = 0; i < gn; i++) {
            if (tot != i * per) memmove(req + tot, req + i * per, (size_t)req_n[i] * sizeof(RevReq));
            tot += req_n[i];
        }
        free(req_n);
        sort_requests(req, tot, (int32_t)g0);
        /* group boundaries per target */
        if (tot + 1 > b->grp_cap) {
            free(b->grp_buf);
            b->grp_buf = (int64_t *)malloc((size_t)(tot + 1) * sizeof(int64_t));
            b->grp_cap = tot + 1;
        }
        int64_t *grp = b->grp_buf;
        int64_t ng = 0;
        for (int64_t i = 0; i < tot; i++)
            if (i == 0 || req[i].target != req[i - 1].target) { grp[ng++] = i; touch(b, req[i].target); }
        grp[ng] = tot;
        for (int64_t s = g0; s < g1; s++) touch(b, (int32_t)s);
        /* phase B: one writer per target list */
        double t2 = now_s();
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads)
        for (int64_t gi = 0; gi < ng; gi++) {
            for (int64_t i = grp[gi]; i < grp[gi + 1]; i++) {
                int32_t t = req[i].target; int l = req[i].level;
                int M_max = l == 0 ? b->M0 : b->M;
                int32_t *tc;
                int32_t *tl = list_ptr(b, t, l, &tc);
                list_add(tl, dlist_ptr(b, t, l), tc, req[i].src, req[i].d);
                if (*tc > M_max) prune(b, t, l, M_max, 0);
            }
        }
        b->n = g1;
        b->t_phase[0] += t1 - t0; b->t_phase[1] += t2 - t1; b->t_phase[2] += now_s() - t2;
    }
    if (raise) {
        /* the pending node: level already drawn and allocated; insert it exactly */
        int32_t s = (int32_t)b->n;
        int lv = b->level[s];
        /* undo alloc so insert_exact can redo it identically */
        if (lv > 0) b->n_upper -= lv;
        insert_exact_logged(b, s, lv);
        b->n++;
    }
}

/* Insert the next `count` vectors (slots n .. n+count-1, in order).
 * batch = 0 → exact sequential mode; batch > 0 → batch-synchronous mode once the graph
 * holds at least `warm` nodes (earlier nodes are inserted exactly). */
int muninn_build_insert(void *h, int64_t count, int batch, int64_t warm, int nthreads) {
    Builder *b = (Builder *)h;
    if (b->n + count > b->cap || b->pending_g1 >= 0) return -1;
    if (nthreads < 1) nthreads = 1;
    ensure_scratch(b, nthreads);
    int64_t end = b->n + count;
    while (b->n < end) {
        if (batch <= 0 || b->n < warm) {
            int lv = next_level(b);
            insert_exact_logged(b, (int32_t)b->n, lv);
            b->n++;
            continue;
        }
        int64_t g0, g1; int raise;
        form_group(b, end, batch, &g0, &g1, &raise);
        link_group(b, g0, g1, raise, nthreads, NULL, NULL, 0);
    }
    return 0;
}

/* ───────────── group-at-a-time interface for a caller that runs the layer-0 searches itself ─────────────
 *
 * The batch mode's phase A for a level-0 node is: greedy descent max_level..1, ef_construction-wide beam on
 * layer 0, keep the closest min(found, 2M) (find_links; src/hnsw_algo.c:553-579 for a node of level 0) —
 * which is hnsw_search(k = min(2M, efc), ef = efc) on the graph as it stood before the group. A caller that
 * keeps a device copy of the graph (muninn_cuda_index_upload / _patch) can therefore run those searches as
 * one muninn_cuda_search_batch and hand the results in:
 *
 *   muninn_build_log_touched(h, 1)
 *   loop:  muninn_build_group_begin(h, batch, &g0, &gn)        levels drawn, nodes allocated
 *          [patch the device copy if muninn_build_touched_count(h) > 0: a level-raising first node was
 *           inserted exactly inside group_begin]
 *          search the level-0 nodes of [g0, g0+gn) on the device
 *          muninn_build_group_link(h, cand, cand_cnt, stride, threads)   nodes with cand_cnt < 0 (and all
 *                                                                       nodes of level > 0) search on the host
 *          muninn_build_touched_take → muninn_build_export_nodes → muninn_cuda_index_patch
 *
 * Without bit-equal candidate distances the graph equals muninn_build_insert's batch mode bit for bit
 * (tests/test_device_build.py); with ties the device's (distance, slot) order replaces the heap order. */

void muninn_build_log_touched(void *h, int on) {
    Builder *b = (Builder *)h;
    if (on && !b->touch_stamp) b->touch_stamp = (uint32_t *)calloc((size_t)b->cap, sizeof(uint32_t));
    b->log_touched = on && b->touch_stamp;
    b->n_touched = 0;
    if (++b->touch_epoch == 0) { memset(b->touch_stamp, 0, (size_t)b->cap * sizeof(uint32_t)); b->touch_epoch = 1; }
}

int64_t muninn_build_count(void *h) { return ((Builder *)h)->n; }

/* Returns 0 and the group [*g0, *g0 + *gn); *gn may be 0 when only exact inserts happened (or nothing is
 * left). -1 when a group is already open. */
int muninn_build_group_begin(void *h, int batch, int64_t *g0, int64_t *gn) {
    Builder *b = (Builder *)h;
    if (b->pending_g1 >= 0 || batch < 1) return -1;
    if (b->n >= b->cap) { *g0 = b->n; *gn = 0; return 0; }
    int64_t a, z; int raise;
    form_group(b, b->cap, batch, &a, &z, &raise);
    b->pending_g0 = a; b->pending_g1 = z; b->pending_raise = raise;
    *g0 = a; *gn = z - a;
    return 0;
}

/* levels of the open group's nodes (known after group_begin) */
int muninn_build_group_levels(void *h, int32_t *levels) {
    Builder *b = (Builder *)h;
    if (b->pending_g1 < 0) return -1;
    for (int64_t s = b->pending_g0; s < b->pending_g1; s++) levels[s - b->pending_g0] = b->level[s];
    return 0;
}

int muninn_build_group_link(void *h, const int32_t *cand, const int32_t *cand_cnt, int cand_stride, int nthreads) {
    Builder *b = (Builder *)h;
    if (b->pending_g1 < 0) return -1;
    if (nthreads < 1) nthreads = 1;
    ensure_scratch(b, nthreads);
    if (cand) { /* slots must be stored nodes from before the group */
        int64_t gn = b->pending_g1 - b->pending_g0;
        for (int64_t i = 0; i < gn; i++) {
            if (cand_cnt[i] < 0) continue;
            if (cand_cnt[i] > cand_stride) return -2;
            for (int j = 0; j < cand_cnt[i] && j < b->M0; j++) {
                int32_t t = cand[(size_t)i * (size_t)cand_stride + j];
                if (t < 0 || t >= b->pending_g0) return -2;
            }
        }
    }
    link_group(b, b->pending_g0, b->pending_g1, b->pending_raise, nthreads, cand, cand_cnt, cand_stride);
    b->pending_g0 = b->pending_g1 = -1; b->pending_raise = 0;
    return 0;
}

int64_t muninn_build_touched_count(void *h) { return ((Builder *)h)->n_touched; }

/* Copy the log (each touched slot once, in no particular order) and clear it. */
int64_t muninn_build_touched_take(void *h, int32_t *out, int64_t cap) {
    Builder *b = (Builder *)h;
    int64_t n = b->n_touched < cap ? b->n_touched : cap;
    if (out && n > 0) memcpy(out, b->touched, (size_t)n * sizeof(int32_t));
    b->n_touched = 0;
    if (b->touch_stamp && ++b->touch_epoch == 0) { memset(b->touch_stamp, 0, (size_t)b->cap * sizeof(uint32_t)); b->touch_epoch = 1; }
    return n;
}

/* Records of the given nodes in the list-CSR layout of muninn_cuda_index_patch: neighbour lists in the
 * builder's own (insertion) order. Pass targets = NULL to obtain the sizes only. */
void muninn_build_export_nodes(void *h, int64_t n_nodes, const int32_t *slots, int32_t *levels, int64_t *list_offsets,
                               int32_t *targets, int64_t *n_lists, int64_t *n_edges) {
    Builder *b = (Builder *)h;
    int64_t li = 0, e = 0;
    if (targets) list_offsets[0] = 0;
    for (int64_t i = 0; i < n_nodes; i++) {
        int32_t s = slots[i];
        if (targets) levels[i] = b->level[s];
        for (int l = 0; l <= b->level[s]; l++) {
            int32_t *cnt;
            const int32_t *lst = list_ptr(b, s, l, &cnt);
            if (targets) { memcpy(targets + e, lst, (size_t)*cnt * sizeof(int32_t)); list_offsets[li + 1] = e + *cnt; }
            e += *cnt;
            li++;
        }
    }
    *n_lists = li; *n_edges = e;
}

/* Seconds spent so far in the three phases of the group linker: [0] candidate searches + own lists, [1] sorting the
 * reverse-edge requests, [2] reverse edges + MN-RU prune. */
void muninn_build_phase_seconds(void *h, double out[3]) {
    Builder *b = (Builder *)h;
    for (int i = 0; i < 3; i++) out[i] = b->t_phase[i];
}

void muninn_build_entry(void *h, int32_t *entry, int32_t *max_level) {
    Builder *b = (Builder *)h;
    *entry = b->entry; *max_level = b->max_level;
}

void muninn_build_sizes(void *h, int64_t *n, int64_t *n_lists, int64_t *n_edges, int32_t *entry, int32_t *max_level) {
    Builder *b = (Builder *)h;
    int64_t nl = 0, ne = 0;
    for (int64_t s = 0; s < b->n; s++) {
        nl += b->level[s] + 1;
        ne += b->cnt0[s];
        for (int l = 1; l <= b->level[s]; l++) ne += b->cntU[b->upper_base[s] + l - 1];
    }
    *n = b->n; *n_lists = nl; *n_edges = ne; *entry = b->entry; *max_level = b->max_level;
}

static int cmp_i32(const void *a, const void *b) {
    int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
    return (x > y) - (x < y);
}

/* Export in the list-CSR layout of include/muninn_cuda.h. sort_lists = 1 gives the
 * "reloaded from shadow tables" neighbour order (ascending target id,
 * src/hnsw_vtab.c:163-169 + :322), which is the state the parity contract pins. */
void muninn_build_export(void *h, int32_t *levels, int64_t *list_offsets, int32_t *targets, int sort_lists) {
    Builder *b = (Builder *)h;
    int64_t li = 0, e = 0;
    list_offsets[0] = 0;
    for (int64_t s = 0; s < b->n; s++) {
        levels[s] = b->level[s];
        for (int l = 0; l <= b->level[s]; l++) {
            int32_t *cnt;
            const int32_t *lst = list_ptr(b, (int32_t)s, l, &cnt);
            memcpy(targets + e, lst, (size_t)*cnt * sizeof(int32_t));
            if (sort_lists && *cnt > 1) qsort(targets + e, (size_t)*cnt, sizeof(int32_t), cmp_i32);
            e += *cnt;
            list_offsets[++li] = e;
        }
    }
}

/* ───────────── the group loop (muninn_build.h: muninn_build_run) ───────────── */

typedef struct {
    int32_t *slots, *levels, *targets;
    int64_t *ids, *offs;
    int64_t cap_slots, cap_lists, cap_edges;
} FlushBuf;

/* device copy ← every record the host changed since the last flush; *n_dev = rows the device copy holds */
static int run_flush(Builder *b, const int64_t *ids, const MuninnBuildDevice *dev, FlushBuf *f, int64_t *n_dev,
                     MuninnBuildRunStats *st) {
    const int64_t n_now = b->n, n_new = n_now - *n_dev;
    if (b->n_touched == 0 && n_new == 0) return 0;
    double t0 = now_s();
    const int64_t need = n_new + b->n_touched;
    if (need > f->cap_slots) {
        f->cap_slots = need + need / 2 + 1024;
        f->slots = (int32_t *)realloc(f->slots, (size_t)f->cap_slots * sizeof(int32_t));
        f->levels = (int32_t *)realloc(f->levels, (size_t)f->cap_slots * sizeof(int32_t));
        f->ids = (int64_t *)realloc(f->ids, (size_t)f->cap_slots * sizeof(int64_t));
        if (!f->slots || !f->levels || !f->ids) return -1;
    }
    int64_t ns = 0;
    for (int64_t s = *n_dev; s < n_now; s++) f->slots[ns++] = (int32_t)s; /* new nodes first */
    for (int64_t i = 0; i < b->n_touched; i++)
        if (b->touched[i] < *n_dev) f->slots[ns++] = b->touched[i];
    muninn_build_touched_take(b, NULL, 0); /* clears the log */
    int64_t nl = 0, ne = 0;
    muninn_build_export_nodes(b, ns, f->slots, NULL, NULL, NULL, &nl, &ne);
    if (nl + 1 > f->cap_lists) {
        f->cap_lists = nl + nl / 2 + 1024;
        f->offs = (int64_t *)realloc(f->offs, (size_t)f->cap_lists * sizeof(int64_t));
        if (!f->offs) return -1;
    }
    if (ne + 1 > f->cap_edges) {
        f->cap_edges = ne + ne / 2 + 1024;
        f->targets = (int32_t *)realloc(f->targets, (size_t)f->cap_edges * sizeof(int32_t));
        if (!f->targets) return -1;
    }
    muninn_build_export_nodes(b, ns, f->slots, f->levels, f->offs, f->targets, &nl, &ne);
    for (int64_t i = 0; i < ns; i++) f->ids[i] = ids ? ids[f->slots[i]] : (int64_t)f->slots[i];
    const float *vec = n_new > 0 ? b->vecs + (size_t)*n_dev * (size_t)b->dim : b->vecs;
    int rc = dev->patch(dev->ctx, n_now, (int32_t)ns, f->slots, f->ids, f->levels, vec, f->offs, f->targets, b->entry,
                        b->max_level);
    if (rc) return rc;
    *n_dev = n_now;
    if (st) { st->patches++; st->patched_records += ns; st->patch_s += now_s() - t0; }
    return 0;
}

int muninn_build_run(void *h, const int64_t *ids, int batch, int64_t warm, int nthreads, const MuninnBuildDevice *dev,
                     MuninnBuildRunStats *st) {
    Builder *b = (Builder *)h;
    if (!b || !dev || !dev->upload || !dev->patch || !dev->search || batch < 1 || b->n != 0 || b->pending_g1 >= 0) return -1;
    if (st) memset(st, 0, sizeof(*st));
    const int64_t n = b->cap;
    if (n == 0) return 0;
    const int k = b->M0 < b->efc ? b->M0 : b->efc;
    int rc = -1;
    FlushBuf f;
    memset(&f, 0, sizeof f);
    int32_t *glev = NULL, *cand = NULL, *cand_cnt = NULL, *where = NULL, *cnt = NULL;
    int64_t *found = NULL;
    float *q = NULL, *fdist = NULL;
    int64_t *ident = NULL;

    /* the first `warm` nodes: exact sequential inserts on the host, then one full upload */
    double t0 = now_s();
    const int64_t n_warm = n < (warm > 1 ? warm : 1) ? n : (warm > 1 ? warm : 1);
    if (muninn_build_insert(b, n_warm, 0, 0, 1) != 0) goto done;
    {
        int64_t nn, nl, ne; int32_t es, ml;
        muninn_build_sizes(b, &nn, &nl, &ne, &es, &ml);
        int32_t *lv = (int32_t *)malloc((size_t)nn * sizeof(int32_t));
        int64_t *offs = (int64_t *)malloc((size_t)(nl + 1) * sizeof(int64_t));
        int32_t *tg = (int32_t *)malloc((size_t)(ne + 1) * sizeof(int32_t));
        if (!ids) {
            ident = (int64_t *)malloc((size_t)n_warm * sizeof(int64_t));
            for (int64_t i = 0; ident && i < n_warm; i++) ident[i] = i;
        }
        if (lv && offs && tg && (ids || ident)) {
            muninn_build_export(b, lv, offs, tg, 0);
            rc = dev->upload(dev->ctx, nn, ids ? ids : ident, lv, b->vecs, offs, tg, es, ml);
        }
        free(lv); free(offs); free(tg);
        if (rc) goto done;
        rc = -1;
    }
    if (st) st->warm_s = now_s() - t0;
    muninn_build_log_touched(b, 1);
    if (!b->log_touched) goto done;
    int64_t n_dev = n_warm;
    glev = (int32_t *)malloc((size_t)batch * sizeof(int32_t));
    where = (int32_t *)malloc((size_t)batch * sizeof(int32_t));
    cnt = (int32_t *)malloc((size_t)batch * sizeof(int32_t));
    cand_cnt = (int32_t *)malloc((size_t)batch * sizeof(int32_t));
    cand = (int32_t *)malloc((size_t)batch * (size_t)k * sizeof(int32_t));
    found = (int64_t *)malloc((size_t)batch * (size_t)k * sizeof(int64_t));
    fdist = (float *)malloc((size_t)batch * (size_t)k * sizeof(float));
    q = (float *)malloc((size_t)batch * (size_t)b->dim * sizeof(float));
    if (!glev || !where || !cnt || !cand_cnt || !cand || !found || !fdist || !q) goto done;

    while (b->n < n) {
        int64_t g0, gn;
        if (muninn_build_group_begin(b, batch, &g0, &gn) != 0) goto done;
        if ((rc = run_flush(b, ids, dev, &f, &n_dev, st)) != 0) goto done; /* a level-raising first node was inserted exactly */
        rc = -1;
        if (gn == 0) continue;
        muninn_build_group_levels(b, glev);
        int nq = 0;
        for (int64_t i = 0; i < gn; i++) {
            cand_cnt[i] = -1;
            if (glev[i] == 0) {
                memcpy(q + (size_t)nq * (size_t)b->dim, vec_of(b, (int32_t)(g0 + i)), (size_t)b->dim * sizeof(float));
                where[nq++] = (int32_t)i;
            }
        }
        t0 = now_s();
        if (nq > 0) {
            if ((rc = dev->search(dev->ctx, q, nq, k, b->efc, found, fdist, cnt)) != 0) goto done;
            rc = -1;
            for (int j = 0; j < nq; j++) {
                int32_t *row = cand + (size_t)where[j] * (size_t)k;
                int c = cnt[j] < 0 ? 0 : (cnt[j] > k ? k : cnt[j]);
                for (int r = 0; r < c; r++) {
                    const int64_t id = found[(size_t)j * (size_t)k + r];
                    int64_t sl = id;
                    if (ids) { /* rowid → slot among the nodes stored before the group */
                        int64_t lo = 0, hi = g0 - 1;
                        sl = -1;
                        while (lo <= hi) {
                            int64_t mid = lo + (hi - lo) / 2;
                            if (ids[mid] == id) { sl = mid; break; }
                            if (ids[mid] < id) lo = mid + 1; else hi = mid - 1;
                        }
                    }
                    row[r] = (int32_t)sl; /* out-of-graph values are refused by muninn_build_group_link */
                }
                cand_cnt[where[j]] = c;
            }
        }
        if (st) st->search_s += now_s() - t0;
        t0 = now_s();
        if (muninn_build_group_link(b, cand, cand_cnt, k, nthreads) != 0) goto done;
        if (st) { st->link_s += now_s() - t0; st->groups++; st->device_queries += nq; st->host_searches += gn - nq; }
        if ((rc = run_flush(b, ids, dev, &f, &n_dev, st)) != 0) goto done;
        rc = -1;
    }
    rc = 0;
done:
    free(glev); free(where); free(cnt); free(cand_cnt); free(cand); free(found); free(fdist); free(q); free(ident);
    free(f.slots); free(f.levels); free(f.targets); free(f.ids); free(f.offs);
    return rc;
}

void muninn_build_destroy(void *h) {
    Builder *b = (Builder *)h;
    if (!b) return;
    for (int t = 0; t < b->n_scratch; t++) {
        Scratch *sc = &b->scratch[t];
        free(sc->stamp); free(sc->cand.a); free(sc->res.a); free(sc->r_slot); free(sc->r_dist); free(sc->sel); free(sc->sel_d); free(sc->sel_cnt); free(sc->fresh); free(sc->fresh_d);
    }
    free(b->scratch); free(b->touched); free(b->touch_stamp); free(b->req_buf); free(b->grp_buf);
    free(b->level); free(b->nbr0); free(b->cnt0); free(b->d0); free(b->dU); free(b->upper_base); free(b->nbrU); free(b->cntU);
    free(b);
}
