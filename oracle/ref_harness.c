/*
 * ref_harness.c — flat-array test harness around the UNMODIFIED reference sources.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is compiled together with
 * /root/reference/src/{hnsw_algo.c,vec_math.c,priority_queue.c} (from where they
 * lie; nothing is copied into this repo) into oracle/_ref/libmuninn_ref.so by
 * oracle/Makefile.  It exposes the reference's own hnsw_insert / hnsw_search /
 * hnsw_delete through plain pointers so that tests (ctypes) and bench.py's
 * `--impl reference` leg can (a) build graphs with the real reference, (b) export
 * and import them in the flat "list CSR" format that include/muninn_cuda.h
 * documents, and (c) run and time the reference's search on the host cores.
 *
 * Nothing in the product path (sqlite_muninn_b200/) links or loads this file.
 */
#define _POSIX_C_SOURCE 200809L
#include "hnsw_algo.h" /* from -I/root/reference/src */

#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ── distance-call counter (the reference calls idx->dist_func through a pointer,
 *    src/hnsw_algo.c:262,273,362,411, so a counting wrapper needs no source change) ── */
static __thread int64_t tl_ndist;
static float count_l2(const float *a, const float *b, int d) { tl_ndist++; return vec_l2_distance(a, b, d); }
static float count_cos(const float *a, const float *b, int d) { tl_ndist++; return vec_cosine_distance(a, b, d); }
static float count_ip(const float *a, const float *b, int d) { tl_ndist++; return vec_inner_product_distance(a, b, d); }

static VecDistanceFunc counting_func(VecMetric m) {
    switch (m) {
    case VEC_METRIC_L2: return count_l2;
    case VEC_METRIC_COSINE: return count_cos;
    case VEC_METRIC_INNER_PRODUCT: return count_ip;
    }
    return count_l2;
}

void *ref_create(int dim, int metric, int M, int efc) { return hnsw_create(dim, (VecMetric)metric, M, efc); }
void ref_destroy(void *h) { hnsw_destroy((HnswIndex *)h); }

int ref_insert_many(void *h, int64_t n, const int64_t *ids, const float *vecs) {
    HnswIndex *idx = (HnswIndex *)h;
    for (int64_t i = 0; i < n; i++)
        if (hnsw_insert(idx, ids[i], vecs + (size_t)i * (size_t)idx->dim) != 0)
            return -1;
    return 0;
}

int ref_delete(void *h, int64_t id) { return hnsw_delete((HnswIndex *)h, id); }

static int cmp_i64(const void *a, const void *b) {
    int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
    return (x > y) - (x < y);
}

/* Put every neighbour list into ascending target-id order: the order in which
 * load_index_from_shadow (src/hnsw_vtab.c:322-338) re-creates them from the
 * WITHOUT ROWID edges table whose primary key is (source_id, level, target_id)
 * (src/hnsw_vtab.c:163-169).  SURVEY.md finding 5. */
void ref_sort_lists(void *h) {
    HnswIndex *idx = (HnswIndex *)h;
    for (int i = 0; i < idx->node_capacity; i++) {
        HnswNode *nd = idx->nodes[i];
        if (!nd) continue;
        for (int l = 0; l <= nd->level; l++)
            if (nd->neighbor_count[l] > 1)
                qsort(nd->neighbors[l], (size_t)nd->neighbor_count[l], sizeof(int64_t), cmp_i64);
    }
}

/* Emulate xDisconnect + xConnect after DELETEs (src/hnsw_vtab.c:686-716 only writes
 * `deleted = 1`; edges in the shadow table are whatever persist_node last wrote). The
 * harness cannot know the shadow state, so this is only the list-order part. */

void ref_export_sizes(void *h, int64_t *n_nodes, int64_t *n_lists, int64_t *n_edges) {
    HnswIndex *idx = (HnswIndex *)h;
    int64_t nn = 0, nl = 0, ne = 0;
    for (int i = 0; i < idx->node_capacity; i++) {
        HnswNode *nd = idx->nodes[i];
        if (!nd) continue;
        nn++;
        nl += nd->level + 1;
        for (int l = 0; l <= nd->level; l++) ne += nd->neighbor_count[l];
    }
    *n_nodes = nn; *n_lists = nl; *n_edges = ne;
}

static int64_t find_slot(const int64_t *ids, int64_t n, int64_t id) {
    int64_t lo = 0, hi = n - 1;
    while (lo <= hi) {
        int64_t mid = lo + (hi - lo) / 2;
        if (ids[mid] == id) return mid;
        if (ids[mid] < id) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

/* Flat export: slot = rank of id among all stored nodes (deleted ones included, since
 * load_index_from_shadow keeps them in the table, src/hnsw_vtab.c:309-317). Lists are
 * node-major, level-minor; a target that is not a stored node becomes -1. */
void ref_export(void *h, int64_t *ids, int32_t *levels, uint8_t *deleted, float *vecs, int64_t *list_offsets,
                int32_t *targets, int32_t *entry_slot, int32_t *max_level) {
    HnswIndex *idx = (HnswIndex *)h;
    int64_t n = 0;
    for (int i = 0; i < idx->node_capacity; i++)
        if (idx->nodes[i]) ids[n++] = idx->nodes[i]->id;
    qsort(ids, (size_t)n, sizeof(int64_t), cmp_i64);
    int64_t li = 0, e = 0;
    list_offsets[0] = 0;
    for (int64_t s = 0; s < n; s++) {
        HnswNode *nd = ht_find(idx->nodes, idx->node_capacity, ids[s]);
        levels[s] = nd->level;
        deleted[s] = (uint8_t)(nd->deleted != 0);
        memcpy(vecs + (size_t)s * (size_t)idx->dim, nd->vector, (size_t)idx->dim * sizeof(float));
        for (int l = 0; l <= nd->level; l++) {
            for (int j = 0; j < nd->neighbor_count[l]; j++)
                targets[e++] = (int32_t)find_slot(ids, n, nd->neighbors[l][j]);
            list_offsets[++li] = e;
        }
    }
    *entry_slot = (int32_t)(idx->entry_point == -1 ? -1 : find_slot(ids, n, idx->entry_point));
    *max_level = idx->max_level;
}

/* Flat import: the inverse of ref_export, using the reference's own node_create /
 * ht_insert / node_add_neighbor exactly as load_index_from_shadow does. */
void *ref_import(int dim, int metric, int M, int efc, int64_t n, const int64_t *ids, const int32_t *levels,
                 const uint8_t *deleted, const float *vecs, const int64_t *list_offsets, const int32_t *targets,
                 int32_t entry_slot, int32_t max_level) {
    HnswIndex *idx = hnsw_create(dim, (VecMetric)metric, M, efc);
    if (!idx) return NULL;
    for (int64_t s = 0; s < n; s++) {
        if (idx->node_count * 10 > idx->node_capacity * 7) ht_resize(idx);
        /* node_count below counts every stored node so that the table keeps growing;
         * corrected after the loop to the reference's "non-deleted only" meaning. */
        HnswNode *nd = node_create(ids[s], vecs + (size_t)s * (size_t)dim, dim, levels[s]);
        if (!nd) { hnsw_destroy(idx); return NULL; }
        nd->deleted = deleted ? deleted[s] : 0;
        ht_insert(idx->nodes, idx->node_capacity, nd);
        idx->node_count++;
    }
    int64_t li = 0, fake = INT64_MIN;
    int live = 0;
    for (int64_t s = 0; s < n; s++) {
        HnswNode *nd = ht_find(idx->nodes, idx->node_capacity, ids[s]);
        if (!nd->deleted) live++;
        for (int l = 0; l <= levels[s]; l++, li++)
            for (int64_t e = list_offsets[li]; e < list_offsets[li + 1]; e++)
                node_add_neighbor(nd, l, targets[e] >= 0 ? ids[targets[e]] : fake++);
    }
    idx->node_count = live;
    idx->entry_point = entry_slot >= 0 ? ids[entry_slot] : -1;
    idx->max_level = max_level;
    return idx;
}

int ref_search(void *h, const float *q, int nq, int k, int ef, int64_t *out_ids, float *out_dist, int32_t *out_count,
               int64_t *out_ndist) {
    HnswIndex *idx = (HnswIndex *)h;
    VecDistanceFunc saved = idx->dist_func;
    if (out_ndist) idx->dist_func = counting_func(idx->metric);
    HnswSearchResult *res = (HnswSearchResult *)malloc((size_t)(k > 0 ? k : 1) * sizeof(HnswSearchResult));
    for (int i = 0; i < nq; i++) {
        tl_ndist = 0;
        int c = hnsw_search(idx, q + (size_t)i * (size_t)idx->dim, k, ef, res);
        out_count[i] = c;
        for (int j = 0; j < k; j++) {
            out_ids[(size_t)i * k + j] = j < c ? res[j].id : -1;
            out_dist[(size_t)i * k + j] = j < c ? res[j].distance : 0.0f;
        }
        if (out_ndist) out_ndist[i] = tl_ndist;
    }
    free(res);
    idx->dist_func = saved;
    return 0;
}

/* ── multi-threaded timing leg: queries partitioned over host threads, each thread
 *    calling the unmodified hnsw_search on the shared read-only index (safe: the search
 *    only reads the index and mallocs its own scratch, SURVEY.md §8b "Threading"). ── */
typedef struct {
    HnswIndex *idx; const float *q; int q0, q1, k, ef;
    int64_t *out_ids; float *out_dist; int32_t *out_count;
} MtArg;

static void *mt_worker(void *p) {
    MtArg *a = (MtArg *)p;
    HnswSearchResult *res = (HnswSearchResult *)malloc((size_t)a->k * sizeof(HnswSearchResult));
    for (int i = a->q0; i < a->q1; i++) {
        int c = hnsw_search(a->idx, a->q + (size_t)i * (size_t)a->idx->dim, a->k, a->ef, res);
        a->out_count[i] = c;
        for (int j = 0; j < a->k; j++) {
            a->out_ids[(size_t)i * a->k + j] = j < c ? res[j].id : -1;
            a->out_dist[(size_t)i * a->k + j] = j < c ? res[j].distance : 0.0f;
        }
    }
    free(res);
    return NULL;
}

double ref_search_mt(void *h, const float *q, int nq, int k, int ef, int nthreads, int64_t *out_ids, float *out_dist,
                     int32_t *out_count) {
    if (nthreads < 1) nthreads = 1;
    pthread_t *th = (pthread_t *)malloc((size_t)nthreads * sizeof(pthread_t));
    MtArg *args = (MtArg *)malloc((size_t)nthreads * sizeof(MtArg));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < nthreads; t++) {
        args[t] = (MtArg){(HnswIndex *)h, q, (int)((int64_t)nq * t / nthreads), (int)((int64_t)nq * (t + 1) / nthreads),
                          k, ef, out_ids, out_dist, out_count};
        pthread_create(&th[t], NULL, mt_worker, &args[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th); free(args);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

float ref_distance(int metric, const float *a, const float *b, int dim) {
    return vec_get_distance_func((VecMetric)metric)(a, b, dim);
}

int ref_parse_metric(const char *name) {
    VecMetric m;
    return vec_parse_metric(name, &m) == 0 ? (int)m : -1;
}

int ref_entry_point(void *h, int64_t *entry, int32_t *max_level, int32_t *node_count) {
    HnswIndex *idx = (HnswIndex *)h;
    *entry = idx->entry_point; *max_level = idx->max_level; *node_count = idx->node_count;
    return 0;
}
