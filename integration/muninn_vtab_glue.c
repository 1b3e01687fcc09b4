/*
 * muninn_vtab_glue.c — the host-side C a maintainer adds to sqlite-muninn to route the
 * hnsw_index KNN plan through libmuninn_cuda. It is compiled INTO the reference's extension
 * (see integration/build_patched_ext.sh and INTEGRATION.md); the vtab, its SQL surface, the
 * shadow tables and muninn.c stay as they are.
 *
 * The reference's call site is one line, /root/reference/src/hnsw_vtab.c:604:
 *     cur->result_count = hnsw_search(vtab->index, query, k, ef_search, cur->results);
 * which becomes
 *     cur->result_count = muninn_glue_search(vtab->index, query, k, ef_search, cur->results, &vtab->base.zErrMsg);
 * plus a muninn_glue_note_insert / muninn_glue_note_delete(vtab->index, id) after each hnsw_insert /
 * hnsw_delete in xUpdate (:698, :748) and a muninn_glue_release(vtab->index) before each hnsw_destroy in
 * xDisconnect / xDestroy (:466, :474).
 *
 * The glue keeps one device handle per HnswIndex in a small side table, so the reference's
 * structs need no new field. The first search flattens the host index (every stored node, deleted
 * ones included, neighbour lists in their live order) and uploads it. Later mutations are NOT
 * re-uploaded wholesale: hnsw_insert changes the new node and the nodes it linked it to, hnsw_delete
 * the deleted node and its former neighbours (src/hnsw_algo.c:520-666, :717-805), so the glue
 * remembers those ids and, on the next search, sends just their records with muninn_cuda_index_patch
 * (SURVEY.md §8f N4). It falls back to a full upload when an inserted rowid is not above every stored
 * one (slots would shift), when the pending change records outnumber the stored nodes, or when a list has
 * outgrown the device rows (hnsw_delete's reconnection has no capacity check: the library refuses the patch). There is no
 * CPU fallback: a failure of the device path is reported through zErrMsg and the query fails.
 */
#include "hnsw_algo.h" /* the reference's header, -I<reference>/src */
#include "muninn_cuda.h"
#include "sqlite3ext.h"

#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

SQLITE_EXTENSION_INIT3

typedef struct {
    const HnswIndex *index;
    MuninnCudaIndex *handle;
    int dirty;          /* full upload needed */
    int64_t *ids;       /* sorted rowids of the device copy: slot = rank */
    int64_t n, cap;
    int64_t *pend;      /* rowids whose record changed since the device copy was made */
    int64_t npend, cappend;
    int64_t max_id;     /* highest rowid on the device or pending */
    int64_t n_patches, n_uploads; /* counters for the tests */
} GlueSlot;

/* One heap-allocated slot per HnswIndex, found by pointer. The table itself is guarded by a mutex: connections on
 * different threads create and drop their indexes concurrently. A slot is only ever used by the connection that owns
 * its index (SQLite serialises the calls into one vtab), so the slot's fields need no lock of their own. */
static GlueSlot **g_slots;
static int g_nslots, g_capslots;
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;

static GlueSlot *slot_find(const HnswIndex *index, int create) {
    GlueSlot *s = NULL;
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < g_nslots; i++)
        if (g_slots[i]->index == index) { s = g_slots[i]; break; }
    if (!s && create) {
        if (g_nslots == g_capslots) {
            int nc = g_capslots ? 2 * g_capslots : 8;
            GlueSlot **ns = (GlueSlot **)realloc(g_slots, (size_t)nc * sizeof(GlueSlot *));
            if (ns) { g_slots = ns; g_capslots = nc; }
        }
        if (g_nslots < g_capslots && (s = (GlueSlot *)calloc(1, sizeof(GlueSlot))) != NULL) {
            s->index = index;
            s->dirty = 1;
            g_slots[g_nslots++] = s;
        }
    }
    pthread_mutex_unlock(&g_lock);
    return s;
}

static int cmp_i64(const void *a, const void *b) {
    int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
    return (x > y) - (x < y);
}

static int64_t slot_of(const int64_t *ids, int64_t n, int64_t id) {
    int64_t lo = 0, hi = n - 1;
    while (lo <= hi) {
        int64_t mid = lo + (hi - lo) / 2;
        if (ids[mid] == id) return mid;
        if (ids[mid] < id) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

void muninn_glue_mark_dirty(const HnswIndex *index) {
    GlueSlot *s = slot_find(index, 0);
    if (s) s->dirty = 1;
}

static void pend_add(GlueSlot *s, int64_t id) {
    if (s->dirty) return;
    if (s->npend == s->cappend) {
        int64_t nc = s->cappend ? 2 * s->cappend : 256;
        int64_t *np = (int64_t *)realloc(s->pend, (size_t)nc * sizeof(int64_t));
        if (!np) { s->dirty = 1; return; }
        s->pend = np; s->cappend = nc;
    }
    s->pend[s->npend++] = id;
    if (s->npend > 64 && s->npend > s->n) s->dirty = 1; /* more change records than nodes: start over */
}

/* the node `id` and every node in its lists: all that hnsw_insert(id) or hnsw_delete(id) modified */
static void note_node_and_neighbours(const HnswIndex *index, int64_t id) {
    GlueSlot *s = slot_find(index, 0);
    if (!s || !s->handle || s->dirty) return; /* nothing on the device yet, or a full upload is due anyway */
    const HnswNode *nd = ht_find(index->nodes, index->node_capacity, id);
    if (!nd) { s->dirty = 1; return; }
    pend_add(s, id);
    for (int l = 0; l <= nd->level; l++)
        for (int j = 0; j < nd->neighbor_count[l]; j++) pend_add(s, nd->neighbors[l][j]);
}

void muninn_glue_note_insert(const HnswIndex *index, int64_t id) {
    GlueSlot *s = slot_find(index, 0);
    if (s && s->handle && !s->dirty) {
        if (s->n > 0 && id <= s->max_id) s->dirty = 1; /* not above every stored rowid: slots would shift */
        else s->max_id = id;
    }
    note_node_and_neighbours(index, id);
}

void muninn_glue_note_delete(const HnswIndex *index, int64_t id) { note_node_and_neighbours(index, id); }

/* process-wide counters (the tests read them through the loaded extension with ctypes) */
static int64_t g_total_patches, g_total_uploads; /* updated atomically */
int64_t muninn_glue_total_patches(void) { return g_total_patches; }
int64_t muninn_glue_total_uploads(void) { return g_total_uploads; }

void muninn_glue_release(const HnswIndex *index) {
    GlueSlot *s = NULL;
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < g_nslots; i++)
        if (g_slots[i]->index == index) { s = g_slots[i]; g_slots[i] = g_slots[--g_nslots]; break; }
    pthread_mutex_unlock(&g_lock);
    if (!s) return;
    if (s->handle) muninn_cuda_index_destroy(s->handle);
    free(s->ids); free(s->pend);
    free(s);
}

/* host HnswIndex → the flat arrays of muninn_cuda_index_upload */
static int glue_upload(GlueSlot *s, char **err) {
    const HnswIndex *idx = s->index;
    int64_t n = 0, n_lists = 0, n_edges = 0;
    for (int i = 0; i < idx->node_capacity; i++) {
        const HnswNode *nd = idx->nodes[i];
        if (!nd) continue;
        n++; n_lists += nd->level + 1;
        for (int l = 0; l <= nd->level; l++) n_edges += nd->neighbor_count[l];
    }
    int64_t *ids = (int64_t *)malloc((size_t)(n + 1) * sizeof(int64_t));
    int32_t *levels = (int32_t *)malloc((size_t)(n + 1) * sizeof(int32_t));
    uint8_t *deleted = (uint8_t *)malloc((size_t)(n + 1));
    float *vecs = (float *)malloc((size_t)(n + 1) * (size_t)idx->dim * sizeof(float));
    int64_t *offs = (int64_t *)malloc((size_t)(n_lists + 1) * sizeof(int64_t));
    int32_t *tg = (int32_t *)malloc((size_t)(n_edges + 1) * sizeof(int32_t));
    int rc = MUNINN_CUDA_E_NOMEM;
    if (ids && levels && deleted && vecs && offs && tg) {
        int64_t k = 0;
        for (int i = 0; i < idx->node_capacity; i++)
            if (idx->nodes[i]) ids[k++] = idx->nodes[i]->id;
        qsort(ids, (size_t)n, sizeof(int64_t), cmp_i64);
        int64_t li = 0, e = 0;
        offs[0] = 0;
        for (int64_t sl = 0; sl < n; sl++) {
            const HnswNode *nd = ht_find(idx->nodes, idx->node_capacity, ids[sl]);
            levels[sl] = nd->level;
            deleted[sl] = (uint8_t)(nd->deleted != 0);
            memcpy(vecs + (size_t)sl * (size_t)idx->dim, nd->vector, (size_t)idx->dim * sizeof(float));
            for (int l = 0; l <= nd->level; l++) {
                /* the LIVE list order is kept: insertion order on a freshly built index, ascending
                 * target id after xConnect reloaded it — so the device returns what the host
                 * hnsw_search would return in either state (SURVEY.md finding 5) */
                for (int j = 0; j < nd->neighbor_count[l]; j++)
                    tg[e++] = (int32_t)slot_of(ids, n, nd->neighbors[l][j]);
                offs[++li] = e;
            }
        }
        int32_t entry = idx->entry_point == -1 ? -1 : (int32_t)slot_of(ids, n, idx->entry_point);
        rc = muninn_cuda_index_upload(s->handle, n, ids, levels, deleted, vecs, offs, tg, entry, idx->max_level);
        if (rc != MUNINN_CUDA_OK && err)
            *err = sqlite3_mprintf("hnsw_index: GPU upload failed: %s", muninn_cuda_last_error(s->handle));
    } else if (err) {
        *err = sqlite3_mprintf("hnsw_index: out of memory flattening the index for upload");
    }
    free(levels); free(deleted); free(vecs); free(offs); free(tg);
    if (rc == MUNINN_CUDA_OK) {
        free(s->ids);
        s->ids = ids; s->n = n; s->cap = n + 1; s->max_id = n ? ids[n - 1] : 0;
        s->dirty = 0; s->npend = 0; s->n_uploads++; __atomic_add_fetch(&g_total_uploads, 1, __ATOMIC_RELAXED);
    } else free(ids);
    return rc;
}

/* pending change records → muninn_cuda_index_patch. Returns MUNINN_CUDA_OK, or an error after which the
 * caller falls back to a full upload. */
static int glue_patch(GlueSlot *s) {
    const HnswIndex *idx = s->index;
    qsort(s->pend, (size_t)s->npend, sizeof(int64_t), cmp_i64);
    int64_t np = 0;
    for (int64_t i = 0; i < s->npend; i++)
        if (np == 0 || s->pend[i] != s->pend[np - 1]) s->pend[np++] = s->pend[i];
    /* new rowids (absent from the device copy) are above every stored one and come out of the sort ascending */
    int64_t n_new = 0;
    for (int64_t i = 0; i < np; i++)
        if (slot_of(s->ids, s->n, s->pend[i]) < 0) n_new++;
    const int64_t n_total = s->n + n_new;
    if (n_total + 1 > s->cap) {
        int64_t nc = n_total + n_total / 2 + 64;
        int64_t *ni = (int64_t *)realloc(s->ids, (size_t)nc * sizeof(int64_t));
        if (!ni) return MUNINN_CUDA_E_NOMEM;
        s->ids = ni; s->cap = nc;
    }
    int64_t k = s->n;
    for (int64_t i = 0; i < np; i++)
        if (slot_of(s->ids, s->n, s->pend[i]) < 0) {
            if (k > 0 && s->pend[i] <= s->ids[k - 1]) return MUNINN_CUDA_E_ARG;
            s->ids[k++] = s->pend[i];
        }
    int64_t n_lists = 0, n_edges = 0;
    for (int64_t i = 0; i < np; i++) {
        const HnswNode *nd = ht_find(idx->nodes, idx->node_capacity, s->pend[i]);
        if (!nd) return MUNINN_CUDA_E_ARG;
        n_lists += nd->level + 1;
        for (int l = 0; l <= nd->level; l++) n_edges += nd->neighbor_count[l];
    }
    int32_t *slots = (int32_t *)malloc((size_t)(np + 1) * sizeof(int32_t));
    int32_t *levels = (int32_t *)malloc((size_t)(np + 1) * sizeof(int32_t));
    uint8_t *deleted = (uint8_t *)malloc((size_t)(np + 1));
    float *vecs = (float *)malloc((size_t)(np + 1) * (size_t)idx->dim * sizeof(float));
    int64_t *offs = (int64_t *)malloc((size_t)(n_lists + 1) * sizeof(int64_t));
    int32_t *tg = (int32_t *)malloc((size_t)(n_edges + 1) * sizeof(int32_t));
    int rc = MUNINN_CUDA_E_NOMEM;
    if (slots && levels && deleted && vecs && offs && tg) {
        int64_t li = 0, e = 0;
        offs[0] = 0;
        for (int64_t i = 0; i < np; i++) {
            const HnswNode *nd = ht_find(idx->nodes, idx->node_capacity, s->pend[i]);
            slots[i] = (int32_t)slot_of(s->ids, n_total, s->pend[i]);
            levels[i] = nd->level;
            deleted[i] = (uint8_t)(nd->deleted != 0);
            memcpy(vecs + (size_t)i * (size_t)idx->dim, nd->vector, (size_t)idx->dim * sizeof(float));
            for (int l = 0; l <= nd->level; l++) {
                for (int j = 0; j < nd->neighbor_count[l]; j++) tg[e++] = (int32_t)slot_of(s->ids, n_total, nd->neighbors[l][j]);
                offs[++li] = e;
            }
        }
        int32_t entry = idx->entry_point == -1 ? -1 : (int32_t)slot_of(s->ids, n_total, idx->entry_point);
        rc = muninn_cuda_index_patch(s->handle, n_total, (int32_t)np, slots, s->pend, levels, deleted, vecs, offs, tg, entry,
                                     idx->max_level);
    }
    free(slots); free(levels); free(deleted); free(vecs); free(offs); free(tg);
    if (getenv("MUNINN_GLUE_TRACE")) fprintf(stderr, "muninn_glue: patch of %lld records (%lld new) -> rc %d\n", (long long)np, (long long)n_new, rc);
    if (rc == MUNINN_CUDA_OK) { s->n = n_total; s->npend = 0; s->n_patches++; __atomic_add_fetch(&g_total_patches, 1, __ATOMIC_RELAXED); }
    return rc;
}

/* handle of `index`, created on first use and brought up to date (patch, else full upload); NULL with *err set */
static GlueSlot *glue_sync(HnswIndex *index, char **err) {
    GlueSlot *s = slot_find(index, 1);
    if (!s) { if (err) *err = sqlite3_mprintf("hnsw_index: out of memory"); return NULL; }
    if (!s->handle) {
        int dev = 0;
        const char *env = getenv("MUNINN_CUDA_DEVICE");
        if (env) dev = atoi(env);
        int rc = muninn_cuda_index_create(dev, index->dim, (int)index->metric, index->M, &s->handle);
        if (rc != MUNINN_CUDA_OK) {
            if (err) *err = sqlite3_mprintf("hnsw_index: %s", muninn_cuda_last_error(NULL));
            return NULL;
        }
        s->dirty = 1;
    }
    if (!s->dirty && s->npend > 0 && glue_patch(s) != MUNINN_CUDA_OK) s->dirty = 1; /* anything a patch cannot express */
    if (s->dirty && glue_upload(s, err) != MUNINN_CUDA_OK) return NULL;
    return s;
}

/* Batched form (the per-entity KNN loop of src/llama_er.c:237-284 as ONE launch): nq queries
 * [nq × dim] → ids [nq × k] (rowids, -1 padded), dist [nq × k], counts [nq]. Returns 0, or -1 with *err set. */
int muninn_glue_search_batch(HnswIndex *index, const float *queries, int nq, int k, int ef_search, int64_t *ids,
                             float *dist, int32_t *counts, char **err) {
    if (nq < 1 || k < 1) return 0;
    if (!index || index->entry_point == -1 || index->node_count == 0) { /* hnsw_algo.c:671 */
        for (int i = 0; i < nq; i++) counts[i] = 0;
        return 0;
    }
    GlueSlot *s = glue_sync(index, err);
    if (!s) return -1;
    int rc = muninn_cuda_search_batch(s->handle, queries, nq, k, ef_search, ids, dist, counts, NULL);
    if (rc != MUNINN_CUDA_OK) {
        if (err) *err = sqlite3_mprintf("hnsw_index: GPU search failed: %s", muninn_cuda_last_error(s->handle));
        return -1;
    }
    return 0;
}

/* Drop-in for hnsw_search (src/hnsw_algo.h:73): same arguments and return value (number of
 * results, 0 for an empty index); on a device failure returns 0 with *err set. */
int muninn_glue_search(HnswIndex *index, const float *query, int k, int ef_search, HnswSearchResult *results,
                       char **err) {
    if (!index || index->entry_point == -1 || index->node_count == 0 || k < 1) return 0; /* hnsw_algo.c:671 */
    GlueSlot *s = glue_sync(index, err);
    if (!s) return 0;
    int64_t *ids = (int64_t *)malloc((size_t)k * sizeof(int64_t));
    float *dist = (float *)malloc((size_t)k * sizeof(float));
    int32_t count = 0;
    if (!ids || !dist) { free(ids); free(dist); if (err) *err = sqlite3_mprintf("hnsw_index: out of memory"); return 0; }
    /* the blob pointer SQLite hands out may be unaligned (hnsw_vtab.c:588 casts it): copy */
    float *q = (float *)malloc((size_t)index->dim * sizeof(float));
    if (!q) { free(ids); free(dist); if (err) *err = sqlite3_mprintf("hnsw_index: out of memory"); return 0; }
    memcpy(q, query, (size_t)index->dim * sizeof(float));
    int rc = muninn_cuda_search_batch(s->handle, q, 1, k, ef_search, ids, dist, &count, NULL);
    if (rc != MUNINN_CUDA_OK) {
        if (err) *err = sqlite3_mprintf("hnsw_index: GPU search failed: %s", muninn_cuda_last_error(s->handle));
        count = 0;
    }
    for (int i = 0; i < count; i++) { results[i].id = ids[i]; results[i].distance = dist[i]; }
    free(ids); free(dist); free(q);
    return count;
}
