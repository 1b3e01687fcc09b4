/*
 * muninn_batch_tvf.c — a table-valued batch form of `vector MATCH ?`.
 *
 * The reference's batched KNN callers issue one SQL KNN per entity: the ER blocking stage
 * (/root/reference/src/llama_er.c:237-284) steps
 *     SELECT rowid, distance FROM <tbl> WHERE vector MATCH ?1 AND k = ?2
 * once per entity, i.e. one launch-latency-bound search per row. This file adds, to the SAME
 * extension, an eponymous table-valued function
 *
 *     SELECT query, id, distance, rank
 *       FROM hnsw_knn_batch('<tbl>', <nq·dim·4-byte blob>, <k> [, <ef_search>]);
 *
 * that answers all nq queries with ONE muninn_cuda_search_batch launch. Row (query = i, rank = j)
 * is exactly row j of the reference statement above run with query i: same rowid, same (double)
 * distance, same order, same count (tests/test_gpu_vtab.py compares against the unmodified
 * extension); `ef_search` defaults to 2k as in the vtab (src/hnsw_vtab.c:591).
 *
 * How the function reaches the HnswIndex that lives inside another virtual table without a
 * registry or a new struct field: it prepares the reference's own KNN statement on <tbl> and binds,
 * instead of a blob, a POINTER value (sqlite3_bind_pointer, type "muninn_knn_batch") to a request
 * block. The patched xFilter hands argv[0] to muninn_glue_batch_request() before its blob-length
 * check; for a pointer of that type the glue runs the batch, fills the request and the statement
 * returns no rows. Any other value falls through to the reference's code unchanged. SQL cannot
 * forge a pointer value, so the hook is unreachable from plain SQL.
 */
#include "muninn_vtab_glue.h"
#include "sqlite3ext.h"

#include <stdint.h>
#include <stdlib.h>
#include <string.h>

SQLITE_EXTENSION_INIT3

#define BATCH_PTR_TYPE "muninn_knn_batch"

typedef struct {
    const float *queries; /* aligned copy of the blob */
    int64_t bytes;
    int nq, k;            /* filled by the hook */
    int64_t *ids;         /* [nq × k], malloc'ed by the hook, freed by the cursor */
    float *dist;
    int32_t *counts;
    int done;
} BatchRequest;

int muninn_glue_batch_request(HnswIndex *index, int dim, sqlite3_value *arg, int k, int ef_search, char **err) {
    BatchRequest *rq = (BatchRequest *)sqlite3_value_pointer(arg, BATCH_PTR_TYPE);
    if (!rq) return 0; /* an ordinary MATCH argument: the reference's code handles it */
    const int64_t row = (int64_t)dim * (int64_t)sizeof(float);
    if (rq->bytes <= 0 || rq->bytes % row != 0) {
        *err = sqlite3_mprintf("hnsw_index: expected a whole number of %d-dim vectors (%d bytes each), got %lld bytes", dim,
                               (int)row, (long long)rq->bytes);
        return 1;
    }
    if (k < 1 || rq->bytes / row > 0x7fffffff / (int64_t)k) {
        *err = sqlite3_mprintf("hnsw_index: batch of %lld queries with k = %d is out of range", (long long)(rq->bytes / row), k);
        return 1;
    }
    rq->nq = (int)(rq->bytes / row);
    rq->k = k;
    rq->ids = (int64_t *)malloc((size_t)rq->nq * (size_t)k * sizeof(int64_t));
    rq->dist = (float *)malloc((size_t)rq->nq * (size_t)k * sizeof(float));
    rq->counts = (int32_t *)malloc((size_t)rq->nq * sizeof(int32_t));
    if (!rq->ids || !rq->dist || !rq->counts) {
        *err = sqlite3_mprintf("hnsw_index: out of memory");
        return 1;
    }
    if (muninn_glue_search_batch(index, rq->queries, rq->nq, k, ef_search, rq->ids, rq->dist, rq->counts, err) == 0)
        rq->done = 1;
    return 1;
}

/* ── the eponymous module ───────────────────────────────────────────────────────────── */

enum { COL_QUERY = 0, COL_ID, COL_DISTANCE, COL_RANK, COL_TABLE, COL_QUERIES, COL_K, COL_EF };

typedef struct {
    sqlite3_vtab base;
    sqlite3 *db;
} BatchVtab;

typedef struct {
    sqlite3_vtab_cursor base;
    BatchRequest rq;
    float *qcopy;
    int q, j; /* current row: query q, rank j */
} BatchCursor;

static int batch_connect(sqlite3 *db, void *aux, int argc, const char *const *argv, sqlite3_vtab **out, char **err) {
    (void)aux; (void)argc; (void)argv; (void)err;
    int rc = sqlite3_declare_vtab(db, "CREATE TABLE x(query INTEGER, id INTEGER, distance REAL, rank INTEGER, "
                                      "index_name HIDDEN, queries HIDDEN, k HIDDEN, ef_search HIDDEN)");
    if (rc != SQLITE_OK) return rc;
    BatchVtab *v = (BatchVtab *)sqlite3_malloc(sizeof(BatchVtab));
    if (!v) return SQLITE_NOMEM;
    memset(v, 0, sizeof(*v));
    v->db = db;
    *out = &v->base;
    return SQLITE_OK;
}

static int batch_disconnect(sqlite3_vtab *v) { sqlite3_free(v); return SQLITE_OK; }

/* idxNum: bit 0 = ef_search supplied. The three required arguments arrive as argv[0..2] in column order. */
static int batch_best_index(sqlite3_vtab *v, sqlite3_index_info *info) {
    int at[4] = {-1, -1, -1, -1};
    for (int i = 0; i < info->nConstraint; i++) {
        const struct sqlite3_index_constraint *c = &info->aConstraint[i];
        if (c->op != SQLITE_INDEX_CONSTRAINT_EQ || c->iColumn < COL_TABLE || c->iColumn > COL_EF) continue;
        if (!c->usable) return SQLITE_CONSTRAINT;
        at[c->iColumn - COL_TABLE] = i;
    }
    if (at[0] < 0 || at[1] < 0 || at[2] < 0) {
        v->zErrMsg = sqlite3_mprintf("hnsw_knn_batch(index_name, queries, k [, ef_search]): missing argument");
        return SQLITE_ERROR;
    }
    int n = 0;
    for (int a = 0; a < 4; a++)
        if (at[a] >= 0) {
            info->aConstraintUsage[at[a]].argvIndex = ++n;
            info->aConstraintUsage[at[a]].omit = 1;
        }
    info->idxNum = at[3] >= 0;
    info->estimatedCost = 1000.0;
    info->estimatedRows = 1000;
    return SQLITE_OK;
}

static int batch_open(sqlite3_vtab *v, sqlite3_vtab_cursor **out) {
    (void)v;
    BatchCursor *c = (BatchCursor *)sqlite3_malloc(sizeof(BatchCursor));
    if (!c) return SQLITE_NOMEM;
    memset(c, 0, sizeof(*c));
    *out = &c->base;
    return SQLITE_OK;
}

static void batch_reset(BatchCursor *c) {
    free(c->rq.ids); free(c->rq.dist); free(c->rq.counts); free(c->qcopy);
    memset(&c->rq, 0, sizeof(c->rq));
    c->qcopy = NULL;
    c->q = c->j = 0;
}

static int batch_close(sqlite3_vtab_cursor *cur) {
    batch_reset((BatchCursor *)cur);
    sqlite3_free(cur);
    return SQLITE_OK;
}

static void batch_skip_empty(BatchCursor *c) {
    while (c->q < c->rq.nq && c->j >= c->rq.counts[c->q]) { c->q++; c->j = 0; }
}

static int batch_filter(sqlite3_vtab_cursor *cur, int idxNum, const char *idxStr, int argc, sqlite3_value **argv) {
    (void)idxStr;
    BatchCursor *c = (BatchCursor *)cur;
    BatchVtab *v = (BatchVtab *)cur->pVtab;
    batch_reset(c);
    if (argc < 3) return SQLITE_ERROR;
    const char *name = (const char *)sqlite3_value_text(argv[0]);
    const void *blob = sqlite3_value_blob(argv[1]);
    const int bytes = sqlite3_value_bytes(argv[1]);
    const int k = sqlite3_value_int(argv[2]);
    const int ef = (idxNum & 1) && argc >= 4 ? sqlite3_value_int(argv[3]) : 2 * k; /* src/hnsw_vtab.c:591 */
    if (!name || sqlite3_value_type(argv[1]) != SQLITE_BLOB || k < 1) {
        v->base.zErrMsg = sqlite3_mprintf("hnsw_knn_batch: expected (index_name TEXT, queries BLOB, k >= 1 [, ef_search])");
        return SQLITE_ERROR;
    }
    if (bytes == 0) return SQLITE_OK; /* no queries, no rows */
    c->qcopy = (float *)malloc((size_t)bytes); /* SQLite's blob pointer may be unaligned */
    if (!c->qcopy) return SQLITE_NOMEM;
    memcpy(c->qcopy, blob, (size_t)bytes);
    c->rq.queries = c->qcopy;
    c->rq.bytes = bytes;

    char *sql = sqlite3_mprintf("SELECT rowid FROM \"%w\" WHERE vector MATCH ?1 AND k = ?2 AND ef_search = ?3", name);
    if (!sql) return SQLITE_NOMEM;
    sqlite3_stmt *st = NULL;
    int rc = sqlite3_prepare_v2(v->db, sql, -1, &st, NULL);
    sqlite3_free(sql);
    if (rc != SQLITE_OK) {
        v->base.zErrMsg = sqlite3_mprintf("hnsw_knn_batch: %s", sqlite3_errmsg(v->db));
        return rc;
    }
    sqlite3_bind_pointer(st, 1, &c->rq, BATCH_PTR_TYPE, NULL);
    sqlite3_bind_int(st, 2, k);
    sqlite3_bind_int(st, 3, ef);
    rc = sqlite3_step(st);
    if (rc != SQLITE_DONE) {
        v->base.zErrMsg = sqlite3_mprintf("hnsw_knn_batch: %s", sqlite3_errmsg(v->db));
        sqlite3_finalize(st);
        batch_reset(c);
        return SQLITE_ERROR;
    }
    sqlite3_finalize(st);
    if (!c->rq.done) {
        v->base.zErrMsg = sqlite3_mprintf("hnsw_knn_batch: \"%s\" is not an hnsw_index table of this extension", name);
        batch_reset(c);
        return SQLITE_ERROR;
    }
    batch_skip_empty(c);
    return SQLITE_OK;
}

static int batch_next(sqlite3_vtab_cursor *cur) {
    BatchCursor *c = (BatchCursor *)cur;
    c->j++;
    batch_skip_empty(c);
    return SQLITE_OK;
}

static int batch_eof(sqlite3_vtab_cursor *cur) {
    BatchCursor *c = (BatchCursor *)cur;
    return c->q >= c->rq.nq;
}

static int batch_column(sqlite3_vtab_cursor *cur, sqlite3_context *ctx, int col) {
    BatchCursor *c = (BatchCursor *)cur;
    const size_t at = (size_t)c->q * (size_t)c->rq.k + (size_t)c->j;
    switch (col) {
    case COL_QUERY: sqlite3_result_int(ctx, c->q); break;
    case COL_ID: sqlite3_result_int64(ctx, c->rq.ids[at]); break;
    case COL_DISTANCE: sqlite3_result_double(ctx, (double)c->rq.dist[at]); break; /* as src/hnsw_vtab.c:648-684 */
    case COL_RANK: sqlite3_result_int(ctx, c->j); break;
    default: sqlite3_result_null(ctx); break;
    }
    return SQLITE_OK;
}

static int batch_rowid(sqlite3_vtab_cursor *cur, sqlite3_int64 *rowid) {
    BatchCursor *c = (BatchCursor *)cur;
    *rowid = (sqlite3_int64)c->q * c->rq.k + c->j;
    return SQLITE_OK;
}

static sqlite3_module batch_module = {
    .iVersion = 0,
    .xCreate = NULL, /* eponymous-only */
    .xConnect = batch_connect,
    .xBestIndex = batch_best_index,
    .xDisconnect = batch_disconnect,
    .xOpen = batch_open,
    .xClose = batch_close,
    .xFilter = batch_filter,
    .xNext = batch_next,
    .xEof = batch_eof,
    .xColumn = batch_column,
    .xRowid = batch_rowid,
};

int muninn_glue_register_batch_tvf(sqlite3 *db) { return sqlite3_create_module(db, "hnsw_knn_batch", &batch_module, NULL); }
