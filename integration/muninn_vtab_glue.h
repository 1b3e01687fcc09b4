/* muninn_vtab_glue.h — what the patched hnsw_vtab.c calls (see muninn_vtab_glue.c). */
#ifndef MUNINN_VTAB_GLUE_H
#define MUNINN_VTAB_GLUE_H
#include "hnsw_algo.h"
int muninn_glue_search(HnswIndex *index, const float *query, int k, int ef_search, HnswSearchResult *results, char **err);
void muninn_glue_mark_dirty(const HnswIndex *index);                 /* force a full re-upload before the next search */
void muninn_glue_note_insert(const HnswIndex *index, int64_t id);    /* after a successful hnsw_insert(index, id, …) */
void muninn_glue_note_delete(const HnswIndex *index, int64_t id);    /* after a successful hnsw_delete(index, id) */
int64_t muninn_glue_total_patches(void);                             /* incremental updates / full uploads sent so far */
int64_t muninn_glue_total_uploads(void);
void muninn_glue_release(const HnswIndex *index);
/* batched callers (integration/muninn_batch_tvf.c) */
typedef struct sqlite3 sqlite3;             /* C11: repeating sqlite3.h's typedefs is allowed */
typedef struct sqlite3_value sqlite3_value;
int muninn_glue_search_batch(HnswIndex *index, const float *queries, int nq, int k, int ef_search, int64_t *ids,
                             float *dist, int32_t *counts, char **err);
/* in xFilter, before the blob-length check: 1 = argv[0] was a batch request (handled; *err set on failure), 0 = not */
int muninn_glue_batch_request(HnswIndex *index, int dim, sqlite3_value *arg, int k, int ef_search, char **err);
int muninn_glue_register_batch_tvf(sqlite3 *db); /* registers the eponymous hnsw_knn_batch(index_name, queries, k[, ef]) */
#endif
