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
#endif
