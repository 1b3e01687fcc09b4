/* muninn_vtab_glue.h — what the patched hnsw_vtab.c calls (see muninn_vtab_glue.c). */
#ifndef MUNINN_VTAB_GLUE_H
#define MUNINN_VTAB_GLUE_H
#include "hnsw_algo.h"
int muninn_glue_search(HnswIndex *index, const float *query, int k, int ef_search, HnswSearchResult *results, char **err);
void muninn_glue_mark_dirty(const HnswIndex *index);
void muninn_glue_release(const HnswIndex *index);
#endif
