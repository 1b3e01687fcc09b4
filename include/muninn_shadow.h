/*
 * muninn_shadow.h — bulk I/O of the hnsw_index shadow tables in host C (libmuninn_host), SURVEY.md §8f N1.
 *
 * Replaces, for bulk use, two reference paths (paths relative to /root/reference/):
 *   muninn_shadow_load    load_index_from_shadow                      src/hnsw_vtab.c:286-341
 *   muninn_shadow_write   the per-insert persistence of xUpdate       src/hnsw_vtab.c:237-283
 *                         + create_shadow_tables / save_config        src/hnsw_vtab.c:138-199
 * The tables they read and write are the reference's own, byte for byte: the unmodified extension opens, queries and
 * extends (its own INSERT / DELETE) what muninn_shadow_write stored, and muninn_shadow_load reads what the reference
 * persisted. The loaded form is the flat "list CSR" muninn_cuda_index_upload (include/muninn_cuda.h) takes, so
 *   muninn_shadow_load(db, "t", &g, &err);  muninn_cuda_index_upload(h, g.n, g.ids, g.levels, g.deleted, g.vectors,
 *                                                                    g.list_offsets, g.targets, g.entry_slot, g.max_level);
 * brings a persisted index onto the device without building the reference's in-memory index first.
 */
#ifndef MUNINN_SHADOW_H
#define MUNINN_SHADOW_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sqlite3 sqlite3; /* the host's database handle (sqlite3.h) */

enum { MUNINN_SHADOW_OK = 0, MUNINN_SHADOW_E_ARG = -1, MUNINN_SHADOW_E_NOMEM = -2, MUNINN_SHADOW_E_SQL = -3 };

/* One stored index, flat. Arrays are owned by the struct after a load (muninn_shadow_free); for a write the caller
 * points them at its own memory. Slots are ranks in ascending rowid order; lists are node-major, level-minor. */
typedef struct {
    int32_t dim, metric, M, ef_construction; /* _config: dimensions, metric (VecMetric as int), m, ef_construction */
    int64_t n, n_lists, n_edges;             /* stored nodes (deleted included), Σ(level+1), edges kept */
    int64_t *ids;                            /* [n] _nodes.id ascending */
    int32_t *levels;                         /* [n] _nodes.level */
    uint8_t *deleted;                        /* [n] _nodes.deleted (may be NULL for a write: none) */
    float *vectors;                          /* [n × dim] _nodes.vector */
    int64_t *list_offsets;                   /* [n_lists + 1] */
    int32_t *targets;                        /* [n_edges] slots, -1 = edge to a rowid that has no node row */
    int32_t entry_slot, max_level;           /* _config.entry_point as a slot (-1 = empty index), _config.max_level */
} MuninnShadowGraph;

/* *err (if given) receives a malloc'ed message on failure: release with muninn_shadow_free_message. */
int muninn_shadow_load(sqlite3 *db, const char *table, MuninnShadowGraph *out, char **err);
/* One transaction: tables, config, nodes and edges in primary-key order, the reverse index built once at the end.
 * register_vtab != 0 also stores `CREATE VIRTUAL TABLE <table> USING hnsw_index(...)` in sqlite_schema (without running
 * xCreate), so the next connection with the extension loaded xConnects to the stored graph. The tables must not hold
 * rows of another index already. */
int muninn_shadow_write(sqlite3 *db, const char *table, const MuninnShadowGraph *g, int register_vtab, char **err);
/* The same on a database file (opened read-only / created if missing; the writer switches the journal off: a bulk
 * load of a new file). */
int muninn_shadow_load_path(const char *db_path, const char *table, MuninnShadowGraph *out, char **err);
int muninn_shadow_write_path(const char *db_path, const char *table, const MuninnShadowGraph *g, int register_vtab, char **err);
void muninn_shadow_free(MuninnShadowGraph *g);
void muninn_shadow_free_message(char *msg);

#ifdef __cplusplus
}
#endif
#endif /* MUNINN_SHADOW_H */
