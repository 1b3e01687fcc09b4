/*
 * shadow_io.c — bulk reader / writer of the hnsw_index shadow tables, host C, against the SQLite C API
 * (SURVEY.md §8f N1). Part of libmuninn_host.
 *
 * Format = the reference's (/root/reference/src/hnsw_vtab.c:138-199):
 *   <t>_config(key TEXT PRIMARY KEY, value TEXT)    dimensions, metric (int), m, ef_construction, entry_point, max_level
 *   <t>_nodes(id INTEGER PRIMARY KEY, vector BLOB, level, deleted)
 *   <t>_edges(source_id, target_id, level, distance, PRIMARY KEY(source_id, level, target_id)) WITHOUT ROWID
 *   index <t>_edges_rev(target_id, level)
 *
 * muninn_shadow_load replaces load_index_from_shadow (src/hnsw_vtab.c:286-341) for bulk use: the reference rebuilds
 * its pointer-per-node index row by row, with an ht_find per row and a linear duplicate scan per edge
 * (src/hnsw_algo.c:147-150); here two ordered scans fill the flat arrays muninn_cuda_index_upload takes — every stored
 * node (deleted ones included), an edge kept only when its source exists and level <= the source's level (:334),
 * lists in ascending target_id order (the order the reference's reload produces), a target without a node row kept
 * as -1 (the reference keeps the dangling id and skips it at search time).
 *
 * muninn_shadow_write replaces the per-insert persistence of hnsw_vtab_update (src/hnsw_vtab.c:237-283: one node row,
 * ~2M edge upserts and a config rewrite PER INSERT) with one transaction of prepared statements in primary-key order,
 * and can register the table in sqlite_schema so that the UNMODIFIED extension xConnects to it.
 *
 * The SQLite entry points are declared here (public, stable C ABI) rather than taken from a header: the image has
 * libsqlite3.so.0 but no sqlite3.h outside the reference tree, and nothing of the reference is compiled into the product.
 */
#include "../../include/muninn_shadow.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ── SQLite C API subset (sqlite3.h) ── */
typedef struct sqlite3_stmt sqlite3_stmt;
typedef long long sqlite3_int64;
#define SQLITE_OK 0
#define SQLITE_ROW 100
#define SQLITE_DONE 101
#define SQLITE_OPEN_READWRITE 0x00000002
#define SQLITE_OPEN_CREATE 0x00000004
#define SQLITE_OPEN_READONLY 0x00000001
#define SQLITE_STATIC ((void (*)(void *))0)
int sqlite3_open_v2(const char *, sqlite3 **, int, const char *);
int sqlite3_close(sqlite3 *);
int sqlite3_exec(sqlite3 *, const char *, int (*)(void *, int, char **, char **), void *, char **);
int sqlite3_prepare_v2(sqlite3 *, const char *, int, sqlite3_stmt **, const char **);
int sqlite3_step(sqlite3_stmt *);
int sqlite3_reset(sqlite3_stmt *);
int sqlite3_finalize(sqlite3_stmt *);
int sqlite3_bind_int64(sqlite3_stmt *, int, sqlite3_int64);
int sqlite3_bind_int(sqlite3_stmt *, int, int);
int sqlite3_bind_double(sqlite3_stmt *, int, double);
int sqlite3_bind_blob(sqlite3_stmt *, int, const void *, int, void (*)(void *));
int sqlite3_bind_text(sqlite3_stmt *, int, const char *, int, void (*)(void *));
sqlite3_int64 sqlite3_column_int64(sqlite3_stmt *, int);
int sqlite3_column_int(sqlite3_stmt *, int);
const void *sqlite3_column_blob(sqlite3_stmt *, int);
int sqlite3_column_bytes(sqlite3_stmt *, int);
const unsigned char *sqlite3_column_text(sqlite3_stmt *, int);
const char *sqlite3_errmsg(sqlite3 *);
char *sqlite3_mprintf(const char *, ...);
void sqlite3_free(void *);

static int set_err(char **err, const char *fmt, const char *a, const char *b) {
    if (err) {
        char buf[512];
        snprintf(buf, sizeof buf, fmt, a ? a : "", b ? b : "");
        const size_t len = strlen(buf) + 1;
        *err = malloc(len);
        if (*err) memcpy(*err, buf, len);
    }
    return MUNINN_SHADOW_E_SQL;
}

void muninn_shadow_free_message(char *msg) { free(msg); }

void muninn_shadow_free(MuninnShadowGraph *g) {
    if (!g) return;
    free(g->ids); free(g->levels); free(g->deleted); free(g->vectors); free(g->list_offsets); free(g->targets);
    memset(g, 0, sizeof *g);
}

static int prep(sqlite3 *db, sqlite3_stmt **st, char **err, const char *fmt, const char *table) {
    char *sql = sqlite3_mprintf(fmt, table, table);
    if (!sql) return MUNINN_SHADOW_E_NOMEM;
    int rc = sqlite3_prepare_v2(db, sql, -1, st, NULL);
    sqlite3_free(sql);
    if (rc != SQLITE_OK) return set_err(err, "hnsw_index shadow tables of %s: %s", table, sqlite3_errmsg(db));
    return MUNINN_SHADOW_OK;
}

static int exec_fmt(sqlite3 *db, char **err, const char *fmt, const char *table) {
    char *sql = sqlite3_mprintf(fmt, table, table);
    if (!sql) return MUNINN_SHADOW_E_NOMEM;
    int rc = sqlite3_exec(db, sql, NULL, NULL, NULL);
    sqlite3_free(sql);
    if (rc != SQLITE_OK) return set_err(err, "hnsw_index shadow tables of %s: %s", table, sqlite3_errmsg(db));
    return MUNINN_SHADOW_OK;
}

/* slot of rowid `id` in the ascending ids[0..n), or -1 */
static int64_t slot_of(const int64_t *ids, int64_t n, int64_t id) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (ids[mid] < id) lo = mid + 1; else hi = mid;
    }
    return (lo < n && ids[lo] == id) ? lo : -1;
}

int muninn_shadow_load(sqlite3 *db, const char *table, MuninnShadowGraph *g, char **err) {
    if (err) *err = NULL;
    if (!db || !table || !g) return MUNINN_SHADOW_E_ARG;
    memset(g, 0, sizeof *g);
    g->entry_slot = -1; g->max_level = -1;
    sqlite3_stmt *st = NULL;
    int rc;
    /* ── config (src/hnsw_vtab.c:296-315) ── */
    int64_t entry_id = -1;
    int have = 0;
    if ((rc = prep(db, &st, err, "SELECT key, value FROM \"%w_config\"", table)) != 0) return rc;
    while (sqlite3_step(st) == SQLITE_ROW) {
        const char *k = (const char *)sqlite3_column_text(st, 0), *v = (const char *)sqlite3_column_text(st, 1);
        if (!k || !v) continue;
        if (!strcmp(k, "dimensions")) { g->dim = atoi(v); have |= 1; }
        else if (!strcmp(k, "metric")) { g->metric = atoi(v); have |= 2; }
        else if (!strcmp(k, "m")) { g->M = atoi(v); have |= 4; }
        else if (!strcmp(k, "ef_construction")) { g->ef_construction = atoi(v); have |= 8; }
        else if (!strcmp(k, "entry_point")) entry_id = atoll(v);
        else if (!strcmp(k, "max_level")) g->max_level = atoi(v);
    }
    sqlite3_finalize(st);
    if ((have & 7) != 7 || g->dim < 1) return set_err(err, "%s_config lacks dimensions / metric / m%s", table, "");
    /* ── nodes, ascending id = slot order ── */
    if ((rc = prep(db, &st, err, "SELECT count(*) FROM \"%w_nodes\"", table)) != 0) return rc;
    int64_t n = sqlite3_step(st) == SQLITE_ROW ? sqlite3_column_int64(st, 0) : 0;
    sqlite3_finalize(st);
    const size_t n1 = n > 0 ? (size_t)n : 1;
    g->ids = malloc(n1 * sizeof(int64_t)); g->levels = malloc(n1 * sizeof(int32_t)); g->deleted = malloc(n1);
    g->vectors = malloc(n1 * (size_t)g->dim * sizeof(float));
    if (!g->ids || !g->levels || !g->deleted || !g->vectors) { muninn_shadow_free(g); return MUNINN_SHADOW_E_NOMEM; }
    if ((rc = prep(db, &st, err, "SELECT id, level, deleted, vector FROM \"%w_nodes\" ORDER BY id", table)) != 0) { muninn_shadow_free(g); return rc; }
    int64_t at = 0, n_lists = 0;
    while (sqlite3_step(st) == SQLITE_ROW && at < n) {
        g->ids[at] = sqlite3_column_int64(st, 0);
        g->levels[at] = sqlite3_column_int(st, 1);
        g->deleted[at] = sqlite3_column_int(st, 2) != 0;
        if (sqlite3_column_bytes(st, 3) != g->dim * 4 || g->levels[at] < 0 || g->levels[at] >= 32) {
            sqlite3_finalize(st); muninn_shadow_free(g);
            return set_err(err, "%s_nodes: a row has a vector blob of the wrong size or a level outside 0..31%s", table, "");
        }
        memcpy(g->vectors + (size_t)at * g->dim, sqlite3_column_blob(st, 3), (size_t)g->dim * 4);
        n_lists += g->levels[at] + 1;
        at++;
    }
    sqlite3_finalize(st);
    g->n = n = at;
    g->n_lists = n_lists;
    /* ── edges, one ordered scan: (source, level, target) order IS list order ── */
    if ((rc = prep(db, &st, err, "SELECT count(*) FROM \"%w_edges\"", table)) != 0) { muninn_shadow_free(g); return rc; }
    int64_t cap_edges = sqlite3_step(st) == SQLITE_ROW ? sqlite3_column_int64(st, 0) : 0;
    sqlite3_finalize(st);
    g->list_offsets = calloc((size_t)n_lists + 1, sizeof(int64_t));
    g->targets = malloc((cap_edges > 0 ? (size_t)cap_edges : 1) * sizeof(int32_t));
    int64_t *first = malloc((n1 + 1) * sizeof(int64_t));
    if (!g->list_offsets || !g->targets || !first) { free(first); muninn_shadow_free(g); return MUNINN_SHADOW_E_NOMEM; }
    first[0] = 0;
    for (int64_t s = 0; s < n; s++) first[s + 1] = first[s] + g->levels[s] + 1;
    if ((rc = prep(db, &st, err, "SELECT source_id, level, target_id FROM \"%w_edges\" ORDER BY source_id, level, target_id", table)) != 0) {
        free(first); muninn_shadow_free(g); return rc;
    }
    int64_t ne = 0, cur_src_id = 0, cur_src = -1;
    int have_src = 0;
    while (sqlite3_step(st) == SQLITE_ROW && ne < cap_edges) {
        const int64_t sid = sqlite3_column_int64(st, 0), tid = sqlite3_column_int64(st, 2);
        const int lvl = sqlite3_column_int(st, 1);
        if (!have_src || sid != cur_src_id) { cur_src_id = sid; cur_src = slot_of(g->ids, n, sid); have_src = 1; }
        if (cur_src < 0 || lvl < 0 || lvl > g->levels[cur_src]) continue; /* src/hnsw_vtab.c:334 */
        g->list_offsets[first[cur_src] + lvl + 1]++;
        g->targets[ne++] = (int32_t)slot_of(g->ids, n, tid);
    }
    sqlite3_finalize(st);
    for (int64_t li = 0; li < n_lists; li++) g->list_offsets[li + 1] += g->list_offsets[li];
    g->n_edges = ne;
    free(first);
    if (entry_id != -1 && n > 0) g->entry_slot = (int32_t)slot_of(g->ids, n, entry_id);
    return MUNINN_SHADOW_OK;
}

int muninn_shadow_write(sqlite3 *db, const char *table, const MuninnShadowGraph *g, int register_vtab, char **err) {
    if (err) *err = NULL;
    if (!db || !table || !g || g->n < 0 || g->dim < 1) return MUNINN_SHADOW_E_ARG;
    int rc;
    sqlite3_stmt *st = NULL;
    if (sqlite3_exec(db, "BEGIN", NULL, NULL, NULL) != SQLITE_OK) return set_err(err, "%s: %s", table, sqlite3_errmsg(db));
#define FAIL(code) do { int c_ = (code); if (st) sqlite3_finalize(st); sqlite3_exec(db, "ROLLBACK", NULL, NULL, NULL); return c_; } while (0)
    /* create_shadow_tables, src/hnsw_vtab.c:138-181 (the reverse index is built once at the end, over the finished table) */
    if ((rc = exec_fmt(db, err, "CREATE TABLE IF NOT EXISTS \"%w_config\" (key TEXT PRIMARY KEY, value TEXT NOT NULL)", table))) FAIL(rc);
    if ((rc = exec_fmt(db, err, "CREATE TABLE IF NOT EXISTS \"%w_nodes\" (id INTEGER PRIMARY KEY, vector BLOB NOT NULL, level INTEGER NOT NULL, deleted INTEGER NOT NULL DEFAULT 0)", table))) FAIL(rc);
    if ((rc = exec_fmt(db, err, "CREATE TABLE IF NOT EXISTS \"%w_edges\" (source_id INTEGER NOT NULL, target_id INTEGER NOT NULL, level INTEGER NOT NULL, distance REAL NOT NULL, PRIMARY KEY (source_id, level, target_id)) WITHOUT ROWID", table))) FAIL(rc);
    /* save_config, src/hnsw_vtab.c:183-199 */
    {
        const long long entry_id = g->entry_slot >= 0 && g->entry_slot < g->n ? (long long)g->ids[g->entry_slot] : -1;
        char *sql = sqlite3_mprintf("INSERT OR REPLACE INTO \"%w_config\" (key, value) VALUES ('dimensions', '%d'), ('metric', '%d'), ('m', '%d'), "
                                    "('ef_construction', '%d'), ('entry_point', '%lld'), ('max_level', '%d')",
                                    table, g->dim, g->metric, g->M, g->ef_construction, entry_id, g->max_level);
        if (!sql) FAIL(MUNINN_SHADOW_E_NOMEM);
        rc = sqlite3_exec(db, sql, NULL, NULL, NULL);
        sqlite3_free(sql);
        if (rc != SQLITE_OK) FAIL(set_err(err, "%s_config: %s", table, sqlite3_errmsg(db)));
    }
    /* nodes in rowid order: appends to the table b-tree */
    if ((rc = prep(db, &st, err, "INSERT INTO \"%w_nodes\" (id, vector, level, deleted) VALUES (?, ?, ?, ?)", table))) FAIL(rc);
    for (int64_t s = 0; s < g->n; s++) {
        sqlite3_bind_int64(st, 1, g->ids[s]);
        sqlite3_bind_blob(st, 2, g->vectors + (size_t)s * g->dim, g->dim * 4, SQLITE_STATIC);
        sqlite3_bind_int(st, 3, g->levels[s]);
        sqlite3_bind_int(st, 4, g->deleted ? g->deleted[s] != 0 : 0);
        if (sqlite3_step(st) != SQLITE_DONE) FAIL(set_err(err, "%s_nodes: %s", table, sqlite3_errmsg(db)));
        sqlite3_reset(st);
    }
    sqlite3_finalize(st); st = NULL;
    /* edges in primary-key order (source, level, target): lists are node-major, level-minor; inside a list the targets are
     * written in ascending rowid order whatever their order in memory, so the WITHOUT ROWID b-tree only ever appends */
    if ((rc = prep(db, &st, err, "INSERT OR IGNORE INTO \"%w_edges\" (source_id, target_id, level, distance) VALUES (?, ?, ?, 0.0)", table))) FAIL(rc);
    {
        int64_t li = 0, cap = 0;
        int64_t *tmp = NULL;
        for (int64_t s = 0; s < g->n; s++) {
            for (int l = 0; l <= g->levels[s]; l++, li++) {
                const int64_t b = g->list_offsets[li], e = g->list_offsets[li + 1];
                if (e - b > cap) { cap = (e - b) * 2; free(tmp); tmp = malloc((size_t)cap * sizeof(int64_t)); if (!tmp) FAIL(MUNINN_SHADOW_E_NOMEM); }
                int64_t m = 0;
                for (int64_t j = b; j < e; j++) if (g->targets[j] >= 0 && g->targets[j] < g->n) tmp[m++] = g->ids[g->targets[j]];
                for (int64_t i = 1; i < m; i++) { /* insertion sort: lists hold at most a few dozen entries */
                    const int64_t v = tmp[i];
                    int64_t p = i;
                    while (p > 0 && tmp[p - 1] > v) { tmp[p] = tmp[p - 1]; p--; }
                    tmp[p] = v;
                }
                for (int64_t i = 0; i < m; i++) {
                    sqlite3_bind_int64(st, 1, g->ids[s]);
                    sqlite3_bind_int64(st, 2, tmp[i]);
                    sqlite3_bind_int(st, 3, l);
                    if (sqlite3_step(st) != SQLITE_DONE) { free(tmp); FAIL(set_err(err, "%s_edges: %s", table, sqlite3_errmsg(db))); }
                    sqlite3_reset(st);
                }
            }
        }
        free(tmp);
    }
    sqlite3_finalize(st); st = NULL;
    if ((rc = exec_fmt(db, err, "CREATE INDEX IF NOT EXISTS \"%w_edges_rev\" ON \"%w_edges\"(target_id, level)", table))) FAIL(rc);
    if (register_vtab) {
        /* the schema row of `CREATE VIRTUAL TABLE <t> USING hnsw_index(...)` (arguments as parsed at src/hnsw_vtab.c:80-134),
         * stored directly: running xCreate would start an empty index over the tables just written */
        static const char *names[3] = {"l2", "cosine", "inner_product"};
        if (g->metric < 0 || g->metric > 2) FAIL(MUNINN_SHADOW_E_ARG);
        char *decl = sqlite3_mprintf("CREATE VIRTUAL TABLE \"%w\" USING hnsw_index(dimensions=%d, metric='%s', m=%d, ef_construction=%d)",
                                     table, g->dim, names[g->metric], g->M, g->ef_construction);
        if (!decl) FAIL(MUNINN_SHADOW_E_NOMEM);
        rc = sqlite3_exec(db, "PRAGMA writable_schema=ON", NULL, NULL, NULL);
        if (rc == SQLITE_OK) {
            rc = sqlite3_prepare_v2(db, "INSERT INTO sqlite_master(type, name, tbl_name, rootpage, sql) VALUES('table', ?, ?, 0, ?)", -1, &st, NULL);
            if (rc == SQLITE_OK) {
                sqlite3_bind_text(st, 1, table, -1, SQLITE_STATIC);
                sqlite3_bind_text(st, 2, table, -1, SQLITE_STATIC);
                sqlite3_bind_text(st, 3, decl, -1, SQLITE_STATIC);
                rc = sqlite3_step(st) == SQLITE_DONE ? SQLITE_OK : 1;
                sqlite3_finalize(st); st = NULL;
            }
            sqlite3_exec(db, "PRAGMA writable_schema=OFF", NULL, NULL, NULL);
        }
        sqlite3_free(decl);
        if (rc != SQLITE_OK) FAIL(set_err(err, "registering %s in sqlite_schema: %s", table, sqlite3_errmsg(db)));
    }
#undef FAIL
    if (sqlite3_exec(db, "COMMIT", NULL, NULL, NULL) != SQLITE_OK) return set_err(err, "%s: %s", table, sqlite3_errmsg(db));
    return MUNINN_SHADOW_OK;
}

int muninn_shadow_load_path(const char *db_path, const char *table, MuninnShadowGraph *g, char **err) {
    if (err) *err = NULL;
    sqlite3 *db = NULL;
    if (!db_path) return MUNINN_SHADOW_E_ARG;
    if (sqlite3_open_v2(db_path, &db, SQLITE_OPEN_READONLY, NULL) != SQLITE_OK) {
        int rc = set_err(err, "cannot open %s: %s", db_path, db ? sqlite3_errmsg(db) : "out of memory");
        if (db) sqlite3_close(db);
        return rc;
    }
    int rc = muninn_shadow_load(db, table, g, err);
    sqlite3_close(db);
    return rc;
}

int muninn_shadow_write_path(const char *db_path, const char *table, const MuninnShadowGraph *g, int register_vtab, char **err) {
    if (err) *err = NULL;
    sqlite3 *db = NULL;
    if (!db_path) return MUNINN_SHADOW_E_ARG;
    if (sqlite3_open_v2(db_path, &db, SQLITE_OPEN_READWRITE | SQLITE_OPEN_CREATE, NULL) != SQLITE_OK) {
        int rc = set_err(err, "cannot open %s: %s", db_path, db ? sqlite3_errmsg(db) : "out of memory");
        if (db) sqlite3_close(db);
        return rc;
    }
    sqlite3_exec(db, "PRAGMA journal_mode=OFF; PRAGMA synchronous=OFF; PRAGMA cache_size=-262144", NULL, NULL, NULL); /* a bulk load of a new file */
    int rc = muninn_shadow_write(db, table, g, register_vtab, err);
    sqlite3_close(db);
    return rc;
}
