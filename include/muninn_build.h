/*
 * muninn_build.h — C ABI of libmuninn_host: host-side HNSW graph construction in the reference's format
 * (sqlite_muninn_b200/csrc/hnsw_build.c). SURVEY.md §8f N2; paths relative to /root/reference/.
 *
 * What it stands in for: the loop `for each row: hnsw_insert(index, id, vector)` (src/hnsw_algo.c:520-666, called
 * from xUpdate at src/hnsw_vtab.c:748) when an index is filled in bulk. The output is the flat "list CSR" that
 * muninn_cuda_index_upload (include/muninn_cuda.h) and the shadow-table writer take.
 *
 *   exact mode   muninn_build_insert(h, count, 0, 0, 1): sequential inserts; entry point, levels and ordered
 *                neighbour lists are the reference's own, bit for bit (level generator src/hnsw_algo.c:19-30,
 *                :240-248 with its seed 42; search :257-282, :347-448; select :497-516; link + MN-RU prune :582-646).
 *   batch mode   muninn_build_insert(h, count, batch, warm, threads): after `warm` exact inserts, groups of `batch`
 *                nodes search the graph as it stood before the group (in parallel) and are then linked in id order.
 *                Deterministic for a given batch size, independent of the thread count.
 *   group mode   the batch mode one group at a time, for a caller that runs the layer-0 candidate searches itself
 *                (one muninn_cuda_search_batch(k = min(2M, efc), ef = efc) per group on a device copy of the graph)
 *                and mirrors the changed records with muninn_cuda_index_patch:
 *                    muninn_build_log_touched(h, 1);
 *                    while (muninn_build_count(h) < capacity) {
 *                        muninn_build_group_begin(h, batch, &g0, &gn);     // levels drawn, nodes allocated
 *                        flush();                                          // a level-raising node may have been inserted
 *                        if (!gn) continue;
 *                        muninn_build_group_levels(h, levels);             // search the level-0 nodes on the device
 *                        muninn_build_group_link(h, cand, cand_cnt, stride, threads);
 *                        flush();      // muninn_build_touched_take → muninn_build_export_nodes → muninn_cuda_index_patch
 *                    }
 *                (sqlite_muninn_b200/builder.py build_graph_device is this loop.) Same graph as the batch mode
 *                wherever no two candidate distances are bit-equal.
 *
 * Slots are dense indices in insertion order (= ascending rowid order of the caller's ids). Plain C types only.
 * Not thread-safe per handle; separate handles are independent.
 */
#ifndef MUNINN_BUILD_H
#define MUNINN_BUILD_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* hnsw_create (src/hnsw_algo.h:56): `vectors` is the caller's [capacity × dim] fp32 array, read in place for the
 * builder's lifetime; metric as VecMetric (0 l2, 1 cosine, 2 inner product). NULL on allocation failure. */
void *muninn_build_create(int dim, int metric, int M, int ef_construction, int64_t capacity, const float *vectors);
void muninn_build_destroy(void *h);
/* Level generator seed (the reference's is 42; 0 is mapped to 1 as in src/hnsw_algo.c:222-224). */
void muninn_build_seed(void *h, uint32_t seed);
/* Two-rows-per-call AVX2 distances on (1) / off (0) / as detected (-1); returns what is in effect. Results do not
 * depend on it. */
int muninn_build_set_avx2(int on);

/* Insert the next `count` vectors. batch = 0: exact mode; batch > 0: batch mode once `warm` nodes are stored.
 * 0 on success, -1 if it would exceed the capacity or a group is open. */
int muninn_build_insert(void *h, int64_t count, int batch, int64_t warm, int nthreads);
/* Nodes stored so far. */
int64_t muninn_build_count(void *h);

/* ── group mode ── */
/* Record which nodes' lists change (the new nodes and every node they link themselves to). */
void muninn_build_log_touched(void *h, int on);
/* Open the next group [*g0, *g0 + *gn): levels drawn, nodes allocated. *gn may be 0 (a level-raising node was
 * inserted exactly, or nothing is left). -1 if a group is already open. */
int muninn_build_group_begin(void *h, int batch, int64_t *g0, int64_t *gn);
/* Levels of the open group's nodes: levels[gn]. */
int muninn_build_group_levels(void *h, int32_t *levels);
/* Link the open group. Node g0+i with cand_cnt[i] >= 0 (level-0 nodes only) takes the first min(cand_cnt[i], 2M)
 * slots of cand[i*cand_stride ..] — the caller's search result, ascending by (distance, slot) — as its layer-0
 * selection; every other node searches on the host. cand may be NULL. -1: no open group; -2: a candidate slot is
 * not a node stored before the group. */
int muninn_build_group_link(void *h, const int32_t *cand, const int32_t *cand_cnt, int cand_stride, int nthreads);
/* The log: each touched slot once, in no particular order; _take copies and clears it. */
int64_t muninn_build_touched_count(void *h);
int64_t muninn_build_touched_take(void *h, int32_t *out, int64_t cap);
/* Records of the given nodes in the layout of muninn_cuda_index_patch (lists in the builder's own order).
 * targets = NULL: sizes only. */
void muninn_build_export_nodes(void *h, int64_t n_nodes, const int32_t *slots, int32_t *levels, int64_t *list_offsets,
                               int32_t *targets, int64_t *n_lists, int64_t *n_edges);
void muninn_build_entry(void *h, int32_t *entry, int32_t *max_level);
/* Seconds spent so far in the group linker: [0] candidate searches + own lists, [1] request sort, [2] reverse
 * edges + MN-RU prune. */
void muninn_build_phase_seconds(void *h, double out[3]);

/* ── the group loop in C ──
 * The searcher a group loop drives: a device copy of the graph behind three calls with the argument meaning of
 * muninn_cuda_index_upload / muninn_cuda_index_patch (no deleted flags: a graph under construction has none) /
 * muninn_cuda_search_batch (no stats). Each returns 0 on success. muninn_cuda_build_device (include/muninn_cuda.h)
 * fills one in for a MuninnCudaIndex; the CPU tests fill one in with callbacks that answer from the oracle. */
typedef struct MuninnBuildDevice {
    void *ctx;
    int (*upload)(void *ctx, int64_t n, const int64_t *ids, const int32_t *levels, const float *vectors,
                  const int64_t *list_offsets, const int32_t *targets, int32_t entry_slot, int32_t max_level);
    int (*patch)(void *ctx, int64_t n_total, int32_t n_patch, const int32_t *slots, const int64_t *ids,
                 const int32_t *levels, const float *vectors, const int64_t *list_offsets, const int32_t *targets,
                 int32_t entry_slot, int32_t max_level);
    int (*search)(void *ctx, const float *queries, int nq, int k, int ef, int64_t *out_ids, float *out_dist,
                  int32_t *out_count);
} MuninnBuildDevice;

typedef struct {
    int64_t groups, device_queries, host_searches, patches, patched_records;
    double warm_s, search_s, link_s, patch_s; /* wall seconds: exact warm-up inserts + upload, device searches,
                                                 host linking, record export + patch */
} MuninnBuildRunStats;

/* Fill the (empty) builder to its capacity with the group loop sketched at the top of this file: `warm` exact
 * inserts, one upload, then per group one dev->search over its level-0 nodes and one dev->patch with the touched
 * records (new nodes first; `vectors` of a patch points at the first new node's row). ids[capacity] are the rowids,
 * strictly ascending (NULL = 0..capacity-1). Returns 0, -1 on a bad argument / allocation failure / a candidate
 * outside the graph, or the device call's own non-zero code. st may be NULL. There is no host fallback: a failing
 * device call ends the build. */
int muninn_build_run(void *h, const int64_t *ids, int batch, int64_t warm, int nthreads, const MuninnBuildDevice *dev,
                     MuninnBuildRunStats *st);

/* ── whole-graph export (the layout of muninn_cuda_index_upload) ── */
void muninn_build_sizes(void *h, int64_t *n, int64_t *n_lists, int64_t *n_edges, int32_t *entry, int32_t *max_level);
/* sort_lists = 1: neighbour lists ascending by target, the order load_index_from_shadow rebuilds
 * (src/hnsw_vtab.c:163-169, :322); 0: the builder's own (insertion) order. */
void muninn_build_export(void *h, int32_t *levels, int64_t *list_offsets, int32_t *targets, int sort_lists);

#ifdef __cplusplus
}
#endif
#endif /* MUNINN_BUILD_H */
