/*
 * exact_knn.cuh — exact brute-force k-NN with the reference's distance arithmetic.
 *
 * Semantics: the brute_force_knn helper of the reference's tests (test/test_hnsw_algo.c:10-34):
 * all n distances with the index's dist_func, output ascending by distance with ties to the
 * lower row index. Distances are evaluated in the reference's summation order
 * (src/vec_math.c:78-143), so ids AND distances are bit-exact against the CPU oracle.
 *
 * Two kernels per base chunk:
 *   exact_dist_tile   64 queries × 64 rows per CTA, 4×4 pairs per thread, four chain
 *                     accumulators per pair (no FMA: sub/mul/add as the SSE path does);
 *                     writes an fp32 distance panel [nq × chunk_rows].
 *   exact_select      one warp per query scans its panel row (coalesced), filters against the
 *                     current k-th key and merges survivors into a sorted list of 64-bit
 *                     (distance, row) keys.
 * The panel costs 8 B of HBM traffic per pair against ≥ 256 FP32 instructions per pair at
 * dim 128, so the split is ~10 % off a fused version; the CUDA-core roofline applies (bit-exact
 * fp32 chains cannot use the tensor pipe).
 */
#pragma once
#include "muninn_dev.cuh"

namespace muninn {

constexpr int XT = 64, QT = 64, KC = 8;   /* rows, queries, 16-byte chunks per smem stage */
constexpr int KROW = KC * 4 + 4;          /* padded smem row, floats */

template <int METRIC>
__global__ void __launch_bounds__(256)
exact_dist_tile(const float *__restrict__ base, const float *__restrict__ norm2, const uint8_t *__restrict__ deleted,
                int ld, int dim, long long row0, int rows, const float *__restrict__ queries, int nq,
                float *__restrict__ panel, int panel_ld) {
    __shared__ __align__(16) float qs[QT * KROW];
    __shared__ __align__(16) float xs[XT * KROW];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int qb = blockIdx.y * QT, xb = blockIdx.x * XT;
    const int cpr = ld >> 2, cfull = dim >> 2, rem = dim & 3;
    float acc[4][4][4]; /* [query a][row b][chain] */
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[a][b][j] = 0.0f;
    float tail[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) tail[a][b] = 0.0f;
    float qn[4] = {0.f, 0.f, 0.f, 0.f}; /* cosine: Σq² chains folded below */
    float qacc[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int j = 0; j < 4; j++) qacc[a][j] = 0.0f;

    for (int c0 = 0; c0 < cpr; c0 += KC) {
        /* stage KC chunks of 64 query rows and 64 base rows: 2 × 512 float4, 256 threads */
        for (int i = tid; i < QT * KC; i += 256) {
            int r = i / KC, c = i % KC;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            int q = qb + r;
            if (q < nq && c0 + c < cpr) {
                const float *src = queries + (size_t)q * dim + 4 * (c0 + c);
                int left = dim - 4 * (c0 + c);
                v.x = src[0];
                if (left > 1) v.y = src[1];
                if (left > 2) v.z = src[2];
                if (left > 3) v.w = src[3];
            }
            *reinterpret_cast<float4 *>(qs + r * KROW + 4 * c) = v;
        }
        for (int i = tid; i < XT * KC; i += 256) {
            int r = i / KC, c = i % KC;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            int x = xb + r;
            if (x < rows && c0 + c < cpr) v = ldg_stream_f4(base + ((size_t)row0 + x) * ld + 4 * (c0 + c));
            *reinterpret_cast<float4 *>(xs + r * KROW + 4 * c) = v;
        }
        __syncthreads();
#pragma unroll
        for (int c = 0; c < KC; c++) {
            const int cg = c0 + c;
            if (cg >= cpr) break;
            float4 qv[4], xv[4];
#pragma unroll
            for (int a = 0; a < 4; a++) qv[a] = *reinterpret_cast<const float4 *>(qs + (ty * 4 + a) * KROW + 4 * c);
#pragma unroll
            for (int b = 0; b < 4; b++) xv[b] = *reinterpret_cast<const float4 *>(xs + (tx + 16 * b) * KROW + 4 * c);
            if (cg < cfull) {
#pragma unroll
                for (int a = 0; a < 4; a++) {
#pragma unroll
                    for (int b = 0; b < 4; b++) {
                        float4 pr = prod4<METRIC>(qv[a], xv[b]);
                        acc[a][b][0] = __fadd_rn(acc[a][b][0], pr.x);
                        acc[a][b][1] = __fadd_rn(acc[a][b][1], pr.y);
                        acc[a][b][2] = __fadd_rn(acc[a][b][2], pr.z);
                        acc[a][b][3] = __fadd_rn(acc[a][b][3], pr.w);
                    }
                    if (METRIC == METRIC_COSINE) {
                        float4 pq = prod4<METRIC_IP>(qv[a], qv[a]);
                        qacc[a][0] = __fadd_rn(qacc[a][0], pq.x); qacc[a][1] = __fadd_rn(qacc[a][1], pq.y);
                        qacc[a][2] = __fadd_rn(qacc[a][2], pq.z); qacc[a][3] = __fadd_rn(qacc[a][3], pq.w);
                    }
                }
            } else { /* the partial chunk: its first `rem` products are added after the horizontal sum */
#pragma unroll
                for (int a = 0; a < 4; a++) {
#pragma unroll
                    for (int b = 0; b < 4; b++) {
                        float4 pr = prod4<METRIC>(qv[a], xv[b]);
                        /* stash: sequential tail add needs the horizontal sum first; keep the three terms */
                        acc[a][b][0] = __fadd_rn(__fadd_rn(__fadd_rn(acc[a][b][0], acc[a][b][1]), acc[a][b][2]), acc[a][b][3]);
                        acc[a][b][0] = __fadd_rn(acc[a][b][0], pr.x);
                        if (rem > 1) acc[a][b][0] = __fadd_rn(acc[a][b][0], pr.y);
                        if (rem > 2) acc[a][b][0] = __fadd_rn(acc[a][b][0], pr.z);
                        tail[a][b] = 1.0f; /* marks acc[a][b][0] as already final */
                    }
                    if (METRIC == METRIC_COSINE) {
                        float4 pq = prod4<METRIC_IP>(qv[a], qv[a]);
                        qacc[a][0] = __fadd_rn(__fadd_rn(__fadd_rn(qacc[a][0], qacc[a][1]), qacc[a][2]), qacc[a][3]);
                        qacc[a][0] = __fadd_rn(qacc[a][0], pq.x);
                        if (rem > 1) qacc[a][0] = __fadd_rn(qacc[a][0], pq.y);
                        if (rem > 2) qacc[a][0] = __fadd_rn(qacc[a][0], pq.z);
                        qn[a] = 1.0f;
                    }
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < 4; a++) {
        const int q = qb + ty * 4 + a;
        float na = 0.0f;
        if (METRIC == METRIC_COSINE)
            na = qn[a] != 0.0f ? qacc[a][0]
                               : __fadd_rn(__fadd_rn(__fadd_rn(qacc[a][0], qacc[a][1]), qacc[a][2]), qacc[a][3]);
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int x = xb + tx + 16 * b;
            if (q >= nq || x >= rows) continue;
            float sum = tail[a][b] != 0.0f
                            ? acc[a][b][0]
                            : __fadd_rn(__fadd_rn(__fadd_rn(acc[a][b][0], acc[a][b][1]), acc[a][b][2]), acc[a][b][3]);
            float d;
            if (METRIC == METRIC_L2) d = sum;
            else if (METRIC == METRIC_IP) d = -sum;
            else d = cosine_finish(sum, na, norm2[row0 + x]);
            if (deleted && deleted[row0 + x]) d = __int_as_float(0x7fc00000); /* NaN marks a skipped row */
            panel[(size_t)q * panel_ld + x] = d;
        }
    }
}

/* one warp per query: merge a panel row into the running sorted top-k (keys in global memory) */
__global__ void __launch_bounds__(128)
exact_select(const float *__restrict__ panel, int panel_ld, int rows, unsigned row_base, int nq, int k,
             unsigned long long *__restrict__ topk, int *__restrict__ topk_size) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int q = blockIdx.x * 4 + wib;
    unsigned long long *R = reinterpret_cast<unsigned long long *>(smem_raw) + (size_t)wib * (k + 32);
    unsigned long long *S = R + k;
    if (q >= nq) return;
    int size = topk_size[q];
    for (int i = lane; i < size; i += 32) R[i] = topk[(size_t)q * k + i];
    __syncwarp();
    const float *row = panel + (size_t)q * panel_ld;
    int ev = 0, mp;
    for (int x0 = 0; x0 < rows; x0 += 32) {
        int x = x0 + lane;
        float d = x < rows ? row[x] : __int_as_float(0x7fc00000);
        unsigned long long worst = size >= k ? R[k - 1] : KEY_NONE;
        unsigned long long f = (d == d) ? make_key(d, row_base + (unsigned)x) : KEY_NONE;
        bool keep = f < worst;
        unsigned km = __ballot_sync(FULL, keep);
        int ns = __popc(km);
        if (ns == 0) continue;
        if (keep) S[__popc(km & ((1u << lane) - 1))] = f;
        __syncwarp();
        size = warp_merge_sorted(R, size, k, S, ns, lane, &mp, &ev);
    }
    __syncwarp();
    for (int i = lane; i < size; i += 32) topk[(size_t)q * k + i] = R[i];
    if (lane == 0) topk_size[q] = size;
}

/* keys → (index + base, distance, count); list q goes to output row qmap[q] when qmap is given */
__global__ void exact_emit(const unsigned long long *__restrict__ topk, const int *__restrict__ topk_size, int nq, int k,
                           long long index_base, const int *__restrict__ qmap, int64_t *out_idx, float *out_dist,
                           int32_t *out_count) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq * k) return;
    int q = i / k, j = i % k;
    int sz = topk_size[q];
    const int qo = qmap ? qmap[q] : q;
    const size_t o = (size_t)qo * k + j;
    if (j < sz) {
        unsigned long long e = topk[i];
        out_idx[o] = index_base + (long long)(unsigned)e;
        out_dist[o] = okey_inv((unsigned)(e >> 32));
    } else { out_idx[o] = -1; out_dist[o] = 0.0f; }
    if (j == 0) out_count[qo] = sz;
}

/* per-shard result lists → ONE array of 64-bit keys (ordered distance << 32 | global row index), unused tail =
 * KEY_NONE: the payload of the single all-gather of the sharded exact k-NN (nq·k·8 bytes per rank) */
__global__ void topk_pack_keys(int nq, int k, const int64_t *__restrict__ idx, const float *__restrict__ dist,
                               const int32_t *__restrict__ count, unsigned long long *__restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq * k) return;
    const int q = i / k, j = i % k;
    keys[i] = j < count[q] ? make_key(dist[i], (unsigned)idx[i]) : KEY_NONE;
}

/* merge the gathered key lists [parts × nq × k] (each ascending, KEY_NONE padded) per query; one warp per query */
__global__ void __launch_bounds__(128)
topk_merge_keys(int parts, int nq, int k, const unsigned long long *__restrict__ keys, int64_t *out_idx, float *out_dist,
                int32_t *out_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int q = blockIdx.x * 4 + wib;
    unsigned long long *R = reinterpret_cast<unsigned long long *>(smem_raw) + (size_t)wib * (k + 32);
    unsigned long long *S = R + k;
    if (q >= nq) return;
    int size = 0, ev = 0, mp;
    for (int pth = 0; pth < parts; pth++) {
        const unsigned long long *src = keys + ((size_t)pth * nq + q) * k;
        for (int j0 = 0; j0 < k; j0 += 32) {
            const int j = j0 + lane;
            const unsigned long long worst = size >= k ? R[k - 1] : KEY_NONE;
            const unsigned long long f = j < k ? src[j] : KEY_NONE;
            const bool keep = f < worst; /* KEY_NONE never enters */
            const unsigned km = __ballot_sync(FULL, keep);
            const int ns = __popc(km);
            if (ns == 0) break; /* lists ascend: nothing further of this part can enter */
            if (keep) S[__popc(km & ((1u << lane) - 1))] = f;
            __syncwarp();
            size = warp_merge_sorted(R, size, k, S, ns, lane, &mp, &ev);
        }
    }
    __syncwarp();
    for (int j = lane; j < k; j += 32) {
        const size_t o = (size_t)q * k + j;
        if (j < size) { out_idx[o] = (long long)(unsigned)R[j]; out_dist[o] = okey_inv((unsigned)(R[j] >> 32)); }
        else { out_idx[o] = -1; out_dist[o] = 0.0f; }
    }
    if (lane == 0) out_count[q] = size;
}

/* merge `parts` gathered lists per query under (distance, index); one warp per query */
__global__ void __launch_bounds__(128)
topk_merge_parts(int parts, int nq, int k, const int64_t *__restrict__ idx_parts, const float *__restrict__ dist_parts,
                 const int32_t *__restrict__ count_parts, int64_t *out_idx, float *out_dist, int32_t *out_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int q = blockIdx.x * 4 + wib;
    unsigned long long *R = reinterpret_cast<unsigned long long *>(smem_raw) + (size_t)wib * (k + 32);
    unsigned long long *S = R + k;
    if (q >= nq) return;
    int size = 0, ev = 0, mp;
    for (int pth = 0; pth < parts; pth++) {
        const size_t off = ((size_t)pth * nq + q) * k;
        const int cnt = count_parts[(size_t)pth * nq + q];
        for (int j0 = 0; j0 < cnt; j0 += 32) {
            int j = j0 + lane;
            unsigned long long worst = size >= k ? R[k - 1] : KEY_NONE;
            unsigned long long f = j < cnt ? make_key(dist_parts[off + j], (unsigned)idx_parts[off + j]) : KEY_NONE;
            bool keep = f < worst;
            unsigned km = __ballot_sync(FULL, keep);
            int ns = __popc(km);
            if (ns == 0) continue;
            if (keep) S[__popc(km & ((1u << lane) - 1))] = f;
            __syncwarp();
            size = warp_merge_sorted(R, size, k, S, ns, lane, &mp, &ev);
        }
    }
    __syncwarp();
    for (int j = lane; j < k; j += 32) {
        const size_t o = (size_t)q * k + j;
        if (j < size) { out_idx[o] = (long long)(unsigned)R[j]; out_dist[o] = okey_inv((unsigned)(R[j] >> 32)); }
        else { out_idx[o] = -1; out_dist[o] = 0.0f; }
    }
    if (lane == 0) out_count[q] = size;
}

/* per-row Σx² in chain order (the norm_b accumulator of vec_cosine_distance) */
__global__ void row_norm2_kernel(const float *__restrict__ vec, int ld, int dim, long long n, float *__restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = chained_sum_serial<METRIC_IP>(vec + (size_t)i * ld, vec + (size_t)i * ld, dim);
}

/* n independent pairs, one thread each: known-answer tests of test/test_vec_math.c */
__global__ void distance_pairs_kernel(int metric, int dim, long long n, const float *__restrict__ a,
                                      const float *__restrict__ b, float *__restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *x = a + (size_t)i * dim, *y = b + (size_t)i * dim;
    float d;
    if (metric == METRIC_L2) d = chained_sum_serial<METRIC_L2>(x, y, dim);
    else if (metric == METRIC_IP) d = -chained_sum_serial<METRIC_IP>(x, y, dim);
    else d = cosine_finish(chained_sum_serial<METRIC_IP>(x, y, dim), chained_sum_serial<METRIC_IP>(x, x, dim),
                           chained_sum_serial<METRIC_IP>(y, y, dim));
    out[i] = d;
}

} // namespace muninn
