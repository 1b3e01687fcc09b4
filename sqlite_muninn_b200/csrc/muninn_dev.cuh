/*
 * muninn_dev.cuh — device primitives shared by the search and exact-k-NN kernels.
 *
 *  - ordered 64-bit keys (distance, index) so that "ascending by distance, ties to the lower
 *    index" (test/test_hnsw_algo.c:18-31 of the reference) is one integer compare;
 *  - the reference's distance arithmetic (src/vec_math.c:78-143, x86 SSE path): four running
 *    lanes over i mod 4, multiply then add with no FMA, horizontal ((s0+s1)+s2)+s3, scalar
 *    tail — reproduced bit for bit with __fsub_rn/__fmul_rn/__fadd_rn;
 *  - a warp-parallel merge of up to 32 new keys into a sorted shared-memory list.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace muninn {

constexpr unsigned FULL = 0xffffffffu;
constexpr unsigned long long KEY_NONE = ~0ull;

enum { METRIC_L2 = 0, METRIC_COSINE = 1, METRIC_IP = 2 };

/* order-preserving float → uint32 (all finite values and ±inf; -0 is folded into +0 first so
 * that 0 == -0 stays a tie, as it is for the reference's float compares) */
__device__ __forceinline__ uint32_t okey(float d) {
    d = __fadd_rn(d, 0.0f);
    uint32_t b = __float_as_uint(d);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float okey_inv(uint32_t u) {
    uint32_t b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    return __uint_as_float(b);
}
__device__ __forceinline__ unsigned long long make_key(float d, uint32_t low) {
    return ((unsigned long long)okey(d) << 32) | low;
}

__device__ __forceinline__ float4 ldg_stream_f4(const float *p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}

/* per-component product of the reference's inner loop: (a-b)^2 for L2, a*b otherwise */
template <int METRIC>
__device__ __forceinline__ float4 prod4(float4 a, float4 b) {
    float4 p;
    if (METRIC == METRIC_L2) {
        float d0 = __fsub_rn(a.x, b.x), d1 = __fsub_rn(a.y, b.y), d2 = __fsub_rn(a.z, b.z), d3 = __fsub_rn(a.w, b.w);
        p.x = __fmul_rn(d0, d0); p.y = __fmul_rn(d1, d1); p.z = __fmul_rn(d2, d2); p.w = __fmul_rn(d3, d3);
    } else {
        p.x = __fmul_rn(a.x, b.x); p.y = __fmul_rn(a.y, b.y); p.z = __fmul_rn(a.z, b.z); p.w = __fmul_rn(a.w, b.w);
    }
    return p;
}

/* src/vec_math.c:119-125 */
__device__ __forceinline__ float cosine_finish(float dot, float na, float nb) {
    float denom = __fmul_rn(__fsqrt_rn(na), __fsqrt_rn(nb));
    if (denom < 1e-30f) return 1.0f;
    return __fsub_rn(1.0f, __fdiv_rn(dot, denom));
}

/* One thread, one pair: the reference loop as written. Used by the norm / known-answer
 * kernels where throughput does not matter. Returns the raw chained sum of prod4 terms. */
template <int METRIC>
__device__ inline float chained_sum_serial(const float *a, const float *b, int dim) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int i = 0;
    for (; i + 4 <= dim; i += 4) {
        float4 p = prod4<METRIC>(make_float4(a[i], a[i + 1], a[i + 2], a[i + 3]),
                                 make_float4(b[i], b[i + 1], b[i + 2], b[i + 3]));
        s0 = __fadd_rn(s0, p.x); s1 = __fadd_rn(s1, p.y); s2 = __fadd_rn(s2, p.z); s3 = __fadd_rn(s3, p.w);
    }
    float sum = __fadd_rn(__fadd_rn(__fadd_rn(s0, s1), s2), s3);
    for (; i < dim; i++) {
        float p;
        if (METRIC == METRIC_L2) { float d = __fsub_rn(a[i], b[i]); p = __fmul_rn(d, d); }
        else p = __fmul_rn(a[i], b[i]);
        sum = __fadd_rn(sum, p);
    }
    return sum;
}

/*
 * Merge ns (1..32) new keys S[0..ns) (shared memory, any order, all distinct from the
 * list's) into the ascending list R[0..size) of capacity cap, keeping the cap smallest.
 * Whole warp participates. Returns the new size; *min_pos = lowest position written;
 * *evicted_open |= 1 if an element pushed off the end had bit 0 clear.
 * Caller guarantees a __syncwarp() before (S and R visible) — one is issued after.
 */
__device__ __forceinline__ int warp_merge_sorted(unsigned long long *R, int size, int cap, const unsigned long long *S,
                                                 int ns, int lane, int *min_pos, int *evicted_open) {
    /* rank of each new key: lane i < ns owns S[i] */
    unsigned long long f = lane < ns ? S[lane] : KEY_NONE;
    int rF = 0, rR = 0;
    if (lane < ns) {
        for (int i = 0; i < ns; i++) rF += (S[i] < f) ? 1 : 0;
        int lo = 0, hi = size; /* lower bound in R */
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (R[mid] < f) lo = mid + 1; else hi = mid;
        }
        rR = lo;
    }
    int mp = lane < ns ? rR : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mp = min(mp, __shfl_xor_sync(FULL, mp, o));
    /* shift the tail of R upward, highest 32-chunk first so nothing unread is overwritten. (A fast path that moves
     * whole chunks lying above every new key by ns without the per-element count was measured at ef = 256, where this
     * loop is 11 % of the instructions: 6.81 ms with it, 6.73 ms without — the kernel waits on memory there, not on
     * issue slots. profiles/r2_bench/ab_b_geometry_nospec.jsonl) */
    int ev = 0;
    if (size > 0) {
        for (int cb = (size - 1) & ~31; cb >= (mp & ~31); cb -= 32) {
            int idx = cb + lane;
            unsigned long long e = idx < size ? R[idx] : KEY_NONE;
            int cnt = 0;
            if (idx < size && idx >= mp)
                for (int i = 0; i < ns; i++) cnt += (S[i] < e) ? 1 : 0;
            __syncwarp();
            if (cnt > 0) {
                int np = idx + cnt;
                if (np < cap) R[np] = e;
                else if (!(e & 1ull)) ev = 1;
            }
            __syncwarp();
        }
    }
    if (lane < ns) {
        int np = rR + rF;
        if (np < cap) R[np] = f;
        else ev = 1; /* a new key that does not fit is an open candidate the reference would have queued */
    }
    __syncwarp();
    if (__any_sync(FULL, ev)) *evicted_open = 1;
    *min_pos = mp;
    int nsz = size + ns;
    return nsz < cap ? nsz : cap;
}

} // namespace muninn
