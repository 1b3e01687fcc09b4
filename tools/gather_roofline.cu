// gather_roofline.cu — ceiling of the search kernel's access pattern on this GPU: random ROW gathers
// (row = dim fp32, 16-byte aligned) from a slab much larger than L2, as many loads in flight as the
// SM can hold, no dependent work. Build & run on the box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/gather tools/gather_roofline.cu && /tmp/gather
// Prints GB/s per (row bytes, rows in flight per warp). The streaming-copy peak in MEASURED_PEAKS.json is
// the contract's denominator; this number says how much of it a perfect random-row gather can reach.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

template <int ROWS_IN_FLIGHT>
__global__ void gather(const float4 *__restrict__ slab, long long n_rows, int chunks_per_row, int iters, float *out) {
    const int lane = threadIdx.x & 31;
    unsigned long long s = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 0x9E3779B97F4A7C15ull + 12345;
    float acc = 0.f;
    for (int it = 0; it < iters; it++) {
        float4 v[ROWS_IN_FLIGHT];
#pragma unroll
        for (int r = 0; r < ROWS_IN_FLIGHT; r++) {
            s = s * 6364136223846793005ull + 1442695040888963407ull;
            const long long row = (long long)((s >> 24) & (unsigned long long)(n_rows - 1)); /* n_rows is a power of two */
            const float4 *p = slab + row * chunks_per_row;
            float4 a = make_float4(0, 0, 0, 0);
            for (int c = lane; c < chunks_per_row; c += 32) {
                float4 t;
                asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "l"(p + c));
                a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
            }
            v[r] = a;
        }
#pragma unroll
        for (int r = 0; r < ROWS_IN_FLIGHT; r++) acc += v[r].x + v[r].y + v[r].z + v[r].w;
    }
    if (acc == 123.456f) out[0] = acc;
}

template <int R>
static void run(const float4 *slab, long long n_rows, int dim, int warps_per_cta, int ctas_per_sm, float *out) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int chunks = dim / 4, iters = 400 * 8 / R;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = sms * ctas_per_sm, block = 32 * warps_per_cta;
    gather<R><<<grid, block>>>(slab, n_rows, chunks, 20, out);
    cudaEventRecord(e0);
    gather<R><<<grid, block>>>(slab, n_rows, chunks, iters, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = (double)grid * warps_per_cta * iters * R * dim * 4.0;
    printf("{\"row_bytes\": %d, \"rows_in_flight_per_warp\": %d, \"warps_per_sm\": %d, \"GBps\": %.1f, \"ms\": %.3f}\n", dim * 4, R,
           warps_per_cta * ctas_per_sm, bytes / ms / 1e6, ms);
}

int main() {
    const size_t max_bytes = (size_t)4 << 30;
    float4 *slab; float *out;
    cudaMalloc(&slab, max_bytes); cudaMalloc(&out, 4);
    cudaMemset(slab, 0, max_bytes);
    for (size_t slab_bytes : {(size_t)512 << 20, (size_t)4 << 30}) {
        for (int dim : {96, 128, 768}) {
            long long n_rows = 1;
            while ((size_t)(2 * n_rows) * dim * 4 <= slab_bytes) n_rows *= 2;
            printf("{\"slab_MiB\": %zu, \"rows\": %lld}\n", (size_t)n_rows * dim * 4 >> 20, n_rows);
            for (int wps : {16, 32, 64}) {
                const int wpc = 8, cps = wps / wpc;
                run<1>(slab, n_rows, dim, wpc, cps, out);
                run<2>(slab, n_rows, dim, wpc, cps, out);
                run<4>(slab, n_rows, dim, wpc, cps, out);
            }
        }
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
