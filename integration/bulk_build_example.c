/*
 * bulk_build_example.c — a bulk index build from C: the reference's `for each row: hnsw_insert(...)`
 * (/root/reference/src/hnsw_algo.c:520-666) with the layer-0 candidate searches of every insert group on the GPU.
 * libmuninn_host runs the loop, libmuninn_cuda is its searcher. Build and run on a B200:
 *
 *   gcc -O2 -std=c11 -D_POSIX_C_SOURCE=200809L -Iinclude integration/bulk_build_example.c -Lsqlite_muninn_b200/csrc \
 *       -lmuninn_host -lmuninn_cuda -Wl,-rpath,$PWD/sqlite_muninn_b200/csrc -lm -o /tmp/bulk_build_example
 *   /tmp/bulk_build_example 200000 128
 *
 * tests/test_host.py compiles it (headers and call signatures stay in step); it needs a device to run.
 */
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include "muninn_build.h"
#include "muninn_cuda.h"

int main(int argc, char **argv) {
    const int64_t n = argc > 1 ? atoll(argv[1]) : 100000;
    const int dim = argc > 2 ? atoi(argv[2]) : 128, M = 16, efc = 200;
    float *x = (float *)malloc((size_t)n * (size_t)dim * sizeof(float));
    if (!x) return 1;
    unsigned s = 12345; /* rows on a 16-dimensional latent + noise: a graph the reference's selection rule handles well */
    float lat[16];
    for (int64_t i = 0; i < n; i++) {
        for (int j = 0; j < 16; j++) { s = s * 1103515245u + 12345u; lat[j] = (float)(s >> 8) / 8388608.0f - 1.0f; }
        for (int d = 0; d < dim; d++) {
            s = s * 1103515245u + 12345u;
            x[i * dim + d] = lat[d % 16] * (float)(1 + d / 16) + 0.05f * ((float)(s >> 8) / 8388608.0f - 1.0f);
        }
    }
    MuninnCudaIndex *idx = NULL;
    if (muninn_cuda_index_create(0, dim, MUNINN_METRIC_L2, M, &idx) != MUNINN_CUDA_OK) {
        fprintf(stderr, "no device: %s\n", muninn_cuda_last_error(NULL)); /* there is no host fallback */
        return 2;
    }
    MuninnBuildDevice dev;
    muninn_cuda_build_device(idx, &dev);
    void *b = muninn_build_create(dim, MUNINN_METRIC_L2, M, efc, n, x);
    MuninnBuildRunStats st;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    int rc = muninn_build_run(b, NULL, 1024, 4096, 16, &dev, &st);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (rc != 0) { fprintf(stderr, "build failed (%d): %s\n", rc, muninn_cuda_last_error(idx)); return 3; }
    double secs = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    printf("{\"n\": %lld, \"dim\": %d, \"seconds\": %.3f, \"vectors_per_s\": %.0f, \"groups\": %lld, \"device_queries\": %lld, "
           "\"host_searches\": %lld, \"patched_records\": %lld, \"warm_s\": %.3f, \"search_s\": %.3f, \"link_s\": %.3f, \"patch_s\": %.3f}\n",
           (long long)n, dim, secs, (double)n / secs, (long long)st.groups, (long long)st.device_queries, (long long)st.host_searches,
           (long long)st.patched_records, st.warm_s, st.search_s, st.link_s, st.patch_s);
    /* the handle now answers searches over the finished graph */
    int64_t ids[10]; float dist[10]; int32_t cnt = 0;
    muninn_cuda_search_batch(idx, x, 1, 10, 64, ids, dist, &cnt, NULL);
    printf("nearest of row 0: id %lld at distance %g (%d results)\n", (long long)ids[0], dist[0], cnt);
    muninn_build_destroy(b);
    muninn_cuda_index_destroy(idx);
    free(x);
    return 0;
}
