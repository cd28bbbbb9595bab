/* Plain-C caller of libfcb200.so: what a non-Python host (the cgo / JNI / N-API side of a port of the reference's
 * tools/qdrant_tool.py) binds. Nothing here is CUDA-specific except the two allocation/copy calls a host needs to place
 * its corpus in device memory; every fcb_* signature uses plain pointers and sizes.
 *
 *   gcc -std=c99 -I include examples/ffi_demo.c -L financial-chatbot-llm_b200/csrc -lfcb200 \
 *       -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/financial-chatbot-llm_b200/csrc -o ffi_demo
 *
 * Builds a 4096-row, 64-dimensional float32 corpus of two users on the GPU, asks for user 1's five nearest transactions
 * newer than a date, and prints ids and float64 cosine scores. Needs a B200 to RUN; the CPU test suite only compiles and
 * links it (tests/test_abi.py). */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "fcb200.h"

/* the two CUDA runtime calls a host needs, declared by hand so that no CUDA header is required */
extern int cudaMalloc(void **ptr, size_t size);
extern int cudaMemcpy(void *dst, const void *src, size_t count, int kind);
extern int cudaFree(void *ptr);
enum { H2D = 1, D2H = 2 };

#define ROWS 4096
#define DIM 64
#define K 5

static void *to_device(const void *host, size_t bytes) {
    void *dev = NULL;
    if (cudaMalloc(&dev, bytes) != 0 || cudaMemcpy(dev, host, bytes, H2D) != 0) { fprintf(stderr, "CUDA allocation failed\n"); exit(2); }
    return dev;
}

int main(void) {
    static float vectors[ROWS][DIM];
    static int32_t dates[ROWS];
    const int64_t seg_offsets[3] = {0, 1500, ROWS};              /* user 0 owns rows [0,1500), user 1 the rest */
    uint32_t x = 12345u;
    for (int r = 0; r < ROWS; ++r) {
        for (int d = 0; d < DIM; ++d) { x = x * 1664525u + 1013904223u; vectors[r][d] = (float)(int32_t)(x >> 8) / 8388608.0f - 1.0f; }
        dates[r] = 1700000000 + 3600 * r;
    }
    double query[DIM];
    for (int d = 0; d < DIM; ++d) query[d] = vectors[2000][d] + 0.01 * d;   /* close to row 2000 */

    fcb_config cfg = {0, 1, FCB_MAX_K, 0};                        /* device 0, max_batch 1, max_k 10000 = the tool's default limit */
    fcb_engine *eng = NULL;
    if (fcb_open(&cfg, &eng) != 0) { fprintf(stderr, "fcb_open failed\n"); return 1; }

    fcb_shard shard = {0};
    shard.vectors = to_device(vectors, sizeof vectors);
    shard.dates = (const int32_t *)to_device(dates, sizeof dates);
    shard.seg_offsets = (const int64_t *)to_device(seg_offsets, sizeof seg_offsets);
    shard.n_rows = ROWS; shard.n_segs = 2; shard.row_base = 0; shard.dim = DIM; shard.dtype = 0;   /* float32 */
    if (fcb_load_shard(eng, &shard) != 0) { fprintf(stderr, "fcb_load_shard: %s\n", fcb_last_error(eng)); return 1; }

    /* Filter(must=[user_id == user 1, date >= dates[1800]]) -> row range + inclusive date bounds */
    fcb_filter flt = {seg_offsets[1], seg_offsets[2], dates[1800], INT64_MAX, UINT64_MAX};
    uint64_t ids[K];
    double scores[K];
    int32_t count = 0;
    if (fcb_search_host(eng, query, &flt, 1, K, ids, scores, &count) != 0) { fprintf(stderr, "fcb_search_host: %s\n", fcb_last_error(eng)); return 1; }
    for (int i = 0; i < count; ++i) printf("%d: id %llu score %.17g\n", i, (unsigned long long)ids[i], scores[i]);
    fcb_close(eng);
    cudaFree((void *)shard.vectors); cudaFree((void *)shard.dates); cudaFree((void *)shard.seg_offsets);
    return count == K && ids[0] == 2000 ? 0 : 3;
}
