// Synthetic corpus generator on the device: the same counter-based integer recipe as
// financial-chatbot-llm_b200/synth.py (bit-identical; no libm, no reductions).
#include "fcb_internal.cuh"

#define S_CLUSTER 1ull
#define S_CENTRE 2ull
#define S_NOISE 3ull
#define S_DATE 4ull
#define S_CAT 5ull
#define S_CENTS 6ull
#define SYNTH_DATE_BASE 1640995200ll

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27; x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return x;
}
__device__ __forceinline__ uint64_t hash_u64(uint64_t seed, uint64_t stream, uint64_t idx) {
    return mix64((seed ^ (stream * 0xD6E8FEB86659FD93ull)) + idx * 0x9E3779B97F4A7C15ull);
}
__device__ __forceinline__ int noise_from_bits(uint64_t u) {
    return (int)((u & 0xFFFF) + ((u >> 16) & 0xFFFF) + ((u >> 32) & 0xFFFF) + (u >> 48)) - 131070;
}

// one thread per 4 consecutive elements of a row (dim % 4 == 0)
template <typename T>
__global__ void __launch_bounds__(256) synth_vectors_kernel(T *__restrict__ out, int64_t row_lo, int64_t n_rows,
                                                            int dim, uint64_t seed, uint32_t n_clusters) {
    const int64_t quads_per_row = dim / 4;
    const int64_t total = n_rows * quads_per_row;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t lr = t / quads_per_row;
        const int j0 = (int)(t - lr * quads_per_row) * 4;
        const uint64_t row = (uint64_t)(row_lo + lr);
        const uint64_t cl = hash_u64(seed, S_CLUSTER, row) % n_clusters;
        float v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint64_t j = (uint64_t)(j0 + i);
            const int centre = (int)(hash_u64(seed, S_CENTRE, cl * (uint64_t)dim + j) >> 48) - 32768;
            const int noise = noise_from_bits(hash_u64(seed, S_NOISE, row * (uint64_t)dim + j));
            v[i] = __int2float_rn(4 * centre + noise) * 3.814697265625e-06f;   // * 2^-18, exact
        }
        T *dst = out + lr * (int64_t)dim + j0;
        if (sizeof(T) == 4) {
            *reinterpret_cast<float4 *>(dst) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
            __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
            uint2 w;
            w.x = *reinterpret_cast<uint32_t *>(&a);
            w.y = *reinterpret_cast<uint32_t *>(&b);
            *reinterpret_cast<uint2 *>(dst) = w;
        }
    }
}

__global__ void __launch_bounds__(256) synth_columns_kernel(int32_t *__restrict__ dates, uint8_t *__restrict__ cats,
                                                            int32_t *__restrict__ cents, int64_t row_lo, int64_t n_rows,
                                                            uint64_t seed, uint32_t n_categories, uint64_t span) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n_rows; t += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t row = (uint64_t)(row_lo + t);
        if (dates) dates[t] = (int32_t)(SYNTH_DATE_BASE + (int64_t)(hash_u64(seed, S_DATE, row) % span));
        if (cats) cats[t] = (uint8_t)(hash_u64(seed, S_CAT, row) % n_categories);
        if (cents) {
            const uint64_t u = hash_u64(seed, S_CENTS, row);
            const uint64_t e = (u & 0xFF) % 18;
            const uint64_t mant = 1024 + ((u >> 8) & 1023);
            const int64_t mag = (int64_t)((mant << e) >> 10);
            cents[t] = (int32_t)((((u >> 40) & 7) == 0) ? -mag : mag);
        }
    }
}

// dates non-decreasing within every user: row r of segment [a, b) gets DATE_BASE + floor((r - a) * span / (b - a))
__global__ void __launch_bounds__(256) synth_dates_sorted_kernel(int32_t *__restrict__ dates, const int64_t *__restrict__ seg_offsets,
                                                                 int64_t n_segs, int64_t row_lo, int64_t n_rows, uint64_t span) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n_rows; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t g = row_lo + t;
        int64_t lo = 0, hi = n_segs - 1;                           // segment with seg_offsets[u] <= g < seg_offsets[u + 1]
        while (lo < hi) { const int64_t mid = (lo + hi + 1) >> 1; if (seg_offsets[mid] <= g) lo = mid; else hi = mid - 1; }
        const uint64_t a = (uint64_t)seg_offsets[lo], b = (uint64_t)seg_offsets[lo + 1];
        dates[t] = (int32_t)(SYNTH_DATE_BASE + (int64_t)(((uint64_t)g - a) * span / (b - a)));
    }
}

static int32_t synth_fail(const char *what, cudaError_t st) {
    return fcb_fail(nullptr, FCB_ERR_CUDA, "%s: %s", what, cudaGetErrorString(st));
}

extern "C" int32_t fcb_synth_vectors(int32_t device, void *out, int32_t dtype, int64_t row_lo, int64_t n_rows,
                                     int32_t dim, uint64_t seed, int32_t n_clusters, void *stream) {
    if (!out || n_rows < 0 || dim <= 0 || dim % 4 || n_clusters <= 0 || (dtype != FCB_DTYPE_F32 && dtype != FCB_DTYPE_BF16))
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_synth_vectors: bad argument");
    if (n_rows == 0) return FCB_OK;
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_synth_vectors: cannot select device %d", device);
    const int64_t total = n_rows * (dim / 4);
    const unsigned blocks = (unsigned)((total + 255) / 256 < 148 * 32 ? (total + 255) / 256 : 148 * 32);
    if (dtype == FCB_DTYPE_F32)
        synth_vectors_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>((float *)out, row_lo, n_rows, dim, seed, (uint32_t)n_clusters);
    else
        synth_vectors_kernel<__nv_bfloat16><<<blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16 *)out, row_lo, n_rows, dim, seed, (uint32_t)n_clusters);
    cudaError_t st = cudaGetLastError();
    return st == cudaSuccess ? FCB_OK : synth_fail("fcb_synth_vectors", st);
}

extern "C" int32_t fcb_synth_columns(int32_t device, int32_t *dates, uint8_t *categories, int32_t *cents, int64_t row_lo,
                                     int64_t n_rows, uint64_t seed, int32_t n_categories, int64_t date_span_seconds,
                                     void *stream) {
    if (n_rows < 0 || n_categories <= 0 || n_categories > 64 || date_span_seconds <= 0)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_synth_columns: bad argument");
    if (n_rows == 0) return FCB_OK;
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_synth_columns: cannot select device %d", device);
    const unsigned blocks = (unsigned)((n_rows + 255) / 256 < 148 * 16 ? (n_rows + 255) / 256 : 148 * 16);
    synth_columns_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dates, categories, cents, row_lo, n_rows, seed,
                                                                  (uint32_t)n_categories, (uint64_t)date_span_seconds);
    cudaError_t st = cudaGetLastError();
    return st == cudaSuccess ? FCB_OK : synth_fail("fcb_synth_columns", st);
}

extern "C" int32_t fcb_synth_dates_sorted(int32_t device, int32_t *dates, const int64_t *seg_offsets, int64_t n_segs,
                                          int64_t row_lo, int64_t n_rows, int64_t date_span_seconds, void *stream) {
    if (!dates || !seg_offsets || n_segs <= 0 || n_rows < 0 || date_span_seconds <= 0)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_synth_dates_sorted: bad argument");
    if (n_rows == 0) return FCB_OK;
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_synth_dates_sorted: cannot select device %d", device);
    const unsigned blocks = (unsigned)((n_rows + 255) / 256 < 148 * 16 ? (n_rows + 255) / 256 : 148 * 16);
    synth_dates_sorted_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dates, seg_offsets, n_segs, row_lo, n_rows,
                                                                       (uint64_t)date_span_seconds);
    cudaError_t st = cudaGetLastError();
    return st == cudaSuccess ? FCB_OK : synth_fail("fcb_synth_dates_sorted", st);
}
