// Engine lifetime, error reporting, shard attach + inverse row norms.
#include <stdarg.h>
#include <stdio.h>

#include "fcb_internal.cuh"

static thread_local std::string g_open_error;

int32_t fcb_fail(fcb_engine *e, int32_t code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (e) e->err = buf; else g_open_error = buf;
    return code;
}

extern "C" int32_t fcb_abi_version(void) { return FCB_ABI_VERSION; }

extern "C" int32_t fcb_shard_dates_sorted(const fcb_engine *e) { return e && e->loaded ? e->dates_sorted : -1; }

extern "C" const char *fcb_last_error(const fcb_engine *e) {
    if (!e) return g_open_error.c_str();
    const int code = fcb_dense_error_code(e);                  // a trap in the batched tensor-core path: say so (once)
    if (code != 0 && e->err.find("pipeline protocol") == std::string::npos) {
        fcb_engine *m = const_cast<fcb_engine *>(e);
        char buf[160];
        snprintf(buf, sizeof buf, " [batched tensor-core kernel trapped: pipeline protocol violation, wait code %d]", code);
        m->err += buf;
    }
    return e->err.c_str();
}

static void free_host_staging(fcb_engine *e) {
    cudaFreeHost(e->h_q); cudaFree(e->d_q);
    cudaFreeHost(e->h_ids); cudaFree(e->d_ids);          // one block each: ids | scores | counts
    e->h_q = e->d_q = nullptr; e->h_ids = e->d_ids = nullptr;
    e->h_scores = e->d_scores = nullptr; e->h_counts = e->d_counts = nullptr;
}

extern "C" int32_t fcb_close(fcb_engine *e) {
    if (!e) return FCB_OK;
    FcbDeviceGuard g(e->device);
    if (g.ok) {
        cudaDeviceSynchronize();
        free_host_staging(e);
        cudaFree(e->inv_norm);
        cudaFree(e->row_of_rank);
        cudaFree(e->group_keys);
        fcb_exact_free(e);
        fcb_general_free(e);
        fcb_dense_free(e);
        cudaFree(e->d_filters);
        cudaFree(e->d_cand);
        cudaFree(e->d_counters);
        cudaFreeHost(e->h_filters);
        cudaFree(e->d_route);
        cudaFreeHost(e->h_route);
        for (auto &ev : e->filter_ev) if (ev) cudaEventDestroy(ev);
        for (auto &ev : e->route_ev) if (ev) cudaEventDestroy(ev);
        if (e->ev0) cudaEventDestroy(e->ev0);
        if (e->ev1) cudaEventDestroy(e->ev1);
        if (e->stream) cudaStreamDestroy(e->stream);
    }
    delete e;
    return FCB_OK;
}

static int32_t open_impl(fcb_engine *e, const fcb_config *cfg) {
    e->device = cfg->device;
    e->max_batch = cfg->max_batch;
    e->max_k = cfg->max_k;
    FCB_CUDA(e, cudaSetDevice(e->device));
    cudaDeviceProp prop;
    FCB_CUDA(e, cudaGetDeviceProperties(&prop, e->device));
    if (prop.major < 10)
        return fcb_fail(e, FCB_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only",
                        e->device, prop.major, prop.minor);
    e->num_sms = prop.multiProcessorCount;
    FCB_CUDA(e, cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    FCB_CUDA(e, cudaEventCreate(&e->ev0));
    FCB_CUDA(e, cudaEventCreate(&e->ev1));
    for (auto &ev : e->filter_ev) FCB_CUDA(e, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    for (auto &ev : e->route_ev) FCB_CUDA(e, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    const size_t nr = (size_t)FCB_FILTER_SLOTS * (2 * (size_t)e->max_batch + 1);
    FCB_CUDA(e, cudaMalloc(&e->d_route, nr * sizeof(int32_t)));
    FCB_CUDA(e, cudaMallocHost(&e->h_route, nr * sizeof(int32_t)));

    const size_t nf = (size_t)FCB_FILTER_SLOTS * e->max_batch;
    FCB_CUDA(e, cudaMalloc(&e->d_filters, nf * sizeof(fcb_filter)));
    FCB_CUDA(e, cudaMallocHost(&e->h_filters, nf * sizeof(fcb_filter)));
    // fused path: one 256-byte list per CTA. Every query of a batch may get a full persistent wave
    // (2 CTAs per SM) when its segment is large — e.g. one resident whale per shard in a sharded batch —
    // so size for max_batch full waves, capped at 64 Ki lists (16 MB).
    e->cand_lists = (size_t)e->max_batch * 2 * e->num_sms;
    if (e->cand_lists > 65536) e->cand_lists = 65536;
    if (e->cand_lists < (size_t)8 * e->num_sms) e->cand_lists = (size_t)8 * e->num_sms;
    if (e->cand_lists < (size_t)e->max_batch) e->cand_lists = e->max_batch;
    FCB_CUDA(e, cudaMalloc(&e->d_cand, e->cand_lists * FCB_FUSED_CAND * sizeof(FcbCand)));
    FCB_CUDA(e, cudaMalloc(&e->d_counters, (size_t)e->max_batch * sizeof(unsigned)));
    FCB_CUDA(e, cudaMemset(e->d_counters, 0, (size_t)e->max_batch * sizeof(unsigned)));
    FCB_CUDA(e, cudaDeviceSynchronize());
    return FCB_OK;
}

extern "C" int32_t fcb_open(const fcb_config *cfg, fcb_engine **out) {
    if (!cfg || !out) return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_open: NULL argument");
    *out = nullptr;
    if (cfg->max_batch < 1 || cfg->max_batch > 65535)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_open: max_batch %d outside 1..65535", cfg->max_batch);
    if (cfg->max_k < 1 || cfg->max_k > FCB_MAX_K)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_open: max_k %d outside 1..%d", cfg->max_k, FCB_MAX_K);
    if (cfg->reserved != 0) return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_open: reserved must be 0");
    int ndev = 0;
    cudaError_t st = cudaGetDeviceCount(&ndev);
    if (st != cudaSuccess || ndev == 0)
        return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_open: no CUDA device (%s); this library has no CPU path",
                        st == cudaSuccess ? "device count 0" : cudaGetErrorString(st));
    if (cfg->device < 0 || cfg->device >= ndev)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_open: device %d outside 0..%d", cfg->device, ndev - 1);
    int prev = 0;
    cudaGetDevice(&prev);
    fcb_engine *e = new fcb_engine();
    int32_t rc = open_impl(e, cfg);
    if (rc != FCB_OK) {
        g_open_error = e->err;
        fcb_close(e);
        cudaSetDevice(prev);
        return rc;
    }
    cudaSetDevice(prev);
    *out = e;
    return FCB_OK;
}

// ---- inverse row norms ------------------------------------------------------------------
// One warp per row, 128-bit loads, float32 accumulation: the first-pass ranking only needs
// ~1e-6 relative accuracy (the float64 rescore recomputes norms canonically).
// Also counts the rows whose squared norm is outside [2^-100, 2^100] or not finite (zero rows are fine: they score 0 in
// both passes): the float32 first pass has no error bound for those, so a shard holding any sends every query to the
// exact float64 path (fcb_exact.cu).
template <typename T>
__global__ void __launch_bounds__(256) inv_norm_kernel(const T *__restrict__ vectors, int64_t n_rows,
                                                       int dim, float *__restrict__ inv_norm,
                                                       unsigned long long *__restrict__ out_of_domain) {
    constexpr int VEC = 16 / sizeof(T);
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int nv = dim / VEC;
    for (int64_t r = warp0; r < n_rows; r += nwarps) {
        const uint4 *p = reinterpret_cast<const uint4 *>(vectors + r * (int64_t)dim);
        float ss = 0.f;
        unsigned nz = 0;                                       // any non-zero element (sign bits masked)
        for (int c = lane; c < nv; c += 32) {
            const uint4 v = fcb_ldg_stream(p + c, true);
            nz |= (v.x | v.y | v.z | v.w) & (sizeof(T) == 4 ? 0x7fffffffu : 0x7fff7fffu);
            if (sizeof(T) == 4) {
                const float a = __uint_as_float(v.x), b = __uint_as_float(v.y),
                            cc = __uint_as_float(v.z), d = __uint_as_float(v.w);
                ss = fmaf(a, a, ss); ss = fmaf(b, b, ss); ss = fmaf(cc, cc, ss); ss = fmaf(d, d, ss);
            } else {
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float lo = fcb_bf16lo(w[i]), hi = fcb_bf16hi(w[i]);
                    ss = fmaf(lo, lo, ss); ss = fmaf(hi, hi, ss);
                }
            }
        }
        ss = fcb_warp_sum(ss);
        nz = __reduce_or_sync(FCB_FULL, nz);
        if (lane == 0) {
            inv_norm[r] = ss > 0.f ? rsqrtf(ss) : 0.f;
            // a true zero row is fine; a sum of squares that underflowed to zero is not
            if (!((ss == 0.f && nz == 0) || (ss >= 7.8886091e-31f && ss <= 1.2676506e30f))) atomicAdd(out_of_domain, 1ull);
        }
    }
}

extern "C" int32_t fcb_load_shard(fcb_engine *e, const fcb_shard *s) {
    if (!e) return FCB_ERR_INVALID;
    if (!s) return fcb_fail(e, FCB_ERR_INVALID, "fcb_load_shard: NULL shard");
    if (s->dtype != FCB_DTYPE_F32 && s->dtype != FCB_DTYPE_BF16)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_load_shard: dtype %d is not FCB_DTYPE_F32/BF16", s->dtype);
    const int vec = s->dtype == FCB_DTYPE_F32 ? 4 : 8;
    if (s->dim <= 0 || s->dim > 8192 || s->dim % vec != 0)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_load_shard: dim %d must be a multiple of %d in 1..8192", s->dim, vec);
    if (s->n_rows <= 0 || s->n_rows >= (int64_t)FCB_NO_ROW)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_load_shard: n_rows %lld outside 1..2^32-2", (long long)s->n_rows);
    if (!s->vectors || ((uintptr_t)s->vectors & 15))
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_load_shard: vectors must be a 16-byte aligned device pointer");
    if (s->id_rank && !s->ids)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_load_shard: id_rank given without ids");
    if (s->row_base < 0 || s->n_segs < 0)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_load_shard: negative row_base / n_segs");
    FcbDeviceGuard g(e->device);
    if (!g.ok) return fcb_fail(e, FCB_ERR_CUDA, "fcb_load_shard: cannot select device %d", e->device);

    e->loaded = false;
    e->scan_r = e->scan_u = 0;
    cudaFree(e->inv_norm); e->inv_norm = nullptr;
    cudaFree(e->row_of_rank); e->row_of_rank = nullptr;
    cudaFree(e->group_keys); e->group_keys = nullptr;
    fcb_exact_free(e);
    fcb_general_free(e);
    fcb_dense_free(e);
    free_host_staging(e);
    if (cudaMalloc(&e->inv_norm, (size_t)s->n_rows * sizeof(float)) != cudaSuccess) {
        cudaGetLastError();
        return fcb_fail(e, FCB_ERR_NOMEM, "fcb_load_shard: cannot allocate %zu bytes of row norms",
                        (size_t)s->n_rows * sizeof(float));
    }
    const int blocks = e->num_sms * 8;
    unsigned long long *d_ood = nullptr;
    FCB_CUDA(e, cudaMalloc(&d_ood, sizeof(unsigned long long)));
    FCB_CUDA(e, cudaMemsetAsync(d_ood, 0, sizeof(unsigned long long), e->stream));
    if (s->dtype == FCB_DTYPE_F32)
        inv_norm_kernel<float><<<blocks, 256, 0, e->stream>>>((const float *)s->vectors, s->n_rows, s->dim, e->inv_norm, d_ood);
    else
        inv_norm_kernel<__nv_bfloat16><<<blocks, 256, 0, e->stream>>>((const __nv_bfloat16 *)s->vectors, s->n_rows,
                                                                       s->dim, e->inv_norm, d_ood);
    FCB_CUDA(e, cudaGetLastError());
    FCB_CUDA(e, cudaMemcpyAsync(&e->rows_out_of_domain, d_ood, sizeof(unsigned long long), cudaMemcpyDeviceToHost, e->stream));
    FCB_CUDA(e, cudaStreamSynchronize(e->stream));
    cudaFree(d_ood);

    const size_t B = e->max_batch, K = e->max_k;
    FCB_CUDA(e, cudaMallocHost(&e->h_q, B * s->dim * sizeof(double)));
    FCB_CUDA(e, cudaMalloc(&e->d_q, B * s->dim * sizeof(double)));
    // results of the host entry point: ONE block (ids | scores | counts) so that one copy brings them back
    const size_t res_bytes = B * K * 16 + 2 * B * sizeof(int32_t);        // ... | certificate flags
    FCB_CUDA(e, cudaMallocHost(&e->h_ids, res_bytes));
    FCB_CUDA(e, cudaMalloc(&e->d_ids, res_bytes));
    e->h_scores = reinterpret_cast<double *>(e->h_ids + B * K);
    e->d_scores = reinterpret_cast<double *>(e->d_ids + B * K);
    e->h_counts = reinterpret_cast<int32_t *>(e->h_scores + B * K);
    e->d_counts = reinterpret_cast<int32_t *>(e->d_scores + B * K);
    e->shard = *s;
    if (e->max_k > FCB_FUSED_MAX_K) {
        const int32_t rc = fcb_general_alloc(e);
        if (rc != FCB_OK) return rc;
    }
    {
        const int32_t rc = fcb_exact_alloc(e);                  // after the general-k scratch: shares its key buffer
        if (rc != FCB_OK) return rc;
    }
    {
        const int32_t rc = fcb_check_dates_sorted(e);
        if (rc != FCB_OK) return rc;
    }
    {
        const int32_t rc = fcb_build_group_keys(e);
        if (rc != FCB_OK) return rc;
    }
    if (fcb_dense_eligible(e)) {
        const int32_t rc = fcb_dense_alloc(e);
        if (rc != FCB_OK) return rc;
    }
    {
        const int32_t rc = fcb_tune_scan(e);                   // large shards: pick the scan's loop shape by measurement
        if (rc != FCB_OK) return rc;
    }
    FCB_CUDA(e, cudaStreamSynchronize(e->stream));
    e->loaded = true;
    return FCB_OK;
}
