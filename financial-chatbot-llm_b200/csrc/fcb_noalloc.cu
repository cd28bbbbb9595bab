// fcb_search_noalloc — the filtered top-k search for callers that are not allowed to allocate.
//
// Who needs this: a process that shares a GPU whose memory is already spoken for — e.g. a
// gunicorn worker attached to a corpus that another, resident process owns and has sized to
// fill HBM, so a cudaMalloc in the worker would fail or evict someone — and request threads
// that must never run into the device-wide synchronisation cudaMalloc/cudaFree imply. The normal
// route (fcb_open + fcb_load_shard + fcb_search) allocates an engine, inverse-norm column,
// candidate scratch and staging buffers. This entry point allocates NOTHING, on the host or on
// the device: no cudaMalloc / cudaMallocHost / cudaFree, no engine handle, no std::string.
//
// How: it runs a variant of the fused scan kernel of fcb_search.cu whose scratch (the arrival
// counter and the per-CTA candidate lists the last CTA merges) lives in the file-scope device
// array g_noalloc_scratch below instead of engine-owned memory, and which derives each row's
// inverse norm inside the scan instead of reading a precomputed column. The caller passes every
// input and output buffer.
//
// Limits and rules (all checked; violations return FCB_ERR_INVALID / FCB_ERR_UNSUPPORTED with a
// message in the caller's `err` buffer):
//   - ONE query per call, k <= FCB_NOALLOC_MAX_K (16), dim <= FCB_NOALLOC_MAX_DIM (8192).
//   - the scratch array is FCB_NOALLOC_SCRATCH_BYTES (262,400 bytes: a 256-byte header plus
//     1024 candidate lists of 256 bytes). It is part of the library image, so EVERY GPU that
//     loads libfcb200.so holds that much device memory from load until the process exits,
//     whether or not this entry point is ever called.
//   - all calls on one GPU share that array and it keeps whatever the previous call left in it.
//     The one thing that must start at zero — the arrival counter — is cleared on `stream`
//     before every launch. Two calls whose kernels can overlap on the same GPU (different
//     streams, or different host threads) RACE on the array and give wrong results: serialise
//     them (one stream, or an event between them).
//   - the first call on each device looks up the array's device address
//     (cudaGetSymbolAddress, with that device current) and reads the SM count; make that
//     first call before starting any stream capture. Later calls only clear the counter and launch.
#include <stdarg.h>
#include <stdio.h>

#include <mutex>

#include "fcb_internal.cuh"
#include "fcb_topk.cuh"

#define FCB_NOALLOC_MAX_K 16
#define FCB_NOALLOC_MAX_DIM 8192
#define FCB_NOALLOC_MAX_LISTS 1024
#define FCB_NOALLOC_HEADER_BYTES 256
#define FCB_NOALLOC_SCRATCH_BYTES (FCB_NOALLOC_HEADER_BYTES + FCB_NOALLOC_MAX_LISTS * FCB_FUSED_CAND * 8)
#define FCB_NOALLOC_MAX_DEVICES 64

// In host code this name is only a stand-in for the symbol, NOT a device address: the address a
// kernel can use comes from cudaGetSymbolAddress (device_setup below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[FCB_NOALLOC_SCRATCH_BYTES];

namespace {

struct NoallocParams {
    const void *vectors;
    const uint64_t *ids;
    const uint32_t *id_rank;
    const int32_t *dates;
    const uint8_t *cats;
    const double *query;
    fcb_filter filter;
    int64_t n_rows, row_base;
    int dim, k;
    unsigned char *scratch;                // device address of g_noalloc_scratch
    uint64_t *out_ids;
    double *out_scores;
    int32_t *out_count;
};

// Exact answer by ONE CTA, used only when the certificate below fails (more rows within float32 rounding of the k-th
// score than the 32-candidate margin: duplicated rows, a degenerate query). The engine route hands such a query to a
// grid-wide float64 kernel with an 8-byte-per-row key buffer; this entry point may not allocate one, so the last CTA
// walks the segment again on its own: canonical float64 score of every passing row, each warp keeping its 32 best by
// (score desc, id asc) in registers, the 8 lists merged through shared memory. Slow for a large segment (one SM), but
// allocation-free, and only ever run for the queries the fast pass cannot prove.
template <typename T>
__device__ __noinline__ void na_exact_rescan(const NoallocParams &p, int64_t lo, int64_t hi, bool has_date, bool has_cat) {
    __shared__ double x_s[FCB_SCAN_WARPS][32];
    __shared__ uint32_t x_t[FCB_SCAN_WARPS][32], x_r[FCB_SCAN_WARPS][32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const T *__restrict__ vec = reinterpret_cast<const T *>(p.vectors);
    const fcb_filter f = p.filter;
    double bs = -INFINITY;
    uint32_t bt = 0xffffffffu, br = FCB_NO_ROW;
    auto insert = [&](double sc, uint32_t tie, uint32_t row) {            // warp-uniform arguments
        const bool mine_better = br != FCB_NO_ROW && (bs > sc || (bs == sc && bt < tie));
        const int pos = __popc(__ballot_sync(FCB_FULL, mine_better));     // better entries form a prefix
        const double us = __shfl_up_sync(FCB_FULL, bs, 1);
        const uint32_t ut = __shfl_up_sync(FCB_FULL, bt, 1), ur = __shfl_up_sync(FCB_FULL, br, 1);
        if (lane == pos) { bs = sc; bt = tie; br = row; }
        else if (lane > pos) { bs = us; bt = ut; br = ur; }
    };
    for (int64_t row = lo + warp; row < hi; row += FCB_SCAN_WARPS) {
        bool pass = true;
        if (has_date) { const int64_t d = p.dates[row]; pass = d >= f.date_lo && d <= f.date_hi; }
        if (pass && has_cat) { const unsigned c = p.cats[row]; pass = c < 64 && ((f.category_mask >> c) & 1ull); }
        if (!pass) continue;
        const double sc = fcb_canon_score_warp<T>(vec + row * (int64_t)p.dim, p.query, p.dim, lane);
        insert(sc, fcb_tie_key((uint32_t)row, p.id_rank), (uint32_t)row);
    }
    x_s[warp][lane] = bs; x_t[warp][lane] = bt; x_r[warp][lane] = br;
    __syncthreads();
    if (warp != 0) return;
    for (int w = 1; w < FCB_SCAN_WARPS; ++w)
        for (int i = 0; i < 32; ++i) {
            if (x_r[w][i] == FCB_NO_ROW) break;
            insert(x_s[w][i], x_t[w][i], x_r[w][i]);
        }
    const int nvalid = __popc(__ballot_sync(FCB_FULL, br != FCB_NO_ROW));
    const int count = nvalid < p.k ? nvalid : p.k;
    if (lane < p.k) {
        const bool ok = lane < count;
        p.out_ids[lane] = ok ? (p.ids ? p.ids[br] : (uint64_t)(p.row_base + br)) : UINT64_MAX;
        p.out_scores[lane] = ok ? bs : -INFINITY;
    }
    if (lane == 0) *p.out_count = count;
}

// dot and sum of squares of one 16-byte piece. The sum of squares uses the same per-lane order as inv_norm_kernel
// (fcb_engine.cu): lane l accumulates the 16-byte pieces l, l+32, ... with fmaf, then the same xor butterfly — which is
// the only reason the in-scan rsqrtf(ss) equals the precomputed inverse norm bit for bit. The first-pass float32 scores
// only pick the 32 candidates; what is returned is always the canonical float64 re-score.
template <typename T> struct NaPiece;
template <> struct NaPiece<float> {
    static constexpr int VEC = 4;
    __device__ static __forceinline__ void acc(uint4 v, const float *q, float &dot, float &ss) {
        const float a = __uint_as_float(v.x), b = __uint_as_float(v.y), c = __uint_as_float(v.z),
                    d = __uint_as_float(v.w);
        dot = fmaf(a, q[0], dot); dot = fmaf(b, q[1], dot); dot = fmaf(c, q[2], dot); dot = fmaf(d, q[3], dot);
        ss = fmaf(a, a, ss); ss = fmaf(b, b, ss); ss = fmaf(c, c, ss); ss = fmaf(d, d, ss);
    }
};
template <> struct NaPiece<__nv_bfloat16> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ void acc(uint4 v, const float *q, float &dot, float &ss) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float lo = fcb_bf16lo(w[i]), hi = fcb_bf16hi(w[i]);
            dot = fmaf(lo, q[2 * i], dot); dot = fmaf(hi, q[2 * i + 1], dot);
            ss = fmaf(lo, lo, ss); ss = fmaf(hi, hi, ss);
        }
    }
};

template <typename T>
__global__ void __launch_bounds__(FCB_SCAN_THREADS, 2)
scan_topk_noalloc_kernel(const __grid_constant__ NoallocParams p) {
    constexpr int R = 4;
    constexpr int VEC = NaPiece<T>::VEC;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *q_s = reinterpret_cast<float *>(smem_raw);
    __shared__ FcbCand m_s[FCB_SCAN_WARPS][FCB_FUSED_CAND];
    __shared__ double f_score[FCB_FUSED_CAND];
    __shared__ uint64_t f_id[FCB_FUSED_CAND];
    __shared__ int s_last;

    unsigned *counter = reinterpret_cast<unsigned *>(p.scratch);
    FcbCand *lists = reinterpret_cast<FcbCand *>(p.scratch + FCB_NOALLOC_HEADER_BYTES);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const fcb_filter f = p.filter;
    const double *__restrict__ qd = p.query;
    const uint32_t *__restrict__ id_rank = p.id_rank;

    for (int i = tid; i < p.dim; i += FCB_SCAN_THREADS) q_s[i] = (float)qd[i];
    __syncthreads();

    int64_t lo = f.row_begin - p.row_base, hi = f.row_end - p.row_base;
    if (lo < 0) lo = 0;
    if (hi > p.n_rows) hi = p.n_rows;
    const int64_t nrows = hi > lo ? hi - lo : 0;
    const int64_t ngroups = (nrows + R - 1) / R;
    const bool has_date = f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX;
    const bool has_cat = f.category_mask != FCB_ALL_CATEGORIES;
    const int nv = p.dim / VEC;
    const T *__restrict__ vec = reinterpret_cast<const T *>(p.vectors);

    float my_s = -INFINITY, thr = -INFINITY;
    uint32_t my_r = FCB_NO_ROW;
    bool ood = false;

    const int64_t stride = (int64_t)gridDim.x * FCB_SCAN_WARPS;
    for (int64_t g = (int64_t)blockIdx.x * FCB_SCAN_WARPS + warp; g < ngroups; g += stride) {
        const int64_t r0 = lo + g * R;
        bool pass = false;
        if (lane < R && r0 + lane < hi) {
            const int64_t row = r0 + lane;
            pass = true;
            if (has_date) { const int64_t d = p.dates[row]; pass = d >= f.date_lo && d <= f.date_hi; }
            if (pass && has_cat) { const unsigned c = p.cats[row]; pass = c < 64 && ((f.category_mask >> c) & 1ull); }
        }
        const unsigned pm = __ballot_sync(FCB_FULL, pass);
        if (!pm) continue;
        const uint4 *__restrict__ base = reinterpret_cast<const uint4 *>(vec + r0 * (int64_t)p.dim);
        float dot[R], ss[R];
#pragma unroll
        for (int r = 0; r < R; ++r) { dot[r] = 0.f; ss[r] = 0.f; }
#pragma unroll 4
        for (int c = lane; c < nv; c += 32) {
            float qv[VEC];
#pragma unroll
            for (int i = 0; i < VEC; i += 4)
                *reinterpret_cast<float4 *>(&qv[i]) = *reinterpret_cast<const float4 *>(&q_s[c * VEC + i]);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint4 v = fcb_ldg_stream(base + (int64_t)r * nv + c, (pm >> r) & 1u);
                NaPiece<T>::acc(v, qv, dot[r], ss[r]);
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) { dot[r] = fcb_warp_sum(dot[r]); ss[r] = fcb_warp_sum(ss[r]); }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float inv = ss[r] > 0.f ? rsqrtf(ss[r]) : 0.f;
            // a passing row whose sum of squares left float32's safe range (or is zero): the first pass has no error
            // bound for it, so the last CTA must answer exactly (the engine route checks this once, at fcb_load_shard)
            if (((pm >> r) & 1u) && !(ss[r] >= 7.8886091e-31f && ss[r] <= 1.2676506e30f)) ood = true;
            const float s = dot[r] * inv;
            if (((pm >> r) & 1u) && s >= thr) {
                fcb_warp_insert(my_s, my_r, s, (uint32_t)(r0 + r), id_rank, lane);
                thr = __shfl_sync(FCB_FULL, my_s, 31);
            }
        }
    }

    fcb_cta_merge_tree(my_s, my_r, m_s, id_rank, warp, lane);
    if (warp == 0) lists[(size_t)blockIdx.x * FCB_FUSED_CAND + lane] = FcbCand{my_s, my_r};
    __syncthreads();
    if (ood && lane == 0) atomicOr(counter + 1, 1u);               // header word 1: "a row outside the float32 range passed"
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    my_s = -INFINITY; my_r = FCB_NO_ROW;
    for (unsigned L = warp; L < gridDim.x; L += FCB_SCAN_WARPS) {
        const uint2 raw = __ldcg(reinterpret_cast<const uint2 *>(lists + (size_t)L * FCB_FUSED_CAND + lane));
        fcb_warp_merge_list(my_s, my_r, FcbCand{__uint_as_float(raw.x), raw.y}, id_rank, lane);
    }
    __syncthreads();
    fcb_cta_merge_tree(my_s, my_r, m_s, id_rank, warp, lane);
    for (int c = warp; c < FCB_FUSED_CAND; c += FCB_SCAN_WARPS) {
        const uint32_t row = m_s[0][c].row;
        double sc = -INFINITY;
        uint64_t id = UINT64_MAX;
        if (row != FCB_NO_ROW) {
            sc = fcb_canon_score_warp<T>(vec + (int64_t)row * p.dim, qd, p.dim, lane);
            id = p.ids ? p.ids[row] : (uint64_t)(p.row_base + row);
        }
        if (lane == 0) { f_score[c] = sc; f_id[c] = id; }
    }
    __syncthreads();
    if (warp == 0) {
        double kth;
        int nvalid;
        fcb_rank_and_write(m_s[0], f_score, f_id, p.k, lane, p.out_ids, p.out_scores, p.out_count, kth, nvalid);
        // the same certificate as the engine route (fcb_search.cu, scan_tail): a dropped row's canonical score is at most
        // worst kept first-pass score / |q| + eps
        int flag = 0;
        if (nvalid == FCB_FUSED_CAND) {
            const double nq = __dsqrt_rn(fcb_canon_qq_warp(qd, p.dim, lane));
            const float worst = m_s[0][FCB_FUSED_CAND - 1].score;
            const bool domain = nq > 8.9e-16 && nq < 1.1e15 && isfinite(worst);
            flag = !(domain && kth > (double)worst / nq + fcb_first_pass_eps(p.dim));
        }
        if (__ldcg(counter + 1) != 0) flag = 1;
        if (lane == 0) s_last = flag;                                  // reuse: 1 = not certified
    }
    __syncthreads();
    if (s_last) na_exact_rescan<T>(p, lo, hi, has_date, has_cat);
}

// ---- per-device set-up state: plain static storage, nothing allocated
struct NoallocDevice {
    bool ready;
    unsigned char *scratch;                // what cudaGetSymbolAddress returned on this device
    int num_sms;
};
NoallocDevice g_dev[FCB_NOALLOC_MAX_DEVICES];
std::mutex g_dev_mutex;

int32_t na_fail(char *err, int32_t err_len, int32_t code, const char *fmt, ...) {
    if (err && err_len > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, (size_t)err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

// Runs once per device, with that device current (the caller holds a FcbDeviceGuard).
int32_t device_setup(int device, char *err, int32_t err_len) {
    std::lock_guard<std::mutex> lock(g_dev_mutex);
    NoallocDevice &d = g_dev[device];
    if (d.ready) return FCB_OK;
    int major = 0, sms = 0;
    cudaError_t st = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (st == cudaSuccess) st = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (st != cudaSuccess)
        return na_fail(err, err_len, FCB_ERR_CUDA, "fcb_search_noalloc: cudaDeviceGetAttribute: %s", cudaGetErrorString(st));
    if (major < 10)
        return na_fail(err, err_len, FCB_ERR_UNSUPPORTED, "fcb_search_noalloc: device %d is not sm_100-class", device);
    unsigned char *ptr = nullptr;
    st = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
    if (st != cudaSuccess || ptr == nullptr)
        return na_fail(err, err_len, FCB_ERR_CUDA, "fcb_search_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s",
                       device, cudaGetErrorString(st));
    d.scratch = ptr;
    d.num_sms = sms;
    d.ready = true;
    return FCB_OK;
}

}  // namespace

extern "C" int32_t fcb_search_noalloc(int32_t device, const fcb_shard *s, const double *query,
                                      const fcb_filter *filter, int32_t k, uint64_t *out_ids,
                                      double *out_scores, int32_t *out_count, void *stream, char *err,
                                      int32_t err_len) {
    if (err && err_len > 0) err[0] = '\0';
    if (!s || !query || !filter || !out_ids || !out_scores || !out_count)
        return na_fail(err, err_len, FCB_ERR_INVALID, "fcb_search_noalloc: NULL argument");
    if (device < 0 || device >= FCB_NOALLOC_MAX_DEVICES)
        return na_fail(err, err_len, FCB_ERR_INVALID, "fcb_search_noalloc: device %d outside 0..%d", device,
                       FCB_NOALLOC_MAX_DEVICES - 1);
    if (k < 1 || k > FCB_NOALLOC_MAX_K)
        return na_fail(err, err_len, FCB_ERR_UNSUPPORTED,
                       "fcb_search_noalloc: k=%d is beyond this entry point's maximum of %d (use fcb_search)", k,
                       FCB_NOALLOC_MAX_K);
    if (s->dtype != FCB_DTYPE_F32 && s->dtype != FCB_DTYPE_BF16)
        return na_fail(err, err_len, FCB_ERR_INVALID, "fcb_search_noalloc: dtype %d is not FCB_DTYPE_F32/BF16", s->dtype);
    const int vec = s->dtype == FCB_DTYPE_F32 ? 4 : 8;
    if (s->dim <= 0 || s->dim % vec != 0)
        return na_fail(err, err_len, FCB_ERR_INVALID, "fcb_search_noalloc: dim %d must be a positive multiple of %d", s->dim, vec);
    if (s->dim > FCB_NOALLOC_MAX_DIM)
        return na_fail(err, err_len, FCB_ERR_UNSUPPORTED,
                       "fcb_search_noalloc: dim=%d is beyond this entry point's maximum of %d", s->dim, FCB_NOALLOC_MAX_DIM);
    if (s->n_rows <= 0 || s->n_rows >= (int64_t)FCB_NO_ROW || !s->vectors || ((uintptr_t)s->vectors & 15) ||
        (s->id_rank && !s->ids) || s->row_base < 0)
        return na_fail(err, err_len, FCB_ERR_INVALID, "fcb_search_noalloc: bad shard (n_rows, vectors alignment, id_rank or row_base)");

    FcbDeviceGuard g(device);
    if (!g.ok) return na_fail(err, err_len, FCB_ERR_CUDA, "fcb_search_noalloc: cannot select device %d", device);
    const int32_t rc = device_setup(device, err, err_len);
    if (rc != FCB_OK) return rc;
    const NoallocDevice &d = g_dev[device];

    NoallocParams p{};
    p.vectors = s->vectors; p.ids = s->ids; p.id_rank = s->id_rank; p.dates = s->dates; p.cats = s->categories;
    p.query = query; p.filter = *filter;
    const bool has_date = p.filter.date_lo != FCB_DATE_MIN || p.filter.date_hi != FCB_DATE_MAX;
    const bool has_cat = p.filter.category_mask != FCB_ALL_CATEGORIES;
    if ((has_date && !s->dates) || (has_cat && !s->categories) || p.filter.date_lo > p.filter.date_hi)
        p.filter.row_end = p.filter.row_begin;           // predicate on an absent column: nothing matches
    p.n_rows = s->n_rows; p.row_base = s->row_base; p.dim = s->dim; p.k = k;
    p.scratch = d.scratch;
    p.out_ids = out_ids; p.out_scores = out_scores; p.out_count = out_count;

    int64_t lo = p.filter.row_begin - s->row_base, hi = p.filter.row_end - s->row_base;
    if (lo < 0) lo = 0;
    if (hi > s->n_rows) hi = s->n_rows;
    const int64_t rows = hi > lo ? hi - lo : 0;
    const int64_t groups = (rows + 3) / 4;
    int64_t ctas = (groups + 2 * FCB_SCAN_WARPS - 1) / (2 * FCB_SCAN_WARPS);
    if (ctas > 2 * (int64_t)d.num_sms) ctas = 2 * (int64_t)d.num_sms;
    if (ctas > FCB_NOALLOC_MAX_LISTS) ctas = FCB_NOALLOC_MAX_LISTS;
    if (ctas < 1) ctas = 1;

    cudaStream_t st = (cudaStream_t)stream;
    // the array still holds the previous call's lists and counter: zero the header (the counter)
    cudaError_t ce = cudaMemsetAsync(d.scratch, 0, FCB_NOALLOC_HEADER_BYTES, st);
    if (ce != cudaSuccess)
        return na_fail(err, err_len, FCB_ERR_CUDA, "fcb_search_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(ce));
    const size_t smem = (size_t)s->dim * sizeof(float);  // <= 32 KB: below the 48 KB default limit
    if (s->dtype == FCB_DTYPE_F32)
        scan_topk_noalloc_kernel<float><<<(unsigned)ctas, FCB_SCAN_THREADS, smem, st>>>(p);
    else
        scan_topk_noalloc_kernel<__nv_bfloat16><<<(unsigned)ctas, FCB_SCAN_THREADS, smem, st>>>(p);
    ce = cudaGetLastError();
    if (ce != cudaSuccess)
        return na_fail(err, err_len, FCB_ERR_CUDA, "fcb_search_noalloc: launch: %s", cudaGetErrorString(ce));
    return FCB_OK;
}
