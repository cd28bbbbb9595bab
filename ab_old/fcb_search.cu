// Payload-filtered exact cosine top-k: the kernel behind fcb_search (include/fcb200.h), i.e.
// the replacement for client.query_points(...).points at reference tools/qdrant_tool.py:147-153.
//
// Fused path (k <= FCB_FUSED_MAX_K), ONE kernel per call:
//   grid = (CTAs per query, batch). Each warp streams groups of R consecutive rows of the
//   user's CSR segment with 128-bit no-allocate loads (HBM-bound: 4*dim bytes per passing row),
//   the date / category predicate is evaluated per row BEFORE its vector is fetched (rows that
//   fail cost 5 bytes, not 4*dim), the float32 dot is reduced by warp shuffles, and the warp
//   keeps its 32 best (score, row) candidates one per lane in registers. Warp lists are merged
//   per CTA through shared memory and written to scratch; the last CTA to arrive for a query
//   (atomic counter) merges the per-CTA lists, re-scores the 32 survivors in float64 with the
//   canonical order of operations (bit-identical to oracle/canon.c), ranks them by
//   (score desc, id asc) and writes the k results. No second launch, no host round trip.
#include <string.h>

#include "fcb_internal.cuh"

struct ScanParams {
    const void *vectors;
    const float *inv_norm;
    const uint64_t *ids;
    const uint32_t *id_rank;
    const int32_t *dates;
    const uint8_t *cats;
    const double *queries;
    const fcb_filter *filters;             // device copy, used when !use_inline
    fcb_filter inl[FCB_INLINE_FILTERS];
    int use_inline;
    const int32_t *only_if;                // optional [batch]: run query b only where only_if[b] != 0
    int64_t n_rows, row_base;
    int dim, k;
    FcbCand *cand;
    unsigned *counters;
    uint64_t *out_ids;
    double *out_scores;
    int32_t *out_counts;
};

__device__ __forceinline__ uint32_t tie_key(uint32_t row, const uint32_t *__restrict__ id_rank) {
    return (row == FCB_NO_ROW || id_rank == nullptr) ? row : id_rank[row];
}

// Insert (s, r) into the warp's sorted list (lane i holds the i-th best). Warp-uniform inputs.
__device__ __forceinline__ void warp_insert(float &my_s, uint32_t &my_r, float s, uint32_t r,
                                            const uint32_t *__restrict__ id_rank, int lane) {
    bool mine_better = my_s > s;
    if (my_s == s) mine_better = tie_key(my_r, id_rank) < tie_key(r, id_rank);
    const int pos = __popc(__ballot_sync(FCB_FULL, mine_better));   // better entries form a prefix
    const float up_s = __shfl_up_sync(FCB_FULL, my_s, 1);
    const uint32_t up_r = __shfl_up_sync(FCB_FULL, my_r, 1);
    if (lane == pos) { my_s = s; my_r = r; }
    else if (lane > pos) { my_s = up_s; my_r = up_r; }
}

// Merge a sorted 32-entry list (entry i in lane i of `e`) into the warp's list.
__device__ __forceinline__ void warp_merge_list(float &my_s, uint32_t &my_r, FcbCand e,
                                                const uint32_t *__restrict__ id_rank, int lane) {
    float thr = __shfl_sync(FCB_FULL, my_s, 31);
    for (int i = 0; i < FCB_FUSED_CAND; ++i) {
        const float s = __shfl_sync(FCB_FULL, e.score, i);
        const uint32_t r = __shfl_sync(FCB_FULL, e.row, i);
        if (r == FCB_NO_ROW || !(s >= thr)) break;      // sorted: nothing further can enter
        warp_insert(my_s, my_r, s, r, id_rank, lane);
        thr = __shfl_sync(FCB_FULL, my_s, 31);
    }
}

template <typename T> struct RowDot;
template <> struct RowDot<float> {
    static constexpr int VEC = 4;
    __device__ static __forceinline__ float dot(uint4 v, const float *q, float acc) {
        acc = fmaf(__uint_as_float(v.x), q[0], acc);
        acc = fmaf(__uint_as_float(v.y), q[1], acc);
        acc = fmaf(__uint_as_float(v.z), q[2], acc);
        acc = fmaf(__uint_as_float(v.w), q[3], acc);
        return acc;
    }
};
template <> struct RowDot<__nv_bfloat16> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ float dot(uint4 v, const float *q, float acc) {
        acc = fmaf(fcb_bf16lo(v.x), q[0], acc); acc = fmaf(fcb_bf16hi(v.x), q[1], acc);
        acc = fmaf(fcb_bf16lo(v.y), q[2], acc); acc = fmaf(fcb_bf16hi(v.y), q[3], acc);
        acc = fmaf(fcb_bf16lo(v.z), q[4], acc); acc = fmaf(fcb_bf16hi(v.z), q[5], acc);
        acc = fmaf(fcb_bf16lo(v.w), q[6], acc); acc = fmaf(fcb_bf16hi(v.w), q[7], acc);
        return acc;
    }
};

template <typename T, int R>
__global__ void __launch_bounds__(FCB_SCAN_THREADS, 2)
scan_topk_kernel(const __grid_constant__ ScanParams p) {
    constexpr int VEC = RowDot<T>::VEC;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *q_s = reinterpret_cast<float *>(smem_raw);                 // [dim] float32 query
    __shared__ FcbCand m_s[FCB_SCAN_WARPS][FCB_FUSED_CAND];
    __shared__ double f_score[FCB_FUSED_CAND];
    __shared__ uint64_t f_id[FCB_FUSED_CAND];
    __shared__ int s_last;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y;
    if (p.only_if != nullptr && p.only_if[b] == 0) return;      // block-uniform: re-run of uncertified queries only
    const fcb_filter f = p.use_inline ? p.inl[b] : p.filters[b];
    const double *__restrict__ qd = p.queries + (size_t)b * p.dim;
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

    float my_s = -INFINITY;
    uint32_t my_r = FCB_NO_ROW;
    float thr = -INFINITY;

    // predicate of row (group g, lane) — evaluated one group ahead of the vector loads
    auto load_pred = [&](int64_t g, bool &pass, float &inv) {
        pass = false; inv = 0.f;
        const int64_t row = lo + g * R + lane;
        if (lane < R && row < hi) {
            pass = true;
            if (has_date) { const int64_t d = p.dates[row]; pass = d >= f.date_lo && d <= f.date_hi; }
            if (pass && has_cat) { const unsigned c = p.cats[row]; pass = c < 64 && ((f.category_mask >> c) & 1ull); }
            if (pass) inv = p.inv_norm[row];
        }
    };

    const int64_t stride = (int64_t)gridDim.x * FCB_SCAN_WARPS;
    int64_t g = (int64_t)blockIdx.x * FCB_SCAN_WARPS + warp;
    bool pass = false; float inv = 0.f;
    if (g < ngroups) load_pred(g, pass, inv);
    while (g < ngroups) {
        const int64_t gn = g + stride;
        bool npass = false; float ninv = 0.f;
        if (gn < ngroups) load_pred(gn, npass, ninv);

        const unsigned pm = __ballot_sync(FCB_FULL, pass);
        if (pm) {
            const int64_t r0 = lo + g * R;
            const uint4 *__restrict__ base = reinterpret_cast<const uint4 *>(vec + r0 * (int64_t)p.dim);
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll 4
            for (int c = lane; c < nv; c += 32) {
                float qv[VEC];
#pragma unroll
                for (int i = 0; i < VEC; i += 4)
                    *reinterpret_cast<float4 *>(&qv[i]) = *reinterpret_cast<const float4 *>(&q_s[c * VEC + i]);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const uint4 v = fcb_ldg_stream(base + (int64_t)r * nv + c, (pm >> r) & 1u);
                    acc[r] = RowDot<T>::dot(v, qv, acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = fcb_warp_sum(acc[r]);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float s = acc[r] * __shfl_sync(FCB_FULL, inv, r);
                if (((pm >> r) & 1u) && s >= thr) {
                    warp_insert(my_s, my_r, s, (uint32_t)(r0 + r), id_rank, lane);
                    thr = __shfl_sync(FCB_FULL, my_s, 31);
                }
            }
        }
        pass = npass; inv = ninv; g = gn;
    }

    // ---- CTA merge: 8 warp lists -> warp 0
    m_s[warp][lane] = FcbCand{my_s, my_r};
    __syncthreads();
    if (warp == 0) {
        for (int w = 1; w < FCB_SCAN_WARPS; ++w) warp_merge_list(my_s, my_r, m_s[w][lane], id_rank, lane);
        p.cand[((size_t)b * gridDim.x + blockIdx.x) * FCB_FUSED_CAND + lane] = FcbCand{my_s, my_r};
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        const unsigned prev = atomicAdd(&p.counters[b], 1u);
        s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // ---- last CTA of this query: merge all per-CTA lists
    my_s = -INFINITY; my_r = FCB_NO_ROW;
    const FcbCand *lists = p.cand + (size_t)b * gridDim.x * FCB_FUSED_CAND;
    for (unsigned L = warp; L < gridDim.x; L += FCB_SCAN_WARPS) {
        const uint2 raw = __ldcg(reinterpret_cast<const uint2 *>(lists + (size_t)L * FCB_FUSED_CAND + lane));
        const FcbCand e{__uint_as_float(raw.x), raw.y};
        warp_merge_list(my_s, my_r, e, id_rank, lane);
    }
    __syncthreads();                       // m_s reuse: every warp finished reading it above
    m_s[warp][lane] = FcbCand{my_s, my_r};
    __syncthreads();
    if (warp == 0) {
        for (int w = 1; w < FCB_SCAN_WARPS; ++w) warp_merge_list(my_s, my_r, m_s[w][lane], id_rank, lane);
        m_s[0][lane] = FcbCand{my_s, my_r};
    }
    __syncthreads();

    // ---- canonical float64 re-score of the 32 survivors (warp per candidate)
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
        const bool valid = m_s[0][lane].row != FCB_NO_ROW;
        const int nvalid = __popc(__ballot_sync(FCB_FULL, valid));
        const double ms = f_score[lane];
        const uint64_t mi = f_id[lane];
        int rank = 0;
        for (int j = 0; j < FCB_FUSED_CAND; ++j)
            rank += (m_s[0][j].row != FCB_NO_ROW) && fcb_better64(f_score[j], f_id[j], ms, mi);
        const int count = nvalid < p.k ? nvalid : p.k;
        if (valid && rank < p.k) {
            p.out_ids[(size_t)b * p.k + rank] = mi;
            p.out_scores[(size_t)b * p.k + rank] = ms;
        }
        if (lane >= count && lane < p.k) {
            p.out_ids[(size_t)b * p.k + lane] = UINT64_MAX;
            p.out_scores[(size_t)b * p.k + lane] = -INFINITY;
        }
        if (lane == 0) {
            p.out_counts[b] = count;
            p.counters[b] = 0;             // ready for the next call
        }
    }
}

// ---- host side ------------------------------------------------------------------------------

static int32_t check_search_args(fcb_engine *e, const void *queries, const fcb_filter *filters,
                                 int32_t batch, int32_t k, const void *oi, const void *os, const void *oc) {
    if (!e->loaded) return fcb_fail(e, FCB_ERR_STATE, "fcb_search: no shard loaded (call fcb_load_shard first)");
    if (!queries || !filters || !oi || !os || !oc) return fcb_fail(e, FCB_ERR_INVALID, "fcb_search: NULL argument");
    if (batch < 1 || batch > e->max_batch)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_search: batch %d outside 1..max_batch=%d", batch, e->max_batch);
    if (k < 1 || k > e->max_k)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_search: k %d outside 1..max_k=%d", k, e->max_k);
    return FCB_OK;
}

template <typename T>
static int32_t launch_fused(fcb_engine *e, ScanParams &p, int batch, int64_t max_rows, cudaStream_t st) {
    constexpr int R = 4;
    const int64_t groups = (max_rows + R - 1) / R;
    // >= 2 groups per warp when the segment allows; never more CTAs than the scratch holds
    int64_t ctas = (groups + 2 * FCB_SCAN_WARPS - 1) / (2 * FCB_SCAN_WARPS);
    int64_t budget = (int64_t)e->cand_lists / batch;          // cand_lists >= max_batch >= batch
    const int64_t resident = (int64_t)e->num_sms * 2;
    // a lone query, or the flag-gated re-run of a few uncertified queries: one persistent wave each
    if ((batch == 1 || p.only_if != nullptr) && budget > resident) budget = resident;
    if (ctas > budget) ctas = budget;
    if (ctas < 1) ctas = 1;
    const size_t smem = (size_t)p.dim * sizeof(float);
    auto kern = scan_topk_kernel<T, R>;
    if (smem > 48 * 1024) FCB_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3((unsigned)ctas, (unsigned)batch), FCB_SCAN_THREADS, smem, st>>>(p);
    FCB_CUDA(e, cudaGetLastError());
    e->last_launches += 1;
    return FCB_OK;
}

static int32_t search_device(fcb_engine *e, const double *d_queries, const fcb_filter *h_filters, int32_t batch,
                             int32_t k, uint64_t *d_out_ids, double *d_out_scores, int32_t *d_out_counts,
                             cudaStream_t st) {
    const fcb_shard &s = e->shard;
    e->last_launches = 0;
    if (k > FCB_FUSED_MAX_K) {
        // general-k path: one query at a time, stream-ordered (fcb_select.cu)
        for (int b = 0; b < batch; ++b) {
            fcb_filter f = h_filters[b];
            const bool has_date = f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX;
            const bool has_cat = f.category_mask != FCB_ALL_CATEGORIES;
            if ((has_date && !s.dates) || (has_cat && !s.categories) || f.date_lo > f.date_hi) f.row_end = f.row_begin;
            const int32_t rc = fcb_search_general(e, d_queries + (size_t)b * s.dim, f, k, d_out_ids + (size_t)b * k,
                                                  d_out_scores + (size_t)b * k, d_out_counts + b, st);
            if (rc != FCB_OK) return rc;
        }
        return FCB_OK;
    }
    ScanParams p{};
    p.vectors = s.vectors; p.inv_norm = e->inv_norm; p.ids = s.ids; p.id_rank = s.id_rank;
    p.dates = s.dates; p.cats = s.categories; p.queries = d_queries;
    p.n_rows = s.n_rows; p.row_base = s.row_base; p.dim = s.dim; p.k = k;
    p.cand = e->d_cand; p.counters = e->d_counters;
    p.out_ids = d_out_ids; p.out_scores = d_out_scores; p.out_counts = d_out_counts;

    // batched dense path: bf16 shard, a large batch whose filters are all the same, a segment big
    // enough to amortise the tensor-core pipeline. Uncertified queries are re-run by the exact scan below.
    if (e->dense && batch >= FCB_DENSE_MIN_BATCH && batch <= 256) {
        bool same = true;
        for (int b = 1; b < batch && same; ++b) same = memcmp(&h_filters[b], &h_filters[0], sizeof(fcb_filter)) == 0;
        fcb_filter f0 = h_filters[0];
        const bool has_date = f0.date_lo != FCB_DATE_MIN || f0.date_hi != FCB_DATE_MAX;
        const bool has_cat = f0.category_mask != FCB_ALL_CATEGORIES;
        if ((has_date && !s.dates) || (has_cat && !s.categories) || f0.date_lo > f0.date_hi) f0.row_end = f0.row_begin;
        int64_t lo = f0.row_begin - s.row_base, hi = f0.row_end - s.row_base;
        if (lo < 0) lo = 0;
        if (hi > s.n_rows) hi = s.n_rows;
        if (same && hi - lo >= 65536) {
            const int32_t *redo = nullptr;
            const int32_t rc = fcb_search_dense(e, d_queries, f0, batch, k, d_out_ids, d_out_scores, d_out_counts, st, &redo);
            if (rc != FCB_OK) return rc;
            p.only_if = redo;
        }
    }

    // resolve filters on the host: clip to the shard, reject predicates on absent columns
    fcb_filter *stage = nullptr;
    p.use_inline = batch <= FCB_INLINE_FILTERS;
    if (!p.use_inline) {
        const int slot = e->filter_slot;
        e->filter_slot = (slot + 1) % FCB_FILTER_SLOTS;
        FCB_CUDA(e, cudaEventSynchronize(e->filter_ev[slot]));      // slot's previous copy has drained
        stage = e->h_filters + (size_t)slot * e->max_batch;
        p.filters = e->d_filters + (size_t)slot * e->max_batch;
    }
    int64_t max_rows = 0;
    for (int b = 0; b < batch; ++b) {
        fcb_filter f = h_filters[b];
        const bool has_date = f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX;
        const bool has_cat = f.category_mask != FCB_ALL_CATEGORIES;
        if ((has_date && !s.dates) || (has_cat && !s.categories) || f.date_lo > f.date_hi) f.row_end = f.row_begin;
        int64_t lo = f.row_begin - s.row_base, hi = f.row_end - s.row_base;
        if (lo < 0) lo = 0;
        if (hi > s.n_rows) hi = s.n_rows;
        if (hi - lo > max_rows) max_rows = hi - lo;
        if (p.use_inline) p.inl[b] = f; else stage[b] = f;
    }
    if (!p.use_inline) {
        const int slot = (e->filter_slot + FCB_FILTER_SLOTS - 1) % FCB_FILTER_SLOTS;
        FCB_CUDA(e, cudaMemcpyAsync(const_cast<fcb_filter *>(p.filters), stage, (size_t)batch * sizeof(fcb_filter),
                                    cudaMemcpyHostToDevice, st));
        FCB_CUDA(e, cudaEventRecord(e->filter_ev[slot], st));
    }
    if (s.dtype == FCB_DTYPE_F32) return launch_fused<float>(e, p, batch, max_rows, st);
    return launch_fused<__nv_bfloat16>(e, p, batch, max_rows, st);
}

extern "C" int32_t fcb_search(fcb_engine *e, const double *queries, const fcb_filter *filters, int32_t batch,
                              int32_t k, uint64_t *out_ids, double *out_scores, int32_t *out_counts, void *stream) {
    if (!e) return FCB_ERR_INVALID;
    int32_t rc = check_search_args(e, queries, filters, batch, k, out_ids, out_scores, out_counts);
    if (rc != FCB_OK) return rc;
    FcbDeviceGuard g(e->device);
    if (!g.ok) return fcb_fail(e, FCB_ERR_CUDA, "fcb_search: cannot select device %d", e->device);
    return search_device(e, queries, filters, batch, k, out_ids, out_scores, out_counts, (cudaStream_t)stream);
}

extern "C" int32_t fcb_search_host(fcb_engine *e, const double *queries, const fcb_filter *filters, int32_t batch,
                                   int32_t k, uint64_t *out_ids, double *out_scores, int32_t *out_counts) {
    if (!e) return FCB_ERR_INVALID;
    int32_t rc = check_search_args(e, queries, filters, batch, k, out_ids, out_scores, out_counts);
    if (rc != FCB_OK) return rc;
    FcbDeviceGuard g(e->device);
    if (!g.ok) return fcb_fail(e, FCB_ERR_CUDA, "fcb_search_host: cannot select device %d", e->device);
    const size_t qn = (size_t)batch * e->shard.dim, rn = (size_t)batch * k;
    memcpy(e->h_q, queries, qn * sizeof(double));
    cudaStream_t st = e->stream;
    FCB_CUDA(e, cudaMemcpyAsync(e->d_q, e->h_q, qn * sizeof(double), cudaMemcpyHostToDevice, st));
    FCB_CUDA(e, cudaEventRecord(e->ev0, st));
    rc = search_device(e, e->d_q, filters, batch, k, e->d_ids, e->d_scores, e->d_counts, st);
    if (rc != FCB_OK) return rc;
    FCB_CUDA(e, cudaEventRecord(e->ev1, st));
    FCB_CUDA(e, cudaMemcpyAsync(e->h_ids, e->d_ids, rn * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    FCB_CUDA(e, cudaMemcpyAsync(e->h_scores, e->d_scores, rn * sizeof(double), cudaMemcpyDeviceToHost, st));
    FCB_CUDA(e, cudaMemcpyAsync(e->h_counts, e->d_counts, (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    FCB_CUDA(e, cudaStreamSynchronize(st));
    memcpy(out_ids, e->h_ids, rn * sizeof(uint64_t));
    memcpy(out_scores, e->h_scores, rn * sizeof(double));
    memcpy(out_counts, e->h_counts, (size_t)batch * sizeof(int32_t));
    float ms = -1.f;
    if (cudaEventElapsedTime(&ms, e->ev0, e->ev1) == cudaSuccess) e->last_ms = ms;
    return FCB_OK;
}

extern "C" float fcb_last_search_ms(const fcb_engine *e) { return e ? e->last_ms : -1.f; }
extern "C" int32_t fcb_last_search_launches(const fcb_engine *e) { return e ? e->last_launches : 0; }
