// Payload-filtered exact cosine top-k: the kernel behind fcb_search (include/fcb200.h), i.e.
// the replacement for client.query_points(...).points at reference tools/qdrant_tool.py:147-153.
//
// Fused path (k <= FCB_FUSED_MAX_K), ONE kernel per call:
//   grid = (CTAs per query, batch). Each warp streams groups of R (2..8) consecutive rows of the
//   user's CSR segment with 128-bit no-allocate loads (HBM-bound: 4*dim bytes per passing row),
//   the date / category predicate is evaluated per row BEFORE its vector is fetched (rows that
//   fail cost 5 bytes, not 4*dim), the float32 dot is reduced by warp shuffles, and the warp
//   keeps its 32 best (score, row) candidates one per lane in registers. Warp lists are merged
//   per CTA through shared memory and written to scratch; the last CTA to arrive for a query
//   (atomic counter) merges the per-CTA lists, re-scores the 32 survivors in float64 with the
//   canonical order of operations (bit-identical to oracle/canon.c), ranks them by
//   (score desc, id asc) and writes the k results. No second launch, no host round trip.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "fcb_internal.cuh"
#include "fcb_topk.cuh"

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
    const int32_t *redo_list;              // optional: device list of query indices to run (re-run of uncertified
    const int32_t *redo_count;             //   queries of the dense path); grid.y rows then stride over that list
    unsigned long long *timing;            // diagnostic: %globaltimer at phase boundaries (NULL in production)
    int32_t *cert_flags;                   // [batch]: 1 = the float32 first pass could not certify this query (fcb_exact.cu re-runs it)
    double cert_eps;                       // fcb_first_pass_eps(dim)
    int64_t n_rows, row_base;
    int dim, k;
    FcbCand *cand;
    unsigned *counters;
    uint64_t *out_ids;
    double *out_scores;
    int32_t *out_counts;
};

__device__ __forceinline__ void stamp(const ScanParams &p, int slot) {
    if (p.timing != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.timing[slot] = t;
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

// Tail of the fused scan, executed by the LAST CTA of a query only.
template <typename T>
__device__ __noinline__ void scan_tail(const ScanParams &p, int b, unsigned char *smem_raw,
                                       FcbCand (*m_s)[FCB_FUSED_CAND], double *f_score, uint64_t *f_id) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t *__restrict__ id_rank = p.id_rank;
    const T *__restrict__ vec = reinterpret_cast<const T *>(p.vectors);
    float my_s;
    uint32_t my_r;
    my_s = -INFINITY; my_r = FCB_NO_ROW;
    const FcbCand *lists = p.cand + (size_t)b * gridDim.x * FCB_FUSED_CAND;
    {
        float s1 = -INFINITY;
        uint32_t r1 = FCB_NO_ROW;
        // two lists per step, and the NEXT step's two lists are requested before this step's are merged: a warp of the last
        // CTA folds up to 37 lists (2 CTAs per SM), and one L2 round trip per step used to be most of the 16 us that took
        unsigned L = warp;
        auto list_of = [&](unsigned l) { return __ldcg(reinterpret_cast<const uint2 *>(lists + (size_t)l * FCB_FUSED_CAND + lane)); };
        bool have = L + FCB_SCAN_WARPS < gridDim.x;
        uint2 a = make_uint2(0u, 0u), c = make_uint2(0u, 0u);
        if (have) { a = list_of(L); c = list_of(L + FCB_SCAN_WARPS); }
        while (have) {
            const unsigned Ln = L + 2 * FCB_SCAN_WARPS;
            const bool have_n = Ln + FCB_SCAN_WARPS < gridDim.x;
            uint2 an = make_uint2(0u, 0u), cn = make_uint2(0u, 0u);
            if (have_n) { an = list_of(Ln); cn = list_of(Ln + FCB_SCAN_WARPS); }
            fcb_warp_merge_list2(my_s, my_r, FcbCand{__uint_as_float(a.x), a.y}, s1, r1, FcbCand{__uint_as_float(c.x), c.y},
                             id_rank, lane);
            L = Ln; a = an; c = cn; have = have_n;
        }
        if (L < gridDim.x) {
            const uint2 a = __ldcg(reinterpret_cast<const uint2 *>(lists + (size_t)L * FCB_FUSED_CAND + lane));
            fcb_warp_merge_list(my_s, my_r, FcbCand{__uint_as_float(a.x), a.y}, id_rank, lane);
        }
        fcb_warp_merge_list(my_s, my_r, FcbCand{s1, r1}, id_rank, lane);
    }
    __syncthreads();                       // m_s reuse: every warp finished reading it above
    fcb_cta_merge_tree(my_s, my_r, m_s, id_rank, warp, lane);
    stamp(p, 5);

    // ---- canonical float64 re-score of the 32 survivors: the float64 query goes to shared memory,
    //      each warp scores its 4 candidates with interleaved chains (one memory latency, not four)
    double *qd_s = reinterpret_cast<double *>(smem_raw + (((size_t)p.dim * sizeof(float) + 15) & ~(size_t)15));
    const double qq = fcb_canon_qq_warp(qd_s, p.dim, lane);       // (staged by scan_one_query when the kernel started)
    {
        const T *rows4[4];
        uint32_t rr[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            rr[c] = m_s[0][warp * 4 + c].row;
            rows4[c] = rr[c] != FCB_NO_ROW ? vec + (int64_t)rr[c] * p.dim : nullptr;
        }
        double sc4[4];
        stamp(p, 6);
        fcb_canon_score_warp4<T>(rows4, qd_s, qq, p.dim, lane, sc4);
        stamp(p, 7);
        if (lane < 4) {
            double sc = sc4[0];
            uint32_t row = rr[0];
#pragma unroll
            for (int c = 1; c < 4; ++c) if (lane == c) { sc = sc4[c]; row = rr[c]; }
            const bool ok = row != FCB_NO_ROW;
            f_score[warp * 4 + lane] = ok ? sc : -INFINITY;
            f_id[warp * 4 + lane] = ok ? (p.ids ? p.ids[row] : (uint64_t)(p.row_base + row)) : UINT64_MAX;
        }
    }
    __syncthreads();
    if (warp == 0) {
        double kth;
        int nvalid;
        fcb_rank_and_write(m_s[0], f_score, f_id, p.k, lane, p.out_ids + (size_t)b * p.k, p.out_scores + (size_t)b * p.k,
                           p.out_counts + b, kth, nvalid);
        if (lane == 0) {
            p.counters[b] = 0;             // ready for the next call
            // Certificate. Rows that are NOT among the 32 survivors have a first-pass score <= the worst kept one, so
            // their canonical score is at most worst / |q| + eps: if the exact k-th score clears that, no such row can
            // belong to the top k (or tie with its last entry). Fewer than 32 survivors: every passing row was kept.
            if (p.cert_flags != nullptr) {
                int flag = 0;
                if (nvalid == FCB_FUSED_CAND) {
                    const double nq = __dsqrt_rn(qq);
                    const float worst = m_s[0][FCB_FUSED_CAND - 1].score;
                    const bool domain = nq > 8.9e-16 && nq < 1.1e15 && isfinite(worst);   // 2^-50 .. 2^50
                    flag = !(domain && kth > (double)worst / nq + p.cert_eps);
                }
                p.cert_flags[b] = flag;
            }
        }
    }
    stamp(p, 8);
}

#ifndef FCB_SCAN_MINCTAS
#define FCB_SCAN_MINCTAS 2
#endif

// R = rows per warp step, U = unroll of the column loop: R * U independent 128-bit loads per lane in
// flight. The streaming rate of this loop is sensitive to (R, U) — see profiles/r01_summary.md — so the
// host picks the pair per storage type and dimension (launch_fused) from the instantiated set.
template <typename T, int R, int U>
__device__ __forceinline__ void scan_one_query(const ScanParams &p, const int b, unsigned char *smem_raw,
                                               FcbCand (*m_s)[FCB_FUSED_CAND], double *f_score, uint64_t *f_id, int &s_last) {
    constexpr int VEC = RowDot<T>::VEC;
    float *q_s = reinterpret_cast<float *>(smem_raw);                 // [dim] float32 query
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const fcb_filter f = p.use_inline ? p.inl[b] : p.filters[b];
    const double *__restrict__ qd = p.queries + (size_t)b * p.dim;
    const uint32_t *__restrict__ id_rank = p.id_rank;

    if (blockIdx.x == 0) stamp(p, 0);
    // (the float64 copy is what the LAST CTA re-scores with; every CTA keeps one — the values are in registers here anyway —
    // so that the tail does not start with a load of its own)
    double *qd_s0 = reinterpret_cast<double *>(smem_raw + (((size_t)p.dim * sizeof(float) + 15) & ~(size_t)15));
    for (int i = tid; i < p.dim; i += FCB_SCAN_THREADS) { const double x = qd[i]; q_s[i] = (float)x; qd_s0[i] = x; }
    __syncthreads();
    if (blockIdx.x == 0) stamp(p, 1);

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
#pragma unroll U
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
                    fcb_warp_insert(my_s, my_r, s, (uint32_t)(r0 + r), id_rank, lane);
                    thr = __shfl_sync(FCB_FULL, my_s, 31);
                }
            }
        }
        pass = npass; inv = ninv; g = gn;
    }

    // ---- CTA merge: 8 warp lists -> warp 0 (tree through shared memory)
    fcb_cta_merge_tree(my_s, my_r, m_s, id_rank, warp, lane);
    if (blockIdx.x == 0) stamp(p, 2);
    if (warp == 0) p.cand[((size_t)b * gridDim.x + blockIdx.x) * FCB_FUSED_CAND + lane] = FcbCand{my_s, my_r};
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        const unsigned prev = atomicAdd(&p.counters[b], 1u);
        s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (blockIdx.x == 0) stamp(p, 3);
    if (!s_last) return;
    __threadfence();
    stamp(p, 4);

    // ---- last CTA of this query: merge, re-score, rank, write. Kept out of line so that its register
    //      needs (float64 chains) do not disturb the allocation / scheduling of the streaming loop above.
    scan_tail<T>(p, b, smem_raw, m_s, f_score, f_id);
}

template <typename T, int R, int U>
__global__ void __launch_bounds__(FCB_SCAN_THREADS, FCB_SCAN_MINCTAS)
scan_topk_kernel(const __grid_constant__ ScanParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ FcbCand m_s[FCB_SCAN_WARPS][FCB_FUSED_CAND];
    __shared__ double f_score[FCB_FUSED_CAND];
    __shared__ uint64_t f_id[FCB_FUSED_CAND];
    __shared__ int s_last;
    if (p.redo_list == nullptr) {
        scan_one_query<T, R, U>(p, blockIdx.y, smem_raw, m_s, f_score, f_id, s_last);
    } else {
        // re-run of the few queries the dense pass could not certify: grid.y rows stride over the device list
        const int n = *p.redo_count;
        for (int i = blockIdx.y; i < n; i += gridDim.y) {
            scan_one_query<T, R, U>(p, p.redo_list[i], smem_raw, m_s, f_score, f_id, s_last);
            __syncthreads();
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

template <typename T, int R, int U>
static int32_t launch_fused_cfg(fcb_engine *e, ScanParams &p, int batch, int64_t max_rows, cudaStream_t st) {
    // a launch over a device list of query indices: the dense path's re-run takes 8 rows (the list is usually empty),
    // the routed search one row per listed query
    const int grid_y = p.redo_list != nullptr ? (e->list_rows > 0 ? e->list_rows : (batch > 8 ? 8 : batch)) : batch;
    const int64_t groups = (max_rows + R - 1) / R;
    // >= 2 groups per warp when the segment allows; never more CTAs than the scratch holds
    int64_t ctas = (groups + 2 * FCB_SCAN_WARPS - 1) / (2 * FCB_SCAN_WARPS);
    int64_t budget = (int64_t)e->cand_lists / batch;          // cand_lists >= max_batch >= batch
    const int64_t resident = (int64_t)e->num_sms * FCB_SCAN_MINCTAS;
    // the grid is (CTAs per query, batch); queries whose segment misses this shard, or that the dense
    // pass already certified, leave at once, so a large segment always gets up to one persistent wave
    if (budget > resident) budget = resident;
    if (ctas > budget) ctas = budget;
    if (ctas < 1) ctas = 1;
    // float32 query for the scan + float64 query for the last CTA's re-score
    const size_t smem = (((size_t)p.dim * sizeof(float) + 15) & ~(size_t)15) + (size_t)p.dim * sizeof(double);
    auto kern = scan_topk_kernel<T, R, U>;
    if (smem > 48 * 1024) FCB_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3((unsigned)ctas, (unsigned)grid_y), FCB_SCAN_THREADS, smem, st>>>(p);
    FCB_CUDA(e, cudaGetLastError());
    e->last_launches += 1;
    return FCB_OK;
}

template <typename T>
static int32_t dispatch_fused(int R, int U, fcb_engine *e, ScanParams &p, int batch, int64_t max_rows, cudaStream_t st) {
#define FCB_TRY(r, u) if (R == r && U == u) return launch_fused_cfg<T, r, u>(e, p, batch, max_rows, st)
    FCB_TRY(2, 4); FCB_TRY(2, 6); FCB_TRY(2, 8); FCB_TRY(4, 2); FCB_TRY(4, 3); FCB_TRY(4, 4); FCB_TRY(8, 2);
#undef FCB_TRY
    return launch_fused_cfg<T, 4, 2>(e, p, batch, max_rows, st);
}

// (R, U) = rows per warp step x unroll of the column loop = loads in flight per lane. Which pair streams fastest depends on
// the storage type, the row length AND on incidental code generation: on B200 12 loads in flight usually win and 16 fall
// off a cliff (10.8 vs 8.28 ms on 10M x 1536 fp32), but the best pair for bf16 x 1024 moved from (8,2) to (4,2) between two
// builds of the same loop (5.0 vs 3.76 ms). So fcb_load_shard MEASURES the candidates on the shard itself (fcb_tune_scan,
// shards of at least 1M rows: ~20 scans) and the engine keeps the winner; every pair computes bit-identical results.
// FCB_SCAN_CFG="R,U" forces a pair, FCB_SCAN_TUNE=0 keeps the table below (what won the last offline sweep).
static int g_env_r = -1, g_env_u = -1, g_env_tune = -1;
static void read_scan_env() {
    if (g_env_r >= 0) return;
    g_env_r = 0;
    if (const char *c = getenv("FCB_SCAN_CFG")) sscanf(c, "%d,%d", &g_env_r, &g_env_u);
    const char *t = getenv("FCB_SCAN_TUNE");
    g_env_tune = !(t && t[0] == '0');
}

template <typename T>
static int32_t launch_fused(fcb_engine *e, ScanParams &p, int batch, int64_t max_rows, cudaStream_t st) {
    constexpr int VEC = 16 / (int)sizeof(T);
    const int steps = (p.dim / VEC + 31) / 32;                // column-loop iterations per row
    int R, U;
    if (sizeof(T) == 4) {
        if (steps % 6 == 0) { R = 2; U = 6; }
        else if (steps % 3 == 0) { R = 4; U = 3; }
        else if (steps % 4 == 0) { R = 2; U = 4; }
        else { R = 4; U = 2; }
    } else {
        if (steps % 3 == 0) { R = 4; U = 3; }
        else { R = 4; U = 2; }
    }
    read_scan_env();
    if (g_env_r > 0) { R = g_env_r; U = g_env_u; }
    else if (e->scan_r > 0) { R = e->scan_r; U = e->scan_u; }
    return dispatch_fused<T>(R, U, e, p, batch, max_rows, st);
}

static unsigned long long *g_scan_timing = nullptr;

// Diagnostic: the next fused-scan launches write %globaltimer (ns) at nine phase boundaries of
// query 0 into `buf` (device uint64[9]); NULL switches it off.
extern "C" int32_t fcb_debug_scan_timing(uint64_t *buf) {
    g_scan_timing = reinterpret_cast<unsigned long long *>(buf);
    return FCB_OK;
}

// The host-side view of one call's filters: clipped to what the shard can answer (a predicate on an absent column or an
// empty date range matches nothing), the largest segment, and where the kernels find them — passed by value for small
// batches, else (or when `need_device`) in a slot of the pinned -> device ring.
struct StagedFilters {
    fcb_filter inl[FCB_INLINE_FILTERS];
    const fcb_filter *d_filters = nullptr;
    const fcb_filter *h_stage = nullptr;   // the pinned copy behind d_filters
    int use_inline = 0;
    int64_t max_rows = 0;
    bool same = true;                      // every filter equals the first
};

static int32_t stage_filters(fcb_engine *e, const fcb_filter *h_filters, int32_t batch, bool need_device, cudaStream_t st,
                             StagedFilters &sf) {
    const fcb_shard &s = e->shard;
    sf.use_inline = batch <= FCB_INLINE_FILTERS;
    const bool to_device = need_device || !sf.use_inline;
    fcb_filter *stage = nullptr;
    int slot = 0;
    if (to_device) {
        slot = e->filter_slot;
        e->filter_slot = (slot + 1) % FCB_FILTER_SLOTS;
        FCB_CUDA(e, cudaEventSynchronize(e->filter_ev[slot]));      // slot's previous copy has drained
        stage = e->h_filters + (size_t)slot * e->max_batch;
        sf.d_filters = e->d_filters + (size_t)slot * e->max_batch;
        sf.h_stage = stage;
    }
    for (int b = 0; b < batch; ++b) {
        fcb_filter f = h_filters[b];
        const bool has_date = f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX;
        const bool has_cat = f.category_mask != FCB_ALL_CATEGORIES;
        if ((has_date && !s.dates) || (has_cat && !s.categories) || f.date_lo > f.date_hi) f.row_end = f.row_begin;
        int64_t lo = f.row_begin - s.row_base, hi = f.row_end - s.row_base;
        if (lo < 0) lo = 0;
        if (hi > s.n_rows) hi = s.n_rows;
        if (hi - lo > sf.max_rows) sf.max_rows = hi - lo;
        if (b > 0 && memcmp(&h_filters[b], &h_filters[0], sizeof(fcb_filter)) != 0) sf.same = false;
        if (sf.use_inline) sf.inl[b] = f;
        if (to_device) stage[b] = f;
    }
    if (to_device) {
        FCB_CUDA(e, cudaMemcpyAsync(const_cast<fcb_filter *>(sf.d_filters), stage, (size_t)batch * sizeof(fcb_filter),
                                    cudaMemcpyHostToDevice, st));
        FCB_CUDA(e, cudaEventRecord(e->filter_ev[slot], st));
    }
    return FCB_OK;
}

// `d_flags` ([batch] device int32) receives the certificate flags; NULL = the engine's own array. With `defer_fallback`
// (the synchronous host entry point, which sees the flags anyway) the exact fallback of the fused path is NOT enqueued:
// *fallback_pending tells the caller to look at the flags and call fcb_exact_enqueue itself if one is set.
static int32_t search_device(fcb_engine *e, const double *d_queries, const fcb_filter *h_filters, int32_t batch,
                             int32_t k, uint64_t *d_out_ids, double *d_out_scores, int32_t *d_out_counts,
                             cudaStream_t st, int32_t *d_flags, bool defer_fallback, bool *fallback_pending,
                             StagedFilters *sf_out) {
    const fcb_shard &s = e->shard;
    e->last_launches = 0;
    if (fallback_pending) *fallback_pending = false;
    if (!d_flags) d_flags = fcb_exact_flags(e);
    StagedFilters sf_local;
    StagedFilters &sf = sf_out ? *sf_out : sf_local;
    {
        const int32_t rc = stage_filters(e, h_filters, batch, k > FCB_FUSED_MAX_K && batch > 1, st, sf);
        if (rc != FCB_OK) return rc;
    }
    if (e->rows_out_of_domain > 0) {
        // rows with norms outside the float32 first pass's range: no error bound, so every query is answered exactly
        FCB_CUDA(e, cudaMemsetAsync(d_flags, 1, (size_t)batch * sizeof(int32_t), st));
        return fcb_exact_enqueue(e, d_queries, sf.d_filters, sf.inl, sf.use_inline ? 1 : 0, batch, k, d_flags, d_out_ids,
                                 d_out_scores, d_out_counts, st);
    }
    if (k > FCB_FUSED_MAX_K) {
        // a batch of small segments (every one fits one sorted chunk): ONE launch for all of them, exact by construction
        if (batch > 1 && e->d_bsel && sf.max_rows <= 1024)
            return fcb_search_general_small_batch(e, d_queries, sf.d_filters, batch, k, sf.max_rows, d_out_ids, d_out_scores,
                                                  d_out_counts, st);
        // general-k path: one query at a time, stream-ordered (fcb_select.cu); certificate + exact fallback per query
        for (int b = 0; b < batch; ++b) {
            const fcb_filter &f = sf.use_inline ? sf.inl[b] : sf.h_stage[b];
            const int32_t rc = fcb_search_general(e, d_queries + (size_t)b * s.dim, f, k, d_out_ids + (size_t)b * k,
                                                  d_out_scores + (size_t)b * k, d_out_counts + b, d_flags + b, st);
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
    p.timing = g_scan_timing;
    p.cert_flags = d_flags; p.cert_eps = fcb_first_pass_eps(s.dim);

    // batched dense path: bf16 shard, a large batch whose filters are all the same, a segment big
    // enough to amortise the tensor-core pipeline. Uncertified queries are re-run by the exact scan below.
    bool dense_ran = false;
    if (e->dense && batch >= FCB_DENSE_MIN_BATCH && batch <= 256 && sf.same && sf.max_rows >= 65536) {
        fcb_filter f0 = h_filters[0];
        const bool has_date = f0.date_lo != FCB_DATE_MIN || f0.date_hi != FCB_DATE_MAX;
        const bool has_cat = f0.category_mask != FCB_ALL_CATEGORIES;
        if ((has_date && !s.dates) || (has_cat && !s.categories) || f0.date_lo > f0.date_hi) f0.row_end = f0.row_begin;
        // the re-run scan writes the flags of the queries it runs only: the certified ones must read 0
        FCB_CUDA(e, cudaMemsetAsync(d_flags, 0, (size_t)batch * sizeof(int32_t), st));
        const int32_t rc = fcb_search_dense(e, d_queries, f0, batch, k, d_out_ids, d_out_scores, d_out_counts, st,
                                            &p.redo_list, &p.redo_count);
        if (rc != FCB_OK) return rc;
        dense_ran = true;
    }

    p.use_inline = sf.use_inline;
    p.filters = sf.d_filters;
    if (sf.use_inline) for (int b = 0; b < batch; ++b) p.inl[b] = sf.inl[b];
    int32_t rc = s.dtype == FCB_DTYPE_F32 ? launch_fused<float>(e, p, batch, sf.max_rows, st)
                                          : launch_fused<__nv_bfloat16>(e, p, batch, sf.max_rows, st);
    if (rc != FCB_OK) return rc;
    if (sf.max_rows < FCB_FUSED_CAND) return FCB_OK;           // every passing row is a survivor: certified by construction
    if (defer_fallback && !dense_ran) {
        if (fallback_pending) *fallback_pending = true;
        return FCB_OK;
    }
    return fcb_exact_enqueue(e, d_queries, sf.d_filters, sf.inl, sf.use_inline ? 1 : 0, batch, k, d_flags, d_out_ids,
                             d_out_scores, d_out_counts, st);
}

// Routed search + owner-computes exchange (include/fcb200.h, csrc/fcb_exchange.cu): scan only the queries whose row range
// lies in THIS rank's shard, then push their results to every rank and wait for everybody else's.
extern "C" int32_t fcb_search_routed(fcb_engine *e, const fcb_exchange *x, const double *queries, const fcb_filter *filters,
                                     int32_t batch, int32_t k, uint32_t seq, int32_t *status, void *stream) {
    if (!e) return FCB_ERR_INVALID;
    if (!x || !x->cuts || !status) return fcb_fail(e, FCB_ERR_INVALID, "fcb_search_routed: NULL argument");
    if (x->world < 1 || x->world > FCB_MAX_WORLD || x->rank < 0 || x->rank >= x->world)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_search_routed: world %d / rank %d outside 1..%d", x->world, x->rank, FCB_MAX_WORLD);
    for (int g = 0; g < x->world; ++g)
        if (!x->bufs[g]) return fcb_fail(e, FCB_ERR_INVALID, "fcb_search_routed: exchange buffer of rank %d is NULL", g);
    if (seq == 0) return fcb_fail(e, FCB_ERR_INVALID, "fcb_search_routed: sequence numbers start at 1");
    int32_t rc = check_search_args(e, queries, filters, batch, k, status, status, status);
    if (rc != FCB_OK) return rc;
    if (k > FCB_FUSED_MAX_K)
        return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_search_routed: k %d > %d (use fcb_search + all-gather + fcb_merge_topk)", k, FCB_FUSED_MAX_K);
    if (e->rows_out_of_domain > 0)
        return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_search_routed: the shard holds rows outside the float32 range (use fcb_search)");
    FcbDeviceGuard g(e->device);
    if (!g.ok) return fcb_fail(e, FCB_ERR_CUDA, "fcb_search_routed: cannot select device %d", e->device);
    cudaStream_t st = (cudaStream_t)stream;
    const fcb_shard &s = e->shard;
    e->last_launches = 0;

    // ---- who owns what: the shard holding the first row of the range; a range spanning two shards needs a merge
    const int rslot = e->route_slot;
    e->route_slot = (rslot + 1) % FCB_FILTER_SLOTS;
    FCB_CUDA(e, cudaEventSynchronize(e->route_ev[rslot]));
    const size_t rstride = 2 * (size_t)e->max_batch + 1;
    int32_t *h_owner = e->h_route + (size_t)rslot * rstride, *h_subset = h_owner + e->max_batch, *h_count = h_subset + e->max_batch;
    int32_t *d_owner = e->d_route + (size_t)rslot * rstride, *d_subset = d_owner + e->max_batch, *d_count = d_subset + e->max_batch;
    int n_mine = 0;
    for (int b = 0; b < batch; ++b) {
        const fcb_filter &f = filters[b];
        int owner = 0;
        if (f.row_begin < f.row_end) {
            int lo = 0, hi = x->world - 1;                         // last shard whose first row is <= row_begin
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (x->cuts[mid] <= f.row_begin) lo = mid; else hi = mid - 1; }
            owner = lo;
            if (f.row_end > x->cuts[owner + 1] || f.row_begin < x->cuts[0])
                return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_search_routed: query %d spans rows [%lld, %lld) of more than one shard "
                                "(answer it with fcb_search + all-gather + fcb_merge_topk)", b, (long long)f.row_begin, (long long)f.row_end);
        }
        h_owner[b] = owner;
        if (owner == x->rank) h_subset[n_mine++] = b;
    }
    *h_count = n_mine;
    FCB_CUDA(e, cudaMemcpyAsync(d_owner, h_owner, rstride * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    FCB_CUDA(e, cudaEventRecord(e->route_ev[rslot], st));

    // local results: the engine's result staging block (ids | scores | counts | flags), sized for max_batch x max_k
    const size_t rn = (size_t)batch * k;
    uint64_t *d_ids = e->d_ids;
    double *d_scores = reinterpret_cast<double *>(d_ids + rn);
    int32_t *d_counts = reinterpret_cast<int32_t *>(d_scores + rn), *d_flags = d_counts + batch;
    if (n_mine > 0) {
        StagedFilters sf;
        rc = stage_filters(e, filters, batch, true, st, sf);
        if (rc != FCB_OK) return rc;
        ScanParams p{};
        p.vectors = s.vectors; p.inv_norm = e->inv_norm; p.ids = s.ids; p.id_rank = s.id_rank;
        p.dates = s.dates; p.cats = s.categories; p.queries = queries;
        p.n_rows = s.n_rows; p.row_base = s.row_base; p.dim = s.dim; p.k = k;
        p.cand = e->d_cand; p.counters = e->d_counters;
        p.out_ids = d_ids; p.out_scores = d_scores; p.out_counts = d_counts;
        p.cert_flags = d_flags; p.cert_eps = fcb_first_pass_eps(s.dim);
        p.use_inline = 0; p.filters = sf.d_filters;
        p.redo_list = d_subset; p.redo_count = d_count;
        e->list_rows = n_mine;
        rc = s.dtype == FCB_DTYPE_F32 ? launch_fused<float>(e, p, batch, sf.max_rows, st)
                                      : launch_fused<__nv_bfloat16>(e, p, batch, sf.max_rows, st);
        e->list_rows = 0;
        if (rc != FCB_OK) return rc;
        if (sf.max_rows >= FCB_FUSED_CAND) {
            rc = fcb_exact_enqueue(e, queries, sf.d_filters, sf.inl, 0, batch, k, d_flags, d_ids, d_scores, d_counts, st, d_subset,
                                   d_count);                          // looks at the flags of the listed queries only
            if (rc != FCB_OK) return rc;
        }
    }
    rc = fcb_exchange_launch(e, x, batch, k, seq, d_owner, d_ids, d_scores, d_counts, status, st);
    if (rc != FCB_OK) return rc;
    e->last_launches += 1;
    return FCB_OK;
}

// Called by fcb_load_shard: time every (R, U) candidate on a full-shard scan with a pseudo-random query (results discarded)
// and keep the fastest for this shard. Needs the host-call staging buffers (query, results), so it runs after they exist.
int32_t fcb_tune_scan(fcb_engine *e) {
    const fcb_shard &s = e->shard;
    e->scan_r = e->scan_u = 0;
    read_scan_env();
    if (!g_env_tune || g_env_r > 0 || s.n_rows < (1 << 20) || !e->d_q || !e->d_ids) return FCB_OK;
    uint64_t x = 0x9E3779B97F4A7C15ull;                       // splitmix64: any query without ties will do
    for (int i = 0; i < s.dim; ++i) {
        x += 0x9E3779B97F4A7C15ull;
        uint64_t z = x;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
        e->h_q[i] = (double)(int64_t)(z >> 11) / 9007199254740992.0 - 0.5;
    }
    cudaStream_t st = e->stream;
    FCB_CUDA(e, cudaMemcpyAsync(e->d_q, e->h_q, (size_t)s.dim * sizeof(double), cudaMemcpyHostToDevice, st));
    const int k = e->max_k < 10 ? e->max_k : 10;
    ScanParams p{};
    p.vectors = s.vectors; p.inv_norm = e->inv_norm; p.ids = s.ids; p.id_rank = s.id_rank;
    p.dates = s.dates; p.cats = s.categories; p.queries = e->d_q;
    p.n_rows = s.n_rows; p.row_base = s.row_base; p.dim = s.dim; p.k = k;
    p.cand = e->d_cand; p.counters = e->d_counters;
    p.out_ids = e->d_ids;
    p.out_scores = reinterpret_cast<double *>(e->d_ids + k);
    p.out_counts = reinterpret_cast<int32_t *>(p.out_scores + k);
    p.use_inline = true;
    p.inl[0] = fcb_filter{s.row_base, s.row_base + s.n_rows, FCB_DATE_MIN, FCB_DATE_MAX, FCB_ALL_CATEGORIES};
    static const int cand[][2] = {{2, 4}, {2, 6}, {2, 8}, {4, 2}, {4, 3}, {4, 4}, {8, 2}};
    // Two rounds over all candidates, each launch timed on its own, a candidate's score = its FASTEST launch: the clocks
    // drift under the power cap right after a load, and a single noisy interval once picked a shape 9 % slower.
    constexpr int N_CAND = sizeof(cand) / sizeof(cand[0]);
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
    for (auto &x : ev) FCB_CUDA(e, cudaEventCreate(&x));
    float fastest[N_CAND];
    for (float &f : fastest) f = 1e30f;
    int32_t rc = FCB_OK;
    for (int round = 0; round < 2 && rc == FCB_OK; ++round) {
        for (int ci = 0; ci < N_CAND && rc == FCB_OK; ++ci) {
            const int r = cand[ci][0], u = cand[ci][1];
            auto run = [&]() { return s.dtype == FCB_DTYPE_F32 ? dispatch_fused<float>(r, u, e, p, 1, s.n_rows, st)
                                                               : dispatch_fused<__nv_bfloat16>(r, u, e, p, 1, s.n_rows, st); };
            rc = run();                                         // warm-up (instruction cache, clocks)
            for (int i = 0; i < 2 && rc == FCB_OK; ++i) {
                cudaEventRecord(ev[i], st);
                rc = run();
            }
            cudaEventRecord(ev[2], st);
            if (rc != FCB_OK) break;
            if (cudaEventSynchronize(ev[2]) != cudaSuccess) { rc = fcb_fail(e, FCB_ERR_CUDA, "fcb_load_shard: scan tuning failed"); break; }
            for (int i = 0; i < 2; ++i) {
                float ms = 1e30f;
                if (cudaEventElapsedTime(&ms, ev[i], ev[i + 1]) == cudaSuccess && ms < fastest[ci]) fastest[ci] = ms;
            }
        }
    }
    for (auto &x : ev) cudaEventDestroy(x);
    int best_r = 0, best_u = 0;
    float best = 1e30f;
    for (int ci = 0; ci < N_CAND; ++ci)
        if (fastest[ci] < best) { best = fastest[ci]; best_r = cand[ci][0]; best_u = cand[ci][1]; }
    e->last_launches = 0;
    if (rc != FCB_OK) return rc;
    e->scan_r = best_r; e->scan_u = best_u;
    return FCB_OK;
}

extern "C" int32_t fcb_scan_config(const fcb_engine *e, int32_t *r, int32_t *u) {
    if (!e || !r || !u) return FCB_ERR_INVALID;
    *r = e->scan_r; *u = e->scan_u;
    return FCB_OK;
}

extern "C" int32_t fcb_search(fcb_engine *e, const double *queries, const fcb_filter *filters, int32_t batch,
                              int32_t k, uint64_t *out_ids, double *out_scores, int32_t *out_counts, void *stream) {
    if (!e) return FCB_ERR_INVALID;
    int32_t rc = check_search_args(e, queries, filters, batch, k, out_ids, out_scores, out_counts);
    if (rc != FCB_OK) return rc;
    FcbDeviceGuard g(e->device);
    if (!g.ok) return fcb_fail(e, FCB_ERR_CUDA, "fcb_search: cannot select device %d", e->device);
    return search_device(e, queries, filters, batch, k, out_ids, out_scores, out_counts, (cudaStream_t)stream, nullptr, false,
                         nullptr, nullptr);
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
    // results are laid out for THIS call's (batch, k) inside the one staging block: ids | scores | counts | certificate flags
    uint64_t *d_ids = e->d_ids, *h_ids = e->h_ids;
    double *d_scores = reinterpret_cast<double *>(d_ids + rn), *h_scores = reinterpret_cast<double *>(h_ids + rn);
    int32_t *d_counts = reinterpret_cast<int32_t *>(d_scores + rn), *h_counts = reinterpret_cast<int32_t *>(h_scores + rn);
    int32_t *d_flags = d_counts + batch, *h_flags = h_counts + batch;
    bool pending = false;
    StagedFilters sf;
    // a few single queries through the fused scan (the latency-sensitive call: one worker, one question): the kernel's last
    // CTA writes ids, scores, counts and certificate flags STRAIGHT into the pinned staging block (mapped into the device's
    // address space like all pinned memory), so no device-to-host copy is enqueued behind the kernel
    static const bool direct_ok = []() { const char *v = getenv("FCB_HOST_DIRECT"); return !(v && v[0] == '0'); }();
    const bool direct = direct_ok && batch <= 4 && k <= FCB_FUSED_MAX_K && e->rows_out_of_domain == 0;
    if (direct) { d_ids = h_ids; d_scores = h_scores; d_counts = h_counts; d_flags = h_flags; }
    rc = search_device(e, e->d_q, filters, batch, k, d_ids, d_scores, d_counts, st, d_flags, true, &pending, &sf);
    if (rc != FCB_OK) return rc;
    FCB_CUDA(e, cudaEventRecord(e->ev1, st));
    const size_t res_bytes = rn * 16 + (size_t)batch * sizeof(int32_t) * (pending ? 2 : 1);
    if (!direct) FCB_CUDA(e, cudaMemcpyAsync(h_ids, d_ids, res_bytes, cudaMemcpyDeviceToHost, st));
    FCB_CUDA(e, cudaStreamSynchronize(st));
    if (pending) {
        // the fused scan could not certify some query (the flags came back with the results): answer those exactly
        bool any = false;
        for (int b = 0; b < batch; ++b) any = any || h_flags[b] != 0;
        if (any) {
            rc = fcb_exact_enqueue(e, e->d_q, sf.d_filters, sf.inl, sf.use_inline ? 1 : 0, batch, k, d_flags, d_ids, d_scores,
                                   d_counts, st);
            if (rc != FCB_OK) return rc;
            FCB_CUDA(e, cudaEventRecord(e->ev1, st));
            if (!direct) FCB_CUDA(e, cudaMemcpyAsync(h_ids, d_ids, rn * 16 + (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            FCB_CUDA(e, cudaStreamSynchronize(st));
        }
    }
    memcpy(out_ids, h_ids, rn * sizeof(uint64_t));
    memcpy(out_scores, h_scores, rn * sizeof(double));
    memcpy(out_counts, h_counts, (size_t)batch * sizeof(int32_t));
    e->last_ms = -2.f;                                          // measured by ev0 / ev1; read out on demand
    return FCB_OK;
}

extern "C" float fcb_last_search_ms(const fcb_engine *e) {
    if (!e) return -1.f;
    if (e->last_ms == -2.f) {                                   // the host call left its two events behind
        fcb_engine *m = const_cast<fcb_engine *>(e);
        float ms = -1.f;
        FcbDeviceGuard g(e->device);
        m->last_ms = g.ok && cudaEventElapsedTime(&ms, e->ev0, e->ev1) == cudaSuccess ? ms : -1.f;
    }
    return e->last_ms;
}
extern "C" int32_t fcb_last_search_launches(const fcb_engine *e) { return e ? e->last_launches : 0; }
