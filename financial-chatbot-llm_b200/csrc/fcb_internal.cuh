// Internal declarations shared by the .cu files of libfcb200.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "fcb200.h"

#define FCB_WARP 32
#define FCB_FULL 0xffffffffu
#define FCB_SCAN_THREADS 256               // 8 warps per CTA
#define FCB_SCAN_WARPS (FCB_SCAN_THREADS / FCB_WARP)
#define FCB_FUSED_CAND 32                  // candidates a warp keeps in registers (one per lane)
#define FCB_FUSED_MAX_K 16                 // fused path: k + margin(>=16) <= 32
#define FCB_INLINE_FILTERS 4               // filters passed by value in the kernel parameters
#define FCB_FILTER_SLOTS 8                 // ring of pinned filter staging slots (batch > inline)
#define FCB_DENSE_MIN_BATCH 32             // batches at least this large may take the tensor-core path
#define FCB_MAX_WORLD 8                    // ranks of one routed exchange (the GPUs of one box)
#define FCB_NO_ROW 0xffffffffu
#define FCB_COS_EPS 1.1920929e-7           // zero-norm guard (float32 eps), as in the oracle

struct FcbCand {                           // one first-pass candidate: float32 score + local row
    float score;
    uint32_t row;
};

struct fcb_engine {
    int device = 0;
    int max_batch = 1;
    int max_k = 10;
    int num_sms = 0;
    cudaStream_t stream = nullptr;         // for the *_host entry points
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;

    bool loaded = false;
    fcb_shard shard{};
    float *inv_norm = nullptr;             // device [n_rows], library-owned

    // scratch (device)
    fcb_filter *d_filters = nullptr;       // [FCB_FILTER_SLOTS][max_batch]
    FcbCand *d_cand = nullptr;             // per-CTA candidate lists of the fused path
    size_t cand_lists = 0;                 // capacity in lists of FCB_FUSED_CAND entries
    unsigned *d_counters = nullptr;        // [max_batch] arrival counters (kept at zero between calls)

    // pinned host staging
    fcb_filter *h_filters = nullptr;       // [FCB_FILTER_SLOTS][max_batch]
    cudaEvent_t filter_ev[FCB_FILTER_SLOTS] = {};
    int filter_slot = 0;
    // routed search (fcb_search_routed): per call, owner[max_batch] | subset[max_batch] | count, staged through the same kind of ring
    int32_t *h_route = nullptr, *d_route = nullptr;
    cudaEvent_t route_ev[FCB_FILTER_SLOTS] = {};
    int route_slot = 0;
    int list_rows = 0;                     // grid rows of a scan launch that strides over a device list (0: the re-run default of 8)

    // staging for fcb_search_host (sized at load: needs dim)
    double *h_q = nullptr, *d_q = nullptr;
    uint64_t *h_ids = nullptr, *d_ids = nullptr;
    double *h_scores = nullptr, *d_scores = nullptr;
    int32_t *h_counts = nullptr, *d_counts = nullptr;

    // general-k path (k > FCB_FUSED_MAX_K), allocated in fcb_load_shard when max_k needs it
    uint64_t *d_keys = nullptr;            // [n_rows] sortable (score32, tie) keys of one query
    struct FcbSelectState *d_sel = nullptr;
    uint32_t *d_sel_rows = nullptr;        // [kprime_max] selected rows
    double *d_sel_scores = nullptr;        // [kprime_max] canonical float64 scores
    uint64_t *d_sel_ids = nullptr;         // [kprime_max]
    int kprime_max = 0;
    bool sel_clean = false;                // the selection state is all zero (left so by the one-launch direct path)
    unsigned *d_bsel = nullptr;            // [max_batch][2] (candidates, tickets) of the batched direct path, kept at zero between calls
    double *d_bsel_scores = nullptr;       // [max_batch][1024]
    uint64_t *d_bsel_ids = nullptr;        // [max_batch][1024]

    int dates_sorted = 0;                  // load-time check: dates non-decreasing within every user
    uint16_t *group_keys = nullptr;        // device [n_rows], library-owned: (category << 10) | (month - key_base) per row, built
    int32_t key_base = 0;                  //   at load when the shard has dates, categories and amounts (fcb_aggregate.cu)
    int scan_r = 0, scan_u = 0;            // fused scan: the (rows per step, unroll) pair measured fastest on this shard (0 = not yet)
    uint32_t *row_of_rank = nullptr;       // inverse of shard.id_rank, built by fcb_aggregate_hits on first use
    void *dense = nullptr;                 // FcbDense* (fcb_dense.cu): batched tensor-core path scratch
    void *exact = nullptr;                 // FcbExact* (fcb_exact.cu): exact float64 fallback scratch + certificate flags
    unsigned long long rows_out_of_domain = 0;   // rows whose norm is outside [2^-50, 2^50] (or not finite): the float32
                                           //   first pass has no error bound for them, every query takes the exact path

    float last_ms = -1.f;
    int last_launches = 0;
    std::string err;
};

// device-resident state of one general-k selection (fcb_select.cu)
struct FcbSelectState {
    unsigned hist[8][256];                 // digit histograms of the 8 radix passes (MSB first)
    unsigned n_pass;                       // rows that passed the filter
    unsigned n_selected;                   // compaction cursor
};

int32_t fcb_general_alloc(fcb_engine *e);  // called by fcb_load_shard
void fcb_general_free(fcb_engine *e);
// every query's segment has at most 1024 rows: the whole batch in one launch (filters already clipped and on the device)
int32_t fcb_search_general_small_batch(fcb_engine *e, const double *d_queries, const fcb_filter *d_filters, int32_t batch,
                                       int32_t k, int64_t max_rows, uint64_t *d_out_ids, double *d_out_scores,
                                       int32_t *d_out_counts, cudaStream_t st);
int32_t fcb_search_general(fcb_engine *e, const double *d_query, const fcb_filter &f, int32_t k,
                           uint64_t *d_out_ids, double *d_out_scores, int32_t *d_out_count, int32_t *d_flag, cudaStream_t st);

int32_t fcb_check_dates_sorted(fcb_engine *e);   // fcb_aggregate.cu
int32_t fcb_build_group_keys(fcb_engine *e);     // fcb_aggregate.cu
int32_t fcb_tune_scan(fcb_engine *e);            // fcb_search.cu: measure the scan's (R, U) candidates on the loaded shard

// batched tensor-core path (fcb_dense.cu)
bool fcb_dense_eligible(const fcb_engine *e);
int32_t fcb_dense_alloc(fcb_engine *e);
void fcb_dense_free(fcb_engine *e);
int fcb_dense_error_code(const fcb_engine *e);   // != 0: a tensor-core kernel trapped on a pipeline-protocol violation
int32_t fcb_search_dense(fcb_engine *e, const double *d_queries, const fcb_filter &f, int32_t batch, int32_t k,
                         uint64_t *d_out_ids, double *d_out_scores, int32_t *d_out_counts, cudaStream_t st,
                         const int32_t **redo_list_out, const int32_t **redo_count_out);

// exact float64 fallback (fcb_exact.cu): answers the queries whose certificate flag a fast path set
int32_t fcb_exact_alloc(fcb_engine *e);
void fcb_exact_free(fcb_engine *e);
int32_t *fcb_exact_flags(fcb_engine *e);   // device [max_batch]
int32_t fcb_exact_enqueue(fcb_engine *e, const double *d_queries, const fcb_filter *d_filters, const fcb_filter *inl,
                          int filters_mode, int32_t batch, int32_t k, const int32_t *d_flags, uint64_t *d_out_ids,
                          double *d_out_scores, int32_t *d_out_counts, cudaStream_t st, const int32_t *d_list = nullptr,
                          const int32_t *d_list_count = nullptr);

// owner-computes exchange over peer memory (fcb_exchange.cu)
__host__ __device__ inline size_t fcb_exchange_parity_bytes(int slots, int k) { return (size_t)slots * k * 16 + (size_t)slots * 8; }
int32_t fcb_exchange_launch(fcb_engine *e, const fcb_exchange *x, int32_t slots, int32_t k, uint32_t seq, const int32_t *d_owner,
                            const uint64_t *d_ids, const double *d_scores, const int32_t *d_counts, int32_t *d_status,
                            cudaStream_t st);

// error plumbing (fcb_engine.cu)
int32_t fcb_fail(fcb_engine *e, int32_t code, const char *fmt, ...);
#define FCB_CUDA(e, call)                                                                   \
    do {                                                                                    \
        cudaError_t _st = (call);                                                           \
        if (_st != cudaSuccess)                                                             \
            return fcb_fail((e), FCB_ERR_CUDA, "%s failed: %s (%s:%d)", #call,              \
                            cudaGetErrorString(_st), __FILE__, __LINE__);                   \
    } while (0)

struct FcbDeviceGuard {                    // make the engine's device current for one call
    int prev = -1;
    bool ok = true;
    explicit FcbDeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; prev = -1; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~FcbDeviceGuard() {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
};

// ---- device helpers ---------------------------------------------------------------------

__device__ __forceinline__ float fcb_bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float fcb_bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

__device__ __forceinline__ float fcb_to_float(float x) { return x; }
__device__ __forceinline__ float fcb_to_float(__nv_bfloat16 x) { return __bfloat162float(x); }

// 128-bit streaming load (read-only path, no L1 allocation), zero when !pred
__device__ __forceinline__ uint4 fcb_ldg_stream(const void *p, bool pred) {
    uint4 v;
    asm("{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\tmov.b32 %2, 0;\n\tmov.b32 %3, 0;\n\t"
        "@p ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];\n\t}"
        : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
        : "l"(p), "r"((int)pred));
    return v;
}

__device__ __forceinline__ float fcb_warp_sum(float v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(FCB_FULL, v, off);
    return v;
}

// Canonical float64 cosine of one row against one query, computed by a full warp.
// Bit-identical to oracle/canon.c (see its header for the order of operations).
template <typename T>
__device__ __forceinline__ double fcb_canon_score_warp(const T *__restrict__ row,
                                                       const double *__restrict__ q, int dim,
                                                       int lane) {
    double dot = 0.0, ss = 0.0, qq = 0.0;
    // 16 row + 16 query loads in flight per lane and round (the order of the fma chain is unchanged): with the default
    // 4 a 1536-wide row costs 12 dependent memory round trips, which is what a small query's latency consisted of
#pragma unroll 16
    for (int j = lane; j < dim; j += FCB_WARP) {
        const double x = (double)fcb_to_float(row[j]);
        const double qv = q[j];
        dot = __fma_rn(x, qv, dot);
        ss = __fma_rn(x, x, ss);
        qq = __fma_rn(qv, qv, qq);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        dot = __dadd_rn(dot, __shfl_xor_sync(FCB_FULL, dot, off));
        ss = __dadd_rn(ss, __shfl_xor_sync(FCB_FULL, ss, off));
        qq = __dadd_rn(qq, __shfl_xor_sync(FCB_FULL, qq, off));
    }
    double nr = __dsqrt_rn(ss), nq = __dsqrt_rn(qq);
    if (nr == 0.0) nr = FCB_COS_EPS;
    if (nq == 0.0) nq = FCB_COS_EPS;
    return __ddiv_rn(dot, __dmul_rn(nr, nq));
}

// The same canonical score for FOUR rows at once (latency-oriented: the last CTA of the fused scan
// re-scores 32 survivors with 8 warps). Per row the order of operations is exactly that of
// fcb_canon_score_warp, so the results are bit-identical; the four chains only interleave. The
// float64 query is read from shared memory and its sum of squares `qq` (same order: lane-strided
// fma, xor butterfly) is computed once by the caller. rows[c] == nullptr skips a slot.
template <typename T>
__device__ __forceinline__ void fcb_canon_score_warp4(const T *const (&rows)[4], const double *__restrict__ q_s,
                                                      double qq, int dim, int lane, double (&out)[4]) {
    double dot[4] = {0.0, 0.0, 0.0, 0.0}, ss[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 8
    for (int j = lane; j < dim; j += FCB_WARP) {
        const double qv = q_s[j];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const double x = rows[c] ? (double)fcb_to_float(rows[c][j]) : 0.0;
            dot[c] = __fma_rn(x, qv, dot[c]);
            ss[c] = __fma_rn(x, x, ss[c]);
        }
    }
    double nq = __dsqrt_rn(qq);
    if (nq == 0.0) nq = FCB_COS_EPS;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            dot[c] = __dadd_rn(dot[c], __shfl_xor_sync(FCB_FULL, dot[c], off));
            ss[c] = __dadd_rn(ss[c], __shfl_xor_sync(FCB_FULL, ss[c], off));
        }
        double nr = __dsqrt_rn(ss[c]);
        if (nr == 0.0) nr = FCB_COS_EPS;
        out[c] = __ddiv_rn(dot[c], __dmul_rn(nr, nq));
    }
}

// sum of squares of the float64 query in the canonical order (lane-strided fma, xor butterfly)
__device__ __forceinline__ double fcb_canon_qq_warp(const double *__restrict__ q, int dim, int lane) {
    double qq = 0.0;
#pragma unroll 8
    for (int j = lane; j < dim; j += FCB_WARP) qq = __fma_rn(q[j], q[j], qq);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) qq = __dadd_rn(qq, __shfl_xor_sync(FCB_FULL, qq, off));
    return qq;
}

// (score desc, id asc)
__device__ __forceinline__ bool fcb_better64(double s1, uint64_t i1, double s2, uint64_t i2) {
    return s1 > s2 || (s1 == s2 && i1 < i2);
}
