// General-k path of fcb_search (16 < k <= 10000; the tool's default limit is 10000,
// reference tools/qdrant_tool.py:145). One query at a time, everything stream-ordered, no host
// round trip:
//   1. scan_keys_kernel      the same predicate + GEMV scan as the fused path, but every row of
//                            the segment gets a sortable 64-bit key (float32 score, tie order);
//                            rows failing the predicate get key 0 and are never fetched.
//                            (+8 bytes written per row next to the 4*dim read: HBM-bound as before)
//   2. radix_hist_kernel x8  MSB-first radix select of the K'-th largest key, K' = k + margin;
//                            each pass re-derives (prefix, remaining) from the earlier
//                            histograms, so no host sync and no extra "pick bucket" launches.
//                            Keys are L2-resident for segments up to ~15M rows.
//   3. compact_kernel        rows with key >= threshold -> candidate list (exactly K' rows)
//   4. rescore_kernel        canonical float64 score per candidate (one warp each)
//   5. sort_chunks_kernel    the candidates sorted by (score desc, id asc) in chunks of 1024 (bitonic, shared memory);
//                            up to 1024 candidates this writes the results and ends the query
//   6. rank_merge_kernel     more than one chunk: rank = position in the own chunk + binary searches in the others
// Segments that fit the candidate buffer whole (an ordinary tenant asked for "everything", the tool's
// default call) skip steps 1-4: direct_rescore_kernel scores every passing row in float64 right away.
#include "fcb_internal.cuh"
#include "fcb_topk.cuh"

#define SEL_THREADS 256
#define SORT_CHUNK 1024                    // candidates sorted per CTA (bitonic, shared memory)

static inline int kprime_for(int k) {
    int margin = k / 16;
    if (margin < 32) margin = 32;
    return k + margin;
}

void fcb_general_free(fcb_engine *e) {
    cudaFree(e->d_keys); cudaFree(e->d_sel); cudaFree(e->d_sel_rows);
    cudaFree(e->d_sel_scores); cudaFree(e->d_sel_ids);
    cudaFree(e->d_bsel); cudaFree(e->d_bsel_scores); cudaFree(e->d_bsel_ids);
    e->d_bsel = nullptr; e->d_bsel_scores = nullptr; e->d_bsel_ids = nullptr;
    e->d_keys = nullptr; e->d_sel = nullptr; e->d_sel_rows = nullptr;
    e->d_sel_scores = nullptr; e->d_sel_ids = nullptr; e->kprime_max = 0;
}

int32_t fcb_general_alloc(fcb_engine *e) {
    e->kprime_max = kprime_for(e->max_k);
    e->sel_clean = false;
    const size_t n = (size_t)e->shard.n_rows, kp = (size_t)e->kprime_max;
    if (cudaMalloc(&e->d_keys, n * sizeof(uint64_t)) != cudaSuccess ||
        cudaMalloc(&e->d_sel, sizeof(FcbSelectState)) != cudaSuccess ||
        cudaMalloc(&e->d_sel_rows, kp * sizeof(uint32_t)) != cudaSuccess ||
        cudaMalloc(&e->d_sel_scores, kp * sizeof(double)) != cudaSuccess ||
        cudaMalloc(&e->d_sel_ids, kp * sizeof(uint64_t)) != cudaSuccess) {
        cudaGetLastError();
        fcb_general_free(e);
        return fcb_fail(e, FCB_ERR_NOMEM, "fcb_load_shard: cannot allocate general-k scratch (%zu bytes of keys)",
                        n * sizeof(uint64_t));
    }
    if (e->max_batch > 1) {                                            // batched direct path: 16 KB of candidates per query
        const size_t b = (size_t)e->max_batch;
        if (cudaMalloc(&e->d_bsel, b * 2 * sizeof(unsigned)) != cudaSuccess ||
            cudaMalloc(&e->d_bsel_scores, b * SORT_CHUNK * sizeof(double)) != cudaSuccess ||
            cudaMalloc(&e->d_bsel_ids, b * SORT_CHUNK * sizeof(uint64_t)) != cudaSuccess ||
            cudaMemset(e->d_bsel, 0, b * 2 * sizeof(unsigned)) != cudaSuccess) {
            cudaGetLastError();
            fcb_general_free(e);
            return fcb_fail(e, FCB_ERR_NOMEM, "fcb_load_shard: cannot allocate the batched general-k scratch");
        }
    }
    return FCB_OK;
}

struct KeyScanParams {
    const void *vectors;
    const float *inv_norm;
    const uint32_t *id_rank;
    const int32_t *dates;
    const uint8_t *cats;
    const double *query;
    fcb_filter filter;
    int64_t n_rows, row_base;
    int dim;
    uint64_t *keys;                        // [segment rows], index = row - lo
    FcbSelectState *state;
};

__device__ __forceinline__ uint64_t make_key(float score, uint32_t tie) {
    uint32_t u = __float_as_uint(score);
    u ^= (u >> 31) ? 0xffffffffu : 0x80000000u;             // monotone float -> unsigned
    return ((uint64_t)u << 32) | (uint64_t)(0xffffffffu - tie);   // larger key = better
}

template <typename T> struct KeyDot;
template <> struct KeyDot<float> {
    static constexpr int VEC = 4;
    __device__ static __forceinline__ float dot(uint4 v, const float *q, float acc) {
        acc = fmaf(__uint_as_float(v.x), q[0], acc); acc = fmaf(__uint_as_float(v.y), q[1], acc);
        acc = fmaf(__uint_as_float(v.z), q[2], acc); acc = fmaf(__uint_as_float(v.w), q[3], acc);
        return acc;
    }
};
template <> struct KeyDot<__nv_bfloat16> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ float dot(uint4 v, const float *q, float acc) {
        acc = fmaf(fcb_bf16lo(v.x), q[0], acc); acc = fmaf(fcb_bf16hi(v.x), q[1], acc);
        acc = fmaf(fcb_bf16lo(v.y), q[2], acc); acc = fmaf(fcb_bf16hi(v.y), q[3], acc);
        acc = fmaf(fcb_bf16lo(v.z), q[4], acc); acc = fmaf(fcb_bf16hi(v.z), q[5], acc);
        acc = fmaf(fcb_bf16lo(v.w), q[6], acc); acc = fmaf(fcb_bf16hi(v.w), q[7], acc);
        return acc;
    }
};

// R rows per warp step x U-fold unroll = R * U loads in flight per lane (12 streams fastest, see fcb_search.cu)
template <typename T, int R, int U>
__global__ void __launch_bounds__(FCB_SCAN_THREADS, 2) scan_keys_kernel(const __grid_constant__ KeyScanParams p) {
    constexpr int VEC = KeyDot<T>::VEC;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *q_s = reinterpret_cast<float *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const fcb_filter f = p.filter;
    for (int i = tid; i < p.dim; i += FCB_SCAN_THREADS) q_s[i] = (float)p.query[i];
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
    unsigned n_pass = 0;

    const int64_t stride = (int64_t)gridDim.x * FCB_SCAN_WARPS;
    for (int64_t g = (int64_t)blockIdx.x * FCB_SCAN_WARPS + warp; g < ngroups; g += stride) {
        const int64_t r0 = lo + g * R;
        bool pass = false, in_seg = false;
        float inv = 0.f;
        uint32_t tie = 0;
        if (lane < R && r0 + lane < hi) {
            const int64_t row = r0 + lane;
            in_seg = pass = true;
            if (has_date) { const int64_t d = p.dates[row]; pass = d >= f.date_lo && d <= f.date_hi; }
            if (pass && has_cat) { const unsigned c = p.cats[row]; pass = c < 64 && ((f.category_mask >> c) & 1ull); }
            if (pass) { inv = p.inv_norm[row]; tie = p.id_rank ? p.id_rank[row] : (uint32_t)row; }
        }
        const unsigned pm = __ballot_sync(FCB_FULL, pass);
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;
        if (pm) {
            const uint4 *__restrict__ base = reinterpret_cast<const uint4 *>(vec + r0 * (int64_t)p.dim);
#pragma unroll U
            for (int c = lane; c < nv; c += 32) {
                float qv[VEC];
#pragma unroll
                for (int i = 0; i < VEC; i += 4)
                    *reinterpret_cast<float4 *>(&qv[i]) = *reinterpret_cast<const float4 *>(&q_s[c * VEC + i]);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const uint4 v = fcb_ldg_stream(base + (int64_t)r * nv + c, (pm >> r) & 1u);
                    acc[r] = KeyDot<T>::dot(v, qv, acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = fcb_warp_sum(acc[r]);
        }
        // lane r (< R) owns row r0 + r: pick its sum without dynamic register indexing
        float mine = acc[0];
#pragma unroll
        for (int r = 1; r < R; ++r) mine = (lane == r) ? acc[r] : mine;
        if (in_seg) p.keys[r0 + lane - lo] = pass ? make_key(mine * inv, tie) : 0ull;
        n_pass += __popc(pm);
    }
    if (lane == 0 && n_pass) atomicAdd(&p.state->n_pass, n_pass);
}

// (prefix, remaining) after `passes` completed radix passes; all threads of the CTA call this.
// K' = min(kprime, n_pass). remaining == 0 means "nothing to select". `settled` is set once the score
// bits (passes 0..3) are resolved AND every key of the final bucket is wanted (remaining == bucket size,
// i.e. no tie straddles the cut): the tie-order passes 4..7 then have nothing to decide and are skipped.
__device__ void resolve_state(const FcbSelectState *st, int passes, unsigned kprime, uint64_t &prefix,
                              unsigned &remaining, bool &settled, unsigned *s_scan /* [256] */, unsigned *s_out /* [3] */) {
    const int tid = threadIdx.x;
    unsigned rem = st->n_pass < kprime ? st->n_pass : kprime;
    uint64_t pre = 0;
    bool done = false;
    for (int q = 0; q < passes && rem > 0 && !done; ++q) {
        // suffix-inclusive sums over digits 255..0 (Hillis-Steele on reversed index)
        const unsigned h = tid < 256 ? st->hist[q][255 - tid] : 0;       // position tid <-> digit 255 - tid
        if (tid < 256) s_scan[tid] = h;
        __syncthreads();
        for (int off = 1; off < 256; off <<= 1) {
            unsigned add = 0;
            if (tid < 256 && tid >= off) add = s_scan[tid - off];
            __syncthreads();
            if (tid < 256) s_scan[tid] += add;
            __syncthreads();
        }
        if (tid < 256) {
            const unsigned incl = s_scan[tid], excl = incl - h;           // keys with a larger digit: excl
            if (excl < rem && rem <= incl) { s_out[0] = 255 - tid; s_out[1] = rem - excl; s_out[2] = h; }
        }
        __syncthreads();
        pre |= (uint64_t)s_out[0] << (56 - 8 * q);
        rem = s_out[1];
        done = q >= 3 && rem == s_out[2];
        __syncthreads();
    }
    prefix = pre;
    remaining = rem;
    settled = done;
}

__global__ void __launch_bounds__(SEL_THREADS) radix_hist_kernel(const uint64_t *__restrict__ keys, int64_t n,
                                                                 FcbSelectState *st, int pass, unsigned kprime) {
    __shared__ unsigned s_scan[256], s_out[3], s_hist[256];
    uint64_t prefix;
    unsigned remaining;
    bool settled;
    resolve_state(st, pass, kprime, prefix, remaining, settled, s_scan, s_out);
    if (remaining == 0 || settled) return;
    const int tid = threadIdx.x;
    s_hist[tid] = 0;
    __syncthreads();
    const int shift = 56 - 8 * pass;
    const uint64_t himask = pass == 0 ? 0ull : ~0ull << (shift + 8);
    // neighbouring keys usually share the digit (pass 0: sign + exponent bits of similar scores), so each
    // thread counts runs locally and issues one shared atomic per run instead of one per key
    unsigned cur = 0xffffffffu, run = 0;
    for (int64_t i = (int64_t)blockIdx.x * SEL_THREADS + tid; i < n; i += (int64_t)gridDim.x * SEL_THREADS) {
        const uint64_t key = keys[i];
        if (key != 0ull && (key & himask) == prefix) {
            const unsigned d = (unsigned)(key >> shift) & 255u;
            if (d == cur) { ++run; }
            else { if (run) atomicAdd(&s_hist[cur], run); cur = d; run = 1; }
        }
    }
    if (run) atomicAdd(&s_hist[cur], run);
    __syncthreads();
    if (s_hist[tid]) atomicAdd(&st->hist[pass][tid], s_hist[tid]);
}

__global__ void __launch_bounds__(SEL_THREADS) compact_kernel(const uint64_t *__restrict__ keys, int64_t n, int64_t lo,
                                                              FcbSelectState *st, unsigned kprime,
                                                              uint32_t *__restrict__ sel_rows) {
    __shared__ unsigned s_scan[256], s_out[3];
    uint64_t thr;
    unsigned remaining;
    bool settled;
    resolve_state(st, 8, kprime, thr, remaining, settled, s_scan, s_out);   // settled: thr has zero tie bits = the whole bucket
    if (remaining == 0) return;                                    // nothing passed the filter
    for (int64_t i = (int64_t)blockIdx.x * SEL_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * SEL_THREADS) {
        const uint64_t key = keys[i];
        if (key != 0ull && key >= thr) {
            const unsigned slot = atomicAdd(&st->n_selected, 1u);
            if (slot < kprime) sel_rows[slot] = (uint32_t)(lo + i);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(SEL_THREADS) rescore_kernel(const T *__restrict__ vectors, const uint64_t *__restrict__ ids,
                                                              int64_t row_base, int dim, const double *__restrict__ q,
                                                              const FcbSelectState *st, unsigned kprime,
                                                              const uint32_t *__restrict__ sel_rows,
                                                              double *__restrict__ sel_scores, uint64_t *__restrict__ sel_ids) {
    const unsigned n_sel = st->n_pass < kprime ? st->n_pass : kprime;
    const int lane = threadIdx.x & 31;
    const unsigned w = blockIdx.x * (SEL_THREADS / 32) + (threadIdx.x >> 5);
    if (w >= n_sel) return;
    const uint32_t row = sel_rows[w];
    const double sc = fcb_canon_score_warp<T>(vectors + (int64_t)row * dim, q, dim, lane);
    if (lane == 0) { sel_scores[w] = sc; sel_ids[w] = ids ? ids[row] : (uint64_t)(row_base + row); }
}

// bitonic sort of `width` (a power of two) entries in shared memory by (score desc, id asc); all threads of the CTA call it
__device__ __forceinline__ void sel_sort_chunk(double *s_s, uint64_t *s_i, unsigned width) {
    for (unsigned size = 2; size <= width; size <<= 1) {
        for (unsigned stride = size >> 1; stride > 0; stride >>= 1) {
            for (unsigned t = threadIdx.x; t < width / 2; t += blockDim.x) {
                const unsigned pos = 2 * t - (t & (stride - 1)), mate = pos + stride;
                const bool first_better = (pos & size) == 0;      // this block is ordered best-first, the next one worst-first
                const double sa = s_s[pos], sb = s_s[mate];
                const uint64_t ia = s_i[pos], ib = s_i[mate];
                if (first_better ? fcb_better64(sb, ib, sa, ia) : fcb_better64(sa, ia, sb, ib)) {
                    s_s[pos] = sb; s_i[pos] = ib; s_s[mate] = sa; s_i[mate] = ia;
                }
            }
            __syncthreads();
        }
    }
}

// load up to 1024 candidates starting at `base`, sort them; returns how many are real (padding sorts last)
__device__ __forceinline__ unsigned sel_load_sorted(double *s_s, uint64_t *s_i, const double *sel_scores, const uint64_t *sel_ids,
                                                    unsigned base, unsigned n_sel) {
    const unsigned here = n_sel > base ? (n_sel - base < SORT_CHUNK ? n_sel - base : SORT_CHUNK) : 0;
    unsigned width = 2;                                            // power of two covering the chunk's entries
    while (width < here) width <<= 1;
    for (unsigned t = threadIdx.x; t < width; t += blockDim.x) {
        const bool in = t < here;
        s_s[t] = in ? __ldcg(sel_scores + base + t) : -INFINITY;      // written by other CTAs of this launch (direct path)
        s_i[t] = in ? __ldcg(sel_ids + base + t) : UINT64_MAX;
    }
    __syncthreads();
    sel_sort_chunk(s_s, s_i, width);
    return here;
}

// the first k of a sorted chunk -> outputs, tail padded, count set (the query's only chunk)
__device__ __forceinline__ void sel_write_single(const double *s_s, const uint64_t *s_i, unsigned n_sel, int k, uint64_t *out_ids,
                                                 double *out_scores, int32_t *out_count) {
    const unsigned count = n_sel < (unsigned)k ? n_sel : (unsigned)k;
    if (threadIdx.x == 0) *out_count = (int32_t)count;
    for (unsigned t = threadIdx.x; t < (unsigned)k; t += blockDim.x) {
        out_ids[t] = t < count ? s_i[t] : UINT64_MAX;
        out_scores[t] = t < count ? s_s[t] : -INFINITY;
    }
}

// A query's only chunk (<= 1024 candidates starting at sel_*[0]) -> results. Up to SEL_RANK_MAX candidates no sort at all:
// each warp takes candidates w, w + warps, ... and counts, 32 at a time, how many others rank before it — that count IS its
// output position (ids are unique, the order is strict). A bitonic step costs ~0.25 us of barrier + shared-memory latency
// whatever the width (28 steps for 128 entries: 7 us, more than scoring them); the counting pass of a typical tenant takes
// well under a microsecond. Above SEL_RANK_MAX the bitonic sort wins (the counting is quadratic).
#define SEL_RANK_MAX 384
__device__ __forceinline__ void sel_finish_single(double *s_s, uint64_t *s_i, const double *sel_scores, const uint64_t *sel_ids,
                                                  unsigned n_sel, int k, uint64_t *out_ids, double *out_scores, int32_t *out_count) {
    if (n_sel > SEL_RANK_MAX) {
        sel_load_sorted(s_s, s_i, sel_scores, sel_ids, 0, n_sel);
        sel_write_single(s_s, s_i, n_sel, k, out_ids, out_scores, out_count);
        return;
    }
    const unsigned count = n_sel < (unsigned)k ? n_sel : (unsigned)k;
    for (unsigned t = threadIdx.x; t < n_sel; t += blockDim.x) { s_s[t] = __ldcg(sel_scores + t); s_i[t] = __ldcg(sel_ids + t); }
    if (threadIdx.x == 0) *out_count = (int32_t)count;
    for (unsigned t = count + threadIdx.x; t < (unsigned)k; t += blockDim.x) { out_ids[t] = UINT64_MAX; out_scores[t] = -INFINITY; }
    __syncthreads();
    const unsigned lane = threadIdx.x & 31, warps = blockDim.x >> 5;
    for (unsigned e = threadIdx.x >> 5; e < n_sel; e += warps) {
        const double ms = s_s[e];
        const uint64_t mi = s_i[e];
        unsigned before = 0;
        for (unsigned j = lane; j < n_sel; j += 32) before += fcb_better64(s_s[j], s_i[j], ms, mi);
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) before += __shfl_xor_sync(FCB_FULL, before, off);
        if (lane == 0 && before < (unsigned)k) { out_ids[before] = mi; out_scores[before] = ms; }
    }
}

// Sorts candidates [1024 * blockIdx.x, +1024) by (score desc, id asc) in shared memory. `single` (the host knows
// that at most one chunk can exist): writes the first k straight to the outputs, pads the tail and sets the count.
__global__ void __launch_bounds__(SORT_CHUNK / 2) sort_chunks_kernel(const FcbSelectState *st, unsigned cap, int k, int single,
                                                                     double *__restrict__ sel_scores, uint64_t *__restrict__ sel_ids,
                                                                     uint64_t *__restrict__ out_ids, double *__restrict__ out_scores,
                                                                     int32_t *__restrict__ out_count) {
    __shared__ double s_s[SORT_CHUNK];
    __shared__ uint64_t s_i[SORT_CHUNK];
    const unsigned n_sel = st->n_pass < cap ? st->n_pass : cap;
    const unsigned base = blockIdx.x * SORT_CHUNK;
    if (!single && base >= n_sel) return;
    if (single) { sel_finish_single(s_s, s_i, sel_scores, sel_ids, n_sel, k, out_ids, out_scores, out_count); return; }
    const unsigned here = sel_load_sorted(s_s, s_i, sel_scores, sel_ids, base, n_sel);
    {
        for (unsigned t = threadIdx.x; t < here; t += blockDim.x) { sel_scores[base + t] = s_s[t]; sel_ids[base + t] = s_i[t]; }
    }
}

// Several sorted chunks: an entry's rank is its position in its own chunk plus, for every other chunk, the number of
// entries there that rank before it (binary search; ids are unique, so the order is strict). Writes the first k.
__global__ void __launch_bounds__(256) rank_merge_kernel(const FcbSelectState *st, unsigned cap, int k,
                                                         const double *__restrict__ sel_scores, const uint64_t *__restrict__ sel_ids,
                                                         uint64_t *__restrict__ out_ids, double *__restrict__ out_scores,
                                                         int32_t *__restrict__ out_count) {
    const unsigned n_sel = st->n_pass < cap ? st->n_pass : cap;
    const unsigned count = n_sel < (unsigned)k ? n_sel : (unsigned)k;
    const unsigned i = blockIdx.x * 256 + threadIdx.x;
    if (i == 0) *out_count = (int32_t)count;
    if (i >= count && i < (unsigned)k) { out_ids[i] = UINT64_MAX; out_scores[i] = -INFINITY; }   // pad the tail
    if (i >= n_sel) return;
    const double ms = sel_scores[i];
    const uint64_t mi = sel_ids[i];
    const unsigned own = i / SORT_CHUNK;
    unsigned rank = i % SORT_CHUNK;
    for (unsigned c = 0, lo0 = 0; lo0 < n_sel; ++c, lo0 += SORT_CHUNK) {
        if (c == own) continue;
        unsigned lo = lo0, hi = lo0 + SORT_CHUNK < n_sel ? lo0 + SORT_CHUNK : n_sel;
        while (lo < hi) {                                          // first entry of the chunk that does NOT rank before mine
            const unsigned mid = (lo + hi) >> 1;
            if (fcb_better64(sel_scores[mid], sel_ids[mid], ms, mi)) lo = mid + 1; else hi = mid;
        }
        rank += lo - lo0;
    }
    if (rank < (unsigned)k) { out_ids[rank] = mi; out_scores[rank] = ms; }
}

// Certificate of the select path (after the results are written): the K' selected rows are the K' best by float32 key,
// so a row that was NOT selected has a first-pass score <= the selection threshold's; its canonical score is therefore at
// most thr / |q| + eps (fcb_first_pass_eps). If the exact k-th score clears that, no dropped row belongs to the top k.
// Otherwise the flag is set and the exact float64 path (fcb_exact.cu), enqueued right behind, answers the query again.
__global__ void __launch_bounds__(SEL_THREADS) general_cert_kernel(const FcbSelectState *st, unsigned kprime, const double *__restrict__ q,
                                                                   int dim, double eps, const double *__restrict__ out_scores,
                                                                   const int32_t *__restrict__ out_count, int32_t *__restrict__ flag) {
    __shared__ unsigned s_scan[256], s_out[3];
    __shared__ double s_qq;
    uint64_t thr;
    unsigned remaining;
    bool settled;
    resolve_state(st, 8, kprime, thr, remaining, settled, s_scan, s_out);
    if (threadIdx.x < 32) {
        const double qq = fcb_canon_qq_warp(q, dim, threadIdx.x);
        if (threadIdx.x == 0) s_qq = qq;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    int f = 0;
    if (st->n_pass > kprime) {                                     // rows were dropped
        uint32_t u = (uint32_t)(thr >> 32);                        // float32 score of the worst selected row (make_key's inverse)
        u ^= (u >> 31) ? 0x80000000u : 0xffffffffu;
        const float worst = __uint_as_float(u);
        const double nq = __dsqrt_rn(s_qq);
        const int cnt = *out_count;
        const double kth = cnt > 0 ? out_scores[cnt - 1] : -INFINITY;
        const bool domain = nq > 8.9e-16 && nq < 1.1e15 && isfinite(worst);
        f = !(domain && kth > (double)worst / nq + eps);
    }
    *flag = f;
}

// A segment that fits the candidate buffer whole: predicate + canonical float64 score for every row, one warp each.
// `fused` (at most 1024 rows): the CTA that finishes last also sorts the candidates, writes the results and leaves the
// selection state zeroed — the whole query is this one launch.
template <typename T>
__global__ void __launch_bounds__(SEL_THREADS) direct_rescore_kernel(const T *__restrict__ vectors, const uint64_t *__restrict__ ids,
                                                                     const int32_t *__restrict__ dates, const uint8_t *__restrict__ cats,
                                                                     const fcb_filter f, int64_t row_base, int64_t lo, int64_t hi, int dim,
                                                                     const double *__restrict__ q, FcbSelectState *st,
                                                                     double *__restrict__ sel_scores, uint64_t *__restrict__ sel_ids,
                                                                     int fused, int k, uint64_t *__restrict__ out_ids,
                                                                     double *__restrict__ out_scores, int32_t *__restrict__ out_count) {
    const int lane = threadIdx.x & 31;
    const bool has_date = dates && (f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX);
    const bool has_cat = cats && f.category_mask != FCB_ALL_CATEGORIES;
    const int64_t warps = (int64_t)gridDim.x * (SEL_THREADS / 32);
    for (int64_t row = lo + (int64_t)blockIdx.x * (SEL_THREADS / 32) + (threadIdx.x >> 5); row < hi; row += warps) {
        bool pass = true;
        if (has_date) { const int64_t d = dates[row]; pass = d >= f.date_lo && d <= f.date_hi; }
        if (pass && has_cat) { const unsigned c = cats[row]; pass = c < 64 && ((f.category_mask >> c) & 1ull); }
        if (!pass) continue;
        const double sc = fcb_canon_score_warp<T>(vectors + row * dim, q, dim, lane);
        if (lane == 0) {
            const unsigned slot = atomicAdd(&st->n_pass, 1u);
            sel_scores[slot] = sc;
            sel_ids[slot] = ids ? ids[row] : (uint64_t)(row_base + row);
        }
    }
    if (!fused) return;
    __shared__ double s_s[SORT_CHUNK];
    __shared__ uint64_t s_i[SORT_CHUNK];
    __shared__ unsigned s_last;
    __threadfence();                                               // this CTA's candidates before its ticket
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&st->n_selected, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const unsigned n_sel = *reinterpret_cast<volatile unsigned *>(&st->n_pass);
    sel_finish_single(s_s, s_i, sel_scores, sel_ids, n_sel, k, out_ids, out_scores, out_count);
    if (threadIdx.x == 0) { st->n_pass = 0; st->n_selected = 0; }  // the state stays clean for the next query
}

// The same for a whole batch of small segments: blockIdx.y = query, every query has its own counters and candidate
// buffer; the last CTA of each query sorts and writes that query's results and re-zeroes its counters.
template <typename T>
__global__ void __launch_bounds__(SEL_THREADS) direct_batch_kernel(const T *__restrict__ vectors, const uint64_t *__restrict__ ids,
                                                                   const int32_t *__restrict__ dates, const uint8_t *__restrict__ cats,
                                                                   const fcb_filter *__restrict__ filters, int64_t row_base, int64_t n_rows,
                                                                   int dim, const double *__restrict__ queries, unsigned *__restrict__ bsel,
                                                                   double *__restrict__ bsel_scores, uint64_t *__restrict__ bsel_ids, int k,
                                                                   uint64_t *__restrict__ out_ids, double *__restrict__ out_scores,
                                                                   int32_t *__restrict__ out_counts) {
    __shared__ double s_s[SORT_CHUNK];
    __shared__ uint64_t s_i[SORT_CHUNK];
    __shared__ unsigned s_last;
    const int b = blockIdx.y, lane = threadIdx.x & 31;
    const fcb_filter f = filters[b];
    int64_t lo = f.row_begin - row_base, hi = f.row_end - row_base;
    if (lo < 0) lo = 0;
    if (hi > n_rows) hi = n_rows;
    const bool has_date = dates && (f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX);
    const bool has_cat = cats && f.category_mask != FCB_ALL_CATEGORIES;
    unsigned *my = bsel + 2 * b;
    double *my_scores = bsel_scores + (size_t)b * SORT_CHUNK;
    uint64_t *my_ids = bsel_ids + (size_t)b * SORT_CHUNK;
    const double *q = queries + (size_t)b * dim;
    const int64_t warps = (int64_t)gridDim.x * (SEL_THREADS / 32);
    for (int64_t row = lo + (int64_t)blockIdx.x * (SEL_THREADS / 32) + (threadIdx.x >> 5); row < hi; row += warps) {
        bool pass = true;
        if (has_date) { const int64_t d = dates[row]; pass = d >= f.date_lo && d <= f.date_hi; }
        if (pass && has_cat) { const unsigned c = cats[row]; pass = c < 64 && ((f.category_mask >> c) & 1ull); }
        if (!pass) continue;
        const double sc = fcb_canon_score_warp<T>(vectors + row * dim, q, dim, lane);
        if (lane == 0) {
            const unsigned slot = atomicAdd(&my[0], 1u);
            my_scores[slot] = sc;
            my_ids[slot] = ids ? ids[row] : (uint64_t)(row_base + row);
        }
    }
    __threadfence();                                               // this CTA's candidates before its ticket
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&my[1], 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const unsigned n_sel = *reinterpret_cast<volatile unsigned *>(&my[0]);
    sel_finish_single(s_s, s_i, my_scores, my_ids, n_sel, k, out_ids + (size_t)b * k, out_scores + (size_t)b * k, out_counts + b);
    if (threadIdx.x == 0) { my[0] = 0; my[1] = 0; }
}

int32_t fcb_search_general_small_batch(fcb_engine *e, const double *d_queries, const fcb_filter *d_filters, int32_t batch,
                                       int32_t k, int64_t max_rows, uint64_t *d_out_ids, double *d_out_scores,
                                       int32_t *d_out_counts, cudaStream_t st) {
    const fcb_shard &s = e->shard;
    if (!e->d_bsel) return fcb_fail(e, FCB_ERR_STATE, "fcb_search: batched general-k scratch missing");
    int64_t per_query = (max_rows + SEL_THREADS / 32 - 1) / (SEL_THREADS / 32);
    if (per_query < 1) per_query = 1;
    const dim3 grid((unsigned)per_query, (unsigned)batch);
    if (s.dtype == FCB_DTYPE_F32)
        direct_batch_kernel<float><<<grid, SEL_THREADS, 0, st>>>((const float *)s.vectors, s.ids, s.dates, s.categories, d_filters,
                                                                 s.row_base, s.n_rows, s.dim, d_queries, e->d_bsel, e->d_bsel_scores,
                                                                 e->d_bsel_ids, k, d_out_ids, d_out_scores, d_out_counts);
    else
        direct_batch_kernel<__nv_bfloat16><<<grid, SEL_THREADS, 0, st>>>((const __nv_bfloat16 *)s.vectors, s.ids, s.dates, s.categories,
                                                                         d_filters, s.row_base, s.n_rows, s.dim, d_queries, e->d_bsel,
                                                                         e->d_bsel_scores, e->d_bsel_ids, k, d_out_ids, d_out_scores,
                                                                         d_out_counts);
    FCB_CUDA(e, cudaGetLastError());
    e->last_launches += 1;
    return FCB_OK;
}

template <typename T>
static int32_t general_impl(fcb_engine *e, const double *d_query, const fcb_filter &f, int32_t k, uint64_t *d_out_ids,
                            double *d_out_scores, int32_t *d_out_count, int32_t *d_flag, cudaStream_t st) {
    const fcb_shard &s = e->shard;
    const unsigned kprime = (unsigned)kprime_for(k);
    int64_t lo = f.row_begin - s.row_base, hi = f.row_end - s.row_base;
    if (lo < 0) lo = 0;
    if (hi > s.n_rows) hi = s.n_rows;
    const int64_t n = hi > lo ? hi - lo : 0;

    const bool direct = n > 0 && n <= (int64_t)e->kprime_max;         // the whole segment fits the candidate buffer
    const bool fused = direct && n <= SORT_CHUNK;                      // one launch does it all and leaves the state clean
    if (!(fused && e->sel_clean)) FCB_CUDA(e, cudaMemsetAsync(e->d_sel, 0, sizeof(FcbSelectState), st));
    e->sel_clean = fused;
    unsigned cap = kprime;                                             // how many candidates can exist
    if (direct) {
        // score every passing row in float64 right away
        cap = (unsigned)n;
        int64_t ctas = (n + SEL_THREADS / 32 - 1) / (SEL_THREADS / 32);
        if (ctas > 8 * (int64_t)e->num_sms) ctas = 8 * (int64_t)e->num_sms;
        direct_rescore_kernel<T><<<(unsigned)ctas, SEL_THREADS, 0, st>>>((const T *)s.vectors, s.ids, s.dates, s.categories, f,
                                                                        s.row_base, lo, hi, s.dim, d_query, e->d_sel,
                                                                        e->d_sel_scores, e->d_sel_ids, fused ? 1 : 0, k,
                                                                        d_out_ids, d_out_scores, d_out_count);
        FCB_CUDA(e, cudaGetLastError());
        e->last_launches += 1;
        if (fused) return FCB_OK;
    } else if (n > 0) {
        KeyScanParams p{};
        p.vectors = s.vectors; p.inv_norm = e->inv_norm; p.id_rank = s.id_rank; p.dates = s.dates; p.cats = s.categories;
        p.query = d_query; p.filter = f; p.n_rows = s.n_rows; p.row_base = s.row_base; p.dim = s.dim;
        p.keys = e->d_keys; p.state = e->d_sel;
        const int steps = (s.dim / KeyDot<T>::VEC + 31) / 32;
        const bool six = steps % 6 == 0 && sizeof(T) == 4, three = steps % 3 == 0;
        const int R = six ? 2 : 4;
        const int64_t groups = (n + R - 1) / R;
        int64_t ctas = (groups + 2 * FCB_SCAN_WARPS - 1) / (2 * FCB_SCAN_WARPS);
        if (ctas > 2 * (int64_t)e->num_sms) ctas = 2 * (int64_t)e->num_sms;
        if (ctas < 1) ctas = 1;
        const size_t smem = (size_t)s.dim * sizeof(float);
        if (six) scan_keys_kernel<T, 2, 6><<<(unsigned)ctas, FCB_SCAN_THREADS, smem, st>>>(p);
        else if (three) scan_keys_kernel<T, 4, 3><<<(unsigned)ctas, FCB_SCAN_THREADS, smem, st>>>(p);
        else scan_keys_kernel<T, 4, 2><<<(unsigned)ctas, FCB_SCAN_THREADS, smem, st>>>(p);
        FCB_CUDA(e, cudaGetLastError());
        int64_t sel_ctas = (n + SEL_THREADS * 16 - 1) / (SEL_THREADS * 16);
        if (sel_ctas > 4 * (int64_t)e->num_sms) sel_ctas = 4 * (int64_t)e->num_sms;
        for (int pass = 0; pass < 8; ++pass)
            radix_hist_kernel<<<(unsigned)sel_ctas, SEL_THREADS, 0, st>>>(e->d_keys, n, e->d_sel, pass, kprime);
        compact_kernel<<<(unsigned)sel_ctas, SEL_THREADS, 0, st>>>(e->d_keys, n, lo, e->d_sel, kprime, e->d_sel_rows);
        rescore_kernel<T><<<(kprime + SEL_THREADS / 32 - 1) / (SEL_THREADS / 32), SEL_THREADS, 0, st>>>(
            (const T *)s.vectors, s.ids, s.row_base, s.dim, d_query, e->d_sel, kprime, e->d_sel_rows, e->d_sel_scores,
            e->d_sel_ids);
        FCB_CUDA(e, cudaGetLastError());
        e->last_launches += 11;
    }
    // order the candidates: one sorted chunk ends the query, several are merged by rank
    const bool single = cap <= SORT_CHUNK;
    sort_chunks_kernel<<<single ? 1 : (cap + SORT_CHUNK - 1) / SORT_CHUNK, SORT_CHUNK / 2, 0, st>>>(
        e->d_sel, cap, k, single ? 1 : 0, e->d_sel_scores, e->d_sel_ids, d_out_ids, d_out_scores, d_out_count);
    e->last_launches += 1;
    if (!single) {
        const unsigned span = cap > (unsigned)k ? cap : (unsigned)k;
        rank_merge_kernel<<<(span + 255) / 256, 256, 0, st>>>(e->d_sel, cap, k, e->d_sel_scores, e->d_sel_ids, d_out_ids,
                                                            d_out_scores, d_out_count);
        e->last_launches += 1;
    }
    FCB_CUDA(e, cudaGetLastError());
    if (!direct && n > 0) {
        // the select path ranked rows by float32 keys: certify, and let the exact path answer what cannot be certified
        general_cert_kernel<<<1, SEL_THREADS, 0, st>>>(e->d_sel, kprime, d_query, s.dim, fcb_first_pass_eps(s.dim), d_out_scores,
                                                       d_out_count, d_flag);
        FCB_CUDA(e, cudaGetLastError());
        e->last_launches += 1;
        return fcb_exact_enqueue(e, d_query, nullptr, &f, 1, 1, k, d_flag, d_out_ids, d_out_scores, d_out_count, st);
    }
    return FCB_OK;
}

int32_t fcb_search_general(fcb_engine *e, const double *d_query, const fcb_filter &f, int32_t k, uint64_t *d_out_ids,
                           double *d_out_scores, int32_t *d_out_count, int32_t *d_flag, cudaStream_t st) {
    if (!e->d_keys) return fcb_fail(e, FCB_ERR_STATE, "fcb_search: general-k scratch missing (max_k <= %d at fcb_open?)", FCB_FUSED_MAX_K);
    if (e->shard.dtype == FCB_DTYPE_F32)
        return general_impl<float>(e, d_query, f, k, d_out_ids, d_out_scores, d_out_count, d_flag, st);
    return general_impl<__nv_bfloat16>(e, d_query, f, k, d_out_ids, d_out_scores, d_out_count, d_flag, st);
}
