// Exact float64 fallback of fcb_search: what makes "bit-exact ids" hold by construction on every path.
//
// The fast paths rank rows by a float32 first pass (fused scan, general-k select) or by a reduced-precision tensor-core
// pass (batched path), keep a bounded number of candidates and re-score those canonically in float64. Each of them ends
// with a CERTIFICATE: the exact k-th score must clear the best score any dropped row could have, first-pass error bound
// included (fcb_first_pass_eps, fcb_topk.cuh). A query that cannot be certified — more rows than the candidate margin
// within rounding distance of the k-th score: duplicated transactions, degenerate embeddings, a query far outside the
// float32 range — sets a device flag, and THIS kernel, enqueued right behind on the same stream, answers it from
// scratch without any float32 step:
//   1. canonical float64 score of every passing row of the segment -> a sortable 64-bit key per row (8 B/row scratch),
//   2. MSB-first radix select of the k-th largest key (8 passes of 8 bits; duplicates allowed),
//   3. when a tie straddles the cut: radix select of the wanted number of SMALLEST tie keys (row, or id rank) among the
//      rows whose key equals the k-th — so ties at the k-th score resolve by ascending id, like everywhere else,
//   4. compaction of exactly min(k, passing rows) rows, ranking by (score desc, id asc), results written.
// It is ONE ordinary launch of two CTAs per SM that leaves at once when no flag is set — the cost on the common path is that
// empty launch. Only a flagged query makes the CTAs meet at grid-wide barriers between the phases (a ticket counter in
// global memory; two small CTAs per SM are co-resident on an otherwise idle GPU and squeeze in beside anything that does not
// fill an SM; the waits are BOUNDED — a grid that cannot assemble gives up, raises the error word and fcb_exact_reruns
// reports it — so nothing can hang). A cooperative launch would give the co-residency guarantee for free, but the driver
// runs a cooperative grid only on an idle device: every search would then serialise against whatever else the GPU is doing.
// No host round trip, no CPU path.
// The keys ARE the canonical scores (order-preserving bit transform), so the scores written are bit-identical to a
// re-score. Replaces nothing in the reference: qdrant's local mode is exact by being brute force in float64.
#include "fcb_internal.cuh"
#include "fcb_topk.cuh"

#define EX_THREADS 256
#define EX_WARPS (EX_THREADS / 32)

#define EX_BARRIER_SPINS (1u << 26)        // x ~0.1 us of back-off: a grid that cannot assemble gives up after several seconds

struct FcbExactState {
    unsigned hist[8][256];                 // digit histograms of the score-key passes
    unsigned thist[4][256];                // digit histograms of the tie-key passes
    unsigned n_pass;                       // rows that passed the filter
    unsigned n_sel;                        // compaction cursor
};

struct FcbExactBarrier {                   // survives the per-query clearing of the state
    unsigned count;                        // CTAs that have arrived at the current barrier
    unsigned generation;                   // bumped by the last arrival
    unsigned error;                        // 1: a barrier timed out (the answers of that launch are not valid)
};

// Grid-wide barrier for a grid whose CTAs are all resident (see the header). Bounded: on a timeout the error word is raised
// and every later wait returns at once.
__device__ __forceinline__ void ex_grid_barrier(FcbExactBarrier *bar) {
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned *gen = &bar->generation;
        volatile unsigned *err = &bar->error;
        const unsigned g = *gen;
        __threadfence();
        if (atomicAdd(&bar->count, 1u) == gridDim.x - 1) {
            bar->count = 0;
            __threadfence();
            atomicAdd(&bar->generation, 1u);
        } else {
            unsigned spins = 0;
            while (*gen == g && *err == 0) {
                if (++spins > EX_BARRIER_SPINS) { *err = 1; break; }
                __nanosleep(100);
            }
        }
        __threadfence();
    }
    __syncthreads();
}

struct FcbExact {                          // engine-owned scratch (fcb_load_shard)
    uint64_t *keys = nullptr;              // [n_rows]; the general-k path's key buffer when that exists
    bool own_keys = false;
    FcbExactState *state = nullptr;
    FcbExactBarrier *barrier = nullptr;
    uint64_t *sel_keys = nullptr;          // [max_k]
    uint32_t *sel_rows = nullptr, *sel_ties = nullptr;
    int32_t *flags = nullptr;              // [max_batch] certificate flags of device-API calls
    unsigned long long *reruns = nullptr;  // queries answered by this path since fcb_load_shard
    int grid_f32 = 0, grid_bf16 = 0;
};

struct ExactParams {
    const void *vectors;
    const uint64_t *ids;
    const uint32_t *id_rank;
    const int32_t *dates;
    const uint8_t *cats;
    const double *queries;
    const fcb_filter *filters;             // device copy, used when !use_inline
    fcb_filter inl[FCB_INLINE_FILTERS];
    int use_inline;                        // 1: inl[b]; 2: inl[0] for every query (batched path: one shared filter)
    int64_t n_rows, row_base;
    int dim, k, batch;
    const int32_t *flags;
    const int32_t *list, *list_count;      // optional (routed search): only these queries of the batch ran, look at their flags only
    uint64_t *keys;
    FcbExactState *st;
    FcbExactBarrier *bar;
    uint64_t *sel_keys;
    uint32_t *sel_rows, *sel_ties;
    uint64_t *out_ids;
    double *out_scores;
    int32_t *out_counts;
    unsigned long long *reruns;
};

__device__ __forceinline__ uint64_t ex_key(double score) {      // monotone double -> unsigned, never 0
    uint64_t u = (uint64_t)__double_as_longlong(score);
    u ^= (u >> 63) ? ~0ull : 0x8000000000000000ull;
    return u ? u : 1ull;
}
__device__ __forceinline__ double ex_unkey(uint64_t u) {
    u ^= (u >> 63) ? 0x8000000000000000ull : ~0ull;
    return __longlong_as_double((long long)u);
}

// Which digit holds the `rem`-th entry counted from the top (descending = true) or from the bottom of a 256-bin
// histogram: returns the digit, how many entries are still wanted inside it, and its size. All threads call it.
__device__ void ex_pick_digit(const unsigned *hist, unsigned rem, bool descending, unsigned *s_scan, unsigned *s_out,
                              unsigned &digit, unsigned &rem_in, unsigned &bucket) {
    const int tid = threadIdx.x;
    const unsigned h = __ldcg(hist + (descending ? 255 - tid : tid));      // position tid <-> digit
    s_scan[tid] = h;
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {
        const unsigned add = tid >= off ? s_scan[tid - off] : 0;
        __syncthreads();
        s_scan[tid] += add;
        __syncthreads();
    }
    const unsigned incl = s_scan[tid], excl = incl - h;
    if (excl < rem && rem <= incl) { s_out[0] = descending ? 255 - tid : tid; s_out[1] = rem - excl; s_out[2] = h; }
    __syncthreads();
    digit = s_out[0]; rem_in = s_out[1]; bucket = s_out[2];
    __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(EX_THREADS) exact_fallback_kernel(const __grid_constant__ ExactParams p) {
    extern __shared__ __align__(16) unsigned char ex_smem[];
    double *q_s = reinterpret_cast<double *>(ex_smem);                     // [dim] float64 query
    __shared__ unsigned s_hist[256], s_scan[256], s_out[3];
    __shared__ int s_any;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t gtid = (int64_t)blockIdx.x * EX_THREADS + tid, gthreads = (int64_t)gridDim.x * EX_THREADS;
    const T *__restrict__ vec = reinterpret_cast<const T *>(p.vectors);

    // the common case: nothing to do. Every CTA reads the same flags, so the decision is grid-uniform.
    const int n_queries = p.list ? __ldcg(p.list_count) : p.batch;
    if (tid == 0) s_any = 0;
    __syncthreads();
    for (int i = tid; i < n_queries; i += EX_THREADS)
        if (__ldcg(p.flags + (p.list ? __ldcg(p.list + i) : i)) != 0) s_any = 1;
    __syncthreads();
    if (!s_any) return;

    for (int i = 0; i < n_queries; ++i) {
        const int b = p.list ? __ldcg(p.list + i) : i;
        if (__ldcg(p.flags + b) == 0) continue;                            // grid-uniform
        const fcb_filter f = p.use_inline == 2 ? p.inl[0] : (p.use_inline ? p.inl[b] : p.filters[b]);
        const double *__restrict__ qd = p.queries + (size_t)b * p.dim;
        int64_t lo = f.row_begin - p.row_base, hi = f.row_end - p.row_base;
        if (lo < 0) lo = 0;
        if (hi > p.n_rows) hi = p.n_rows;
        const int64_t n = hi > lo ? hi - lo : 0;
        const bool has_date = f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX;
        const bool has_cat = f.category_mask != FCB_ALL_CATEGORIES;
        uint64_t *out_ids = p.out_ids + (size_t)b * p.k;
        double *out_scores = p.out_scores + (size_t)b * p.k;

        // ---- phase 0: clean state
        for (int64_t i = gtid; i < (int64_t)(sizeof(FcbExactState) / sizeof(unsigned)); i += gthreads)
            reinterpret_cast<unsigned *>(p.st)[i] = 0;
        for (int i = tid; i < p.dim; i += EX_THREADS) q_s[i] = qd[i];
        s_hist[tid] = 0;
        __syncthreads();
        ex_grid_barrier(p.bar);

        // ---- phase 1: canonical float64 score of every passing row -> key; histogram of the top byte
        {
            const double qq = fcb_canon_qq_warp(q_s, p.dim, lane);
            unsigned n_pass = 0;
            const int64_t ngroups = (n + 3) / 4;
            for (int64_t g = (int64_t)blockIdx.x * EX_WARPS + warp; g < ngroups; g += (int64_t)gridDim.x * EX_WARPS) {
                const int64_t r0 = lo + 4 * g;
                bool pass = false;
                if (lane < 4 && r0 + lane < hi) {
                    const int64_t row = r0 + lane;
                    pass = true;
                    if (has_date) { const int64_t d = p.dates[row]; pass = d >= f.date_lo && d <= f.date_hi; }
                    if (pass && has_cat) { const unsigned c = p.cats[row]; pass = c < 64 && ((f.category_mask >> c) & 1ull); }
                }
                const unsigned pm = __ballot_sync(FCB_FULL, pass);
                double sc4[4] = {0.0, 0.0, 0.0, 0.0};
                if (pm) {
                    const T *rows4[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) rows4[c] = ((pm >> c) & 1u) ? vec + (r0 + c) * (int64_t)p.dim : nullptr;
                    fcb_canon_score_warp4<T>(rows4, q_s, qq, p.dim, lane, sc4);
                }
                if (lane < 4 && r0 + lane < hi) {
                    double sc = sc4[0];
#pragma unroll
                    for (int c = 1; c < 4; ++c) if (lane == c) sc = sc4[c];
                    const uint64_t key = pass ? ex_key(sc) : 0ull;
                    p.keys[r0 + lane - lo] = key;
                    if (pass) atomicAdd(&s_hist[(unsigned)(key >> 56)], 1u);
                }
                n_pass += __popc(pm);
            }
            if (lane == 0 && n_pass) atomicAdd(&p.st->n_pass, n_pass);
            __syncthreads();
            if (s_hist[tid]) atomicAdd(&p.st->hist[0][tid], s_hist[tid]);
        }
        ex_grid_barrier(p.bar);

        const unsigned n_pass = __ldcg(&p.st->n_pass);
        const unsigned want = n_pass < (unsigned)p.k ? n_pass : (unsigned)p.k;
        if (want == 0) {                                                   // grid-uniform
            if (blockIdx.x == 0) {
                for (int t = tid; t < p.k; t += EX_THREADS) { out_ids[t] = UINT64_MAX; out_scores[t] = -INFINITY; }
                if (tid == 0) { p.out_counts[b] = 0; atomicAdd(p.reruns, 1ull); }
            }
            ex_grid_barrier(p.bar);
            continue;
        }

        // ---- phase 2: radix select of the want-th largest key
        uint64_t prefix = 0;
        unsigned rem = want, bucket = 0;
        bool whole_bucket = false;                                         // every key under the prefix is wanted
        int passes_done = 0;
        for (int pass = 0; pass < 8; ++pass) {
            unsigned digit;
            ex_pick_digit(p.st->hist[pass], rem, true, s_scan, s_out, digit, rem, bucket);
            prefix |= (uint64_t)digit << (56 - 8 * pass);
            passes_done = pass + 1;
            if (rem == bucket) { whole_bucket = true; break; }
            if (pass == 7) break;
            const int shift = 48 - 8 * pass;                               // next digit
            const uint64_t himask = ~0ull << (shift + 8);
            s_hist[tid] = 0;
            __syncthreads();
            for (int64_t i = gtid; i < n; i += gthreads) {
                const uint64_t key = p.keys[i];
                if (key != 0ull && (key & himask) == prefix) atomicAdd(&s_hist[(unsigned)(key >> shift) & 255u], 1u);
            }
            __syncthreads();
            if (s_hist[tid]) atomicAdd(&p.st->hist[pass + 1][tid], s_hist[tid]);
            ex_grid_barrier(p.bar);
        }
        // keys strictly above the cut are all wanted; of the keys equal to it (whole_bucket: sharing its prefix) `rem`
        const uint64_t cut_mask = passes_done == 8 ? ~0ull : ~0ull << (64 - 8 * passes_done);

        // ---- phase 3: a tie straddles the cut -> the `rem` smallest tie keys among the rows whose key equals it
        uint32_t tie_cut = 0xffffffffu;
        if (!whole_bucket) {                                               // grid-uniform; passes_done == 8, prefix = the key itself
            uint32_t tprefix = 0;
            unsigned trem = rem, tb = 0;
            for (int pass = 0; pass < 4; ++pass) {
                const int shift = 24 - 8 * pass;
                const uint32_t himask = pass == 0 ? 0u : ~0u << (shift + 8);
                s_hist[tid] = 0;
                __syncthreads();
                for (int64_t i = gtid; i < n; i += gthreads) {
                    if (p.keys[i] == prefix) {
                        const uint32_t t = fcb_tie_key((uint32_t)(lo + i), p.id_rank);
                        if ((t & himask) == tprefix) atomicAdd(&s_hist[(t >> shift) & 255u], 1u);
                    }
                }
                __syncthreads();
                if (s_hist[tid]) atomicAdd(&p.st->thist[pass][tid], s_hist[tid]);
                ex_grid_barrier(p.bar);
                unsigned digit;
                ex_pick_digit(p.st->thist[pass], trem, false, s_scan, s_out, digit, trem, tb);
                tprefix |= digit << shift;
            }
            tie_cut = tprefix;                                             // tie keys are unique: exactly `rem` rows have tie <= tie_cut
        }

        // ---- phase 4: compaction of exactly `want` rows
        for (int64_t i = gtid; i < n; i += gthreads) {
            const uint64_t key = p.keys[i];
            if (key == 0ull) continue;
            bool take;
            if (whole_bucket) take = (key & cut_mask) >= prefix;
            else take = key > prefix || (key == prefix && fcb_tie_key((uint32_t)(lo + i), p.id_rank) <= tie_cut);
            if (take) {
                const unsigned slot = atomicAdd(&p.st->n_sel, 1u);
                if (slot < want) {
                    p.sel_keys[slot] = key;
                    p.sel_rows[slot] = (uint32_t)(lo + i);
                    p.sel_ties[slot] = fcb_tie_key((uint32_t)(lo + i), p.id_rank);
                }
            }
        }
        ex_grid_barrier(p.bar);

        // ---- phase 5: rank by (score desc, id asc) and write
        for (int64_t e = gtid; e < (int64_t)want; e += gthreads) {
            const uint64_t mk = __ldcg(p.sel_keys + e);
            const uint32_t mt = __ldcg(p.sel_ties + e);
            unsigned rank = 0;
            for (unsigned j = 0; j < want; ++j) {
                const uint64_t ok = __ldcg(p.sel_keys + j);
                rank += ok > mk || (ok == mk && __ldcg(p.sel_ties + j) < mt);
            }
            const uint32_t row = __ldcg(p.sel_rows + e);
            out_ids[rank] = p.ids ? p.ids[row] : (uint64_t)(p.row_base + row);
            out_scores[rank] = ex_unkey(mk);
        }
        if (blockIdx.x == 0) {
            for (int t = (int)want + tid; t < p.k; t += EX_THREADS) { out_ids[t] = UINT64_MAX; out_scores[t] = -INFINITY; }
            if (tid == 0) { p.out_counts[b] = (int32_t)want; atomicAdd(p.reruns, 1ull); }
        }
        ex_grid_barrier(p.bar);                                                       // state, keys and selection are reused by the next query
    }
}

// ---- host side ----------------------------------------------------------------------------------

void fcb_exact_free(fcb_engine *e) {
    FcbExact *x = reinterpret_cast<FcbExact *>(e->exact);
    if (!x) return;
    if (x->own_keys) cudaFree(x->keys);
    cudaFree(x->barrier);
    cudaFree(x->state); cudaFree(x->sel_keys); cudaFree(x->sel_rows); cudaFree(x->sel_ties); cudaFree(x->flags); cudaFree(x->reruns);
    delete x;
    e->exact = nullptr;
}

template <typename T>
static int32_t exact_grid(fcb_engine *e, size_t smem, int *grid) {
    auto kern = exact_fallback_kernel<T>;
    if (smem > 48 * 1024) FCB_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    FCB_CUDA(e, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, EX_THREADS, smem));
    if (per_sm < 1) return fcb_fail(e, FCB_ERR_CUDA, "fcb_load_shard: the exact fallback kernel does not fit an SM");
    if (per_sm > 2) per_sm = 2;                                     // leave room on every SM: the grid must be co-resident
    *grid = per_sm * e->num_sms;
    return FCB_OK;
}

// Called by fcb_load_shard after the general-k scratch exists (its key buffer is shared: the two paths never overlap
// on one engine's stream).
int32_t fcb_exact_alloc(fcb_engine *e) {
    FcbExact *x = new FcbExact();
    e->exact = x;
    const size_t n = (size_t)e->shard.n_rows, kmax = (size_t)e->max_k, B = (size_t)e->max_batch;
    if (e->d_keys) x->keys = e->d_keys;
    else {
        if (cudaMalloc(&x->keys, n * sizeof(uint64_t)) != cudaSuccess) {
            cudaGetLastError();
            fcb_exact_free(e);
            return fcb_fail(e, FCB_ERR_NOMEM, "fcb_load_shard: cannot allocate %zu bytes of exact-path keys", n * sizeof(uint64_t));
        }
        x->own_keys = true;
    }
    if (cudaMalloc(&x->state, sizeof(FcbExactState)) != cudaSuccess || cudaMalloc(&x->barrier, sizeof(FcbExactBarrier)) != cudaSuccess ||
        cudaMalloc(&x->sel_keys, kmax * sizeof(uint64_t)) != cudaSuccess ||
        cudaMalloc(&x->sel_rows, kmax * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&x->sel_ties, kmax * sizeof(uint32_t)) != cudaSuccess ||
        cudaMalloc(&x->flags, B * sizeof(int32_t)) != cudaSuccess || cudaMalloc(&x->reruns, sizeof(unsigned long long)) != cudaSuccess) {
        cudaGetLastError();
        fcb_exact_free(e);
        return fcb_fail(e, FCB_ERR_NOMEM, "fcb_load_shard: cannot allocate the exact-path scratch");
    }
    FCB_CUDA(e, cudaMemsetAsync(x->flags, 0, B * sizeof(int32_t), e->stream));
    FCB_CUDA(e, cudaMemsetAsync(x->reruns, 0, sizeof(unsigned long long), e->stream));
    FCB_CUDA(e, cudaMemsetAsync(x->barrier, 0, sizeof(FcbExactBarrier), e->stream));
    const size_t smem = (size_t)e->shard.dim * sizeof(double);
    int32_t rc = e->shard.dtype == FCB_DTYPE_F32 ? exact_grid<float>(e, smem, &x->grid_f32)
                                                 : exact_grid<__nv_bfloat16>(e, smem, &x->grid_bf16);
    return rc;
}

int32_t *fcb_exact_flags(fcb_engine *e) {
    FcbExact *x = reinterpret_cast<FcbExact *>(e->exact);
    return x ? x->flags : nullptr;
}

// Enqueue the fallback behind a search whose kernels wrote `d_flags[0..batch)`. `filters_mode`: 0 = `d_filters` (device),
// 1 = `inl[b]` (batch <= FCB_INLINE_FILTERS), 2 = `inl[0]` shared by every query.
int32_t fcb_exact_enqueue(fcb_engine *e, const double *d_queries, const fcb_filter *d_filters, const fcb_filter *inl,
                          int filters_mode, int32_t batch, int32_t k, const int32_t *d_flags, uint64_t *d_out_ids,
                          double *d_out_scores, int32_t *d_out_counts, cudaStream_t st, const int32_t *d_list,
                          const int32_t *d_list_count) {
    FcbExact *x = reinterpret_cast<FcbExact *>(e->exact);
    if (!x) return fcb_fail(e, FCB_ERR_STATE, "fcb_search: exact-path scratch missing");
    const fcb_shard &s = e->shard;
    ExactParams p{};
    p.vectors = s.vectors; p.ids = s.ids; p.id_rank = s.id_rank; p.dates = s.dates; p.cats = s.categories;
    p.queries = d_queries; p.filters = d_filters; p.use_inline = filters_mode;
    if (filters_mode == 1) for (int b = 0; b < batch && b < FCB_INLINE_FILTERS; ++b) p.inl[b] = inl[b];
    if (filters_mode == 2) p.inl[0] = inl[0];
    p.n_rows = s.n_rows; p.row_base = s.row_base; p.dim = s.dim; p.k = k; p.batch = batch;
    p.flags = d_flags; p.list = d_list; p.list_count = d_list_count; p.keys = x->keys; p.st = x->state; p.bar = x->barrier;
    p.sel_keys = x->sel_keys; p.sel_rows = x->sel_rows; p.sel_ties = x->sel_ties;
    p.out_ids = d_out_ids; p.out_scores = d_out_scores; p.out_counts = d_out_counts; p.reruns = x->reruns;
    const size_t smem = (size_t)s.dim * sizeof(double);
    if (s.dtype == FCB_DTYPE_F32) exact_fallback_kernel<float><<<x->grid_f32, EX_THREADS, smem, st>>>(p);
    else exact_fallback_kernel<__nv_bfloat16><<<x->grid_bf16, EX_THREADS, smem, st>>>(p);
    FCB_CUDA(e, cudaGetLastError());
    e->last_launches += 1;
    return FCB_OK;
}

extern "C" int64_t fcb_exact_reruns(fcb_engine *e) {
    if (!e || !e->exact) return -1;
    FcbDeviceGuard g(e->device);
    FcbExact *x = reinterpret_cast<FcbExact *>(e->exact);
    unsigned long long h = 0;
    FcbExactBarrier hb{};
    if (!g.ok || cudaDeviceSynchronize() != cudaSuccess ||
        cudaMemcpy(&h, x->reruns, sizeof h, cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(&hb, x->barrier, sizeof hb, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    if (hb.error) {
        cudaMemset(x->barrier, 0, sizeof(FcbExactBarrier));           // usable again
        fcb_fail(e, FCB_ERR_STATE, "the exact float64 kernel could not assemble its grid (a barrier timed out): the GPU was "
                 "saturated by other work; the answers of that call are not valid");
        return -2;
    }
    return (int64_t)h;
}
