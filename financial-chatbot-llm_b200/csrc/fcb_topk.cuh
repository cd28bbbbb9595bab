// Warp-level sorted candidate lists and the rank-and-write tail shared by the fused scan (fcb_search.cu), the batched
// tensor-core path (fcb_dense.cu) and the allocation-free entry point (fcb_noalloc.cu), plus the first-pass error
// bound their exactness certificates use. One copy: a fix here reaches all three.
#pragma once

#include "fcb_internal.cuh"

// tie order of the first pass: the id rank when ids do not ascend with the row, else the row itself
__device__ __forceinline__ uint32_t fcb_tie_key(uint32_t row, const uint32_t *__restrict__ id_rank) {
    return (row == FCB_NO_ROW || id_rank == nullptr) ? row : id_rank[row];
}

__device__ __forceinline__ bool fcb_cand_better(float s1, uint32_t r1, float s2, uint32_t r2,
                                                const uint32_t *__restrict__ id_rank) {
    if (s1 != s2) return s1 > s2;
    return fcb_tie_key(r1, id_rank) < fcb_tie_key(r2, id_rank);
}

// Insert (s, r) into the warp's sorted list (lane i holds the i-th best). Warp-uniform inputs.
__device__ __forceinline__ void fcb_warp_insert(float &my_s, uint32_t &my_r, float s, uint32_t r,
                                                const uint32_t *__restrict__ id_rank, int lane) {
    bool mine_better = my_s > s;
    if (my_s == s) mine_better = fcb_tie_key(my_r, id_rank) < fcb_tie_key(r, id_rank);
    const int pos = __popc(__ballot_sync(FCB_FULL, mine_better));   // better entries form a prefix
    const float up_s = __shfl_up_sync(FCB_FULL, my_s, 1);
    const uint32_t up_r = __shfl_up_sync(FCB_FULL, my_r, 1);
    if (lane == pos) { my_s = s; my_r = r; }
    else if (lane > pos) { my_s = up_s; my_r = up_r; }
}

// Merge a sorted 32-entry list (entry i in lane i of `e`) into the warp's sorted list: lane i keeps the
// better of mine[i] and other[31 - i] (a bitonic sequence holding the 32 best of the union), then a
// 5-step bitonic merge restores best-first order. ~50 instructions instead of up to 32 insertions.
__device__ __forceinline__ void fcb_warp_merge_list(float &my_s, uint32_t &my_r, FcbCand e,
                                                    const uint32_t *__restrict__ id_rank, int lane) {
    const float thr = __shfl_sync(FCB_FULL, my_s, 31);
    const float top = __shfl_sync(FCB_FULL, e.score, 0);
    const uint32_t top_r = __shfl_sync(FCB_FULL, e.row, 0);
    if (top_r == FCB_NO_ROW || !(top >= thr)) return;          // nothing in `e` can enter
    const int n_valid = __popc(__ballot_sync(FCB_FULL, e.row != FCB_NO_ROW));
    if (n_valid <= 3) {                                        // short list (small segments): plain insertions
        for (int i = 0; i < n_valid; ++i)
            fcb_warp_insert(my_s, my_r, __shfl_sync(FCB_FULL, e.score, i), __shfl_sync(FCB_FULL, e.row, i), id_rank, lane);
        return;
    }
    const float os = __shfl_sync(FCB_FULL, e.score, 31 - lane);
    const uint32_t orow = __shfl_sync(FCB_FULL, e.row, 31 - lane);
    if (fcb_cand_better(os, orow, my_s, my_r, id_rank)) { my_s = os; my_r = orow; }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        const float ps = __shfl_xor_sync(FCB_FULL, my_s, d);
        const uint32_t pr = __shfl_xor_sync(FCB_FULL, my_r, d);
        const bool want_better = (lane & d) == 0;              // lower lane of the pair keeps the better entry
        if (fcb_cand_better(ps, pr, my_s, my_r, id_rank) == want_better && (ps != my_s || pr != my_r)) { my_s = ps; my_r = pr; }
    }
}

// Two independent merges in one instruction stream (no early exits, so the two shuffle chains
// interleave): used by the last CTA, which has up to 2 * SMs lists to fold.
__device__ __forceinline__ void fcb_warp_merge_list2(float &s0, uint32_t &r0, FcbCand e0, float &s1, uint32_t &r1, FcbCand e1,
                                                     const uint32_t *__restrict__ id_rank, int lane) {
    const float o0 = __shfl_sync(FCB_FULL, e0.score, 31 - lane), o1 = __shfl_sync(FCB_FULL, e1.score, 31 - lane);
    const uint32_t q0 = __shfl_sync(FCB_FULL, e0.row, 31 - lane), q1 = __shfl_sync(FCB_FULL, e1.row, 31 - lane);
    if (fcb_cand_better(o0, q0, s0, r0, id_rank)) { s0 = o0; r0 = q0; }
    if (fcb_cand_better(o1, q1, s1, r1, id_rank)) { s1 = o1; r1 = q1; }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        const float p0 = __shfl_xor_sync(FCB_FULL, s0, d), p1 = __shfl_xor_sync(FCB_FULL, s1, d);
        const uint32_t t0 = __shfl_xor_sync(FCB_FULL, r0, d), t1 = __shfl_xor_sync(FCB_FULL, r1, d);
        const bool want_better = (lane & d) == 0;
        if (fcb_cand_better(p0, t0, s0, r0, id_rank) == want_better && (p0 != s0 || t0 != r0)) { s0 = p0; r0 = t0; }
        if (fcb_cand_better(p1, t1, s1, r1, id_rank) == want_better && (p1 != s1 || t1 != r1)) { s1 = p1; r1 = t1; }
    }
}

// 8 warp lists -> warp 0 through shared memory, as a 3-level tree (3 merges deep instead of 7).
__device__ __forceinline__ void fcb_cta_merge_tree(float &my_s, uint32_t &my_r, FcbCand (*m_s)[FCB_FUSED_CAND],
                                                   const uint32_t *__restrict__ id_rank, int warp, int lane) {
    m_s[warp][lane] = FcbCand{my_s, my_r};
    __syncthreads();
#pragma unroll
    for (int stride = 1; stride < FCB_SCAN_WARPS; stride <<= 1) {
        if ((warp & (2 * stride - 1)) == 0) {
            fcb_warp_merge_list(my_s, my_r, m_s[warp + stride][lane], id_rank, lane);
            m_s[warp][lane] = FcbCand{my_s, my_r};
        }
        __syncthreads();
    }
}

// Warp 0 only: the 32 survivors `cand` (first-pass list, best first) with their canonical float64 scores and ids
// -> the first k by (score desc, id asc) into the outputs, tail padded, count set. Returns through `kth` the exact
// score of the last returned entry (-inf when nothing is returned) and through `nvalid` how many survivors exist;
// the callers' exactness certificates compare `kth` with the worst first-pass score they kept.
__device__ __forceinline__ void fcb_rank_and_write(const FcbCand *cand, const double *f_score, const uint64_t *f_id, int k,
                                                   int lane, uint64_t *out_ids, double *out_scores, int32_t *out_count,
                                                   double &kth, int &nvalid) {
    const bool valid = cand[lane].row != FCB_NO_ROW;
    nvalid = __popc(__ballot_sync(FCB_FULL, valid));
    const double ms = f_score[lane];
    const uint64_t mi = f_id[lane];
    int rank = 0;
#pragma unroll 8
    for (int j = 0; j < FCB_FUSED_CAND; ++j)
        rank += (cand[j].row != FCB_NO_ROW) && fcb_better64(f_score[j], f_id[j], ms, mi);
    const int count = nvalid < k ? nvalid : k;
    if (valid && rank < k) { out_ids[rank] = mi; out_scores[rank] = ms; }
    if (lane >= count && lane < k) { out_ids[lane] = UINT64_MAX; out_scores[lane] = -INFINITY; }
    if (lane == 0) *out_count = count;
    const unsigned kb = __ballot_sync(FCB_FULL, valid && rank == count - 1);
    const double cand_kth = __shfl_sync(FCB_FULL, ms, kb ? __ffs(kb) - 1 : 0);
    kth = kb ? cand_kth : -INFINITY;
}

// Worst-case distance, in cosine units, between a float32 first-pass score (what the scans rank rows by) and the canonical
// float64 score of the same row. The first pass computes fl32(dot32 * inv32) with
//   dot32  lane-strided fmaf chains of n = dim/32 + 8 steps at most, then a 5-step butterfly, over the query rounded to
//          float32:  |dot32 - x.q| <= (n + 6) u |x||q|                     (u = 2^-24; Higham's gamma_n, Cauchy-Schwarz)
//   inv32  rsqrtf (<= 2 ulp = 4 u) of a float32 sum of squares of the same shape (relative error <= (n + 5) u, halved by
//          the square root):  inv32 = (1 + d) / |x|, |d| <= ((n + 5) / 2 + 4) u
//   one rounding of the product.
// Together |cos32 - cos| <= (1.5 n + 18) u; the canonical score is within 1e-12 of the true cosine. The bound holds for
// rows and queries whose norms lie in [2^-50, 2^50] (no float32 overflow / underflow): fcb_load_shard counts the rows
// outside and the kernels check the query; anything outside is answered by the exact float64 path.
__host__ __device__ inline double fcb_first_pass_eps(int dim) {
    const double n = (double)(dim / 32 + 8);
    return (1.5 * n + 18.0) * 5.9604644775390625e-8 + 1e-12;
}
