// Batched dense path of fcb_search (BASELINE config 3): many concurrent queries against ONE shared
// row range of a bfloat16 shard become a real contraction
//     S[row, query] = sum_d X[row, d] * Q[query, d]
// and run on the 5th-generation tensor cores: TMA (cp.async.bulk.tensor, 128-byte swizzle) stages
// corpus and query tiles in shared memory, one elected thread issues tcgen05.mma (kind::f16,
// bf16 x bf16 -> fp32) with the accumulators in TMEM, and the epilogue warps read them back with
// tcgen05.ld and run a fused per-tile top-k filter. The corpus is streamed from HBM exactly once.
//
// One CTA per SM (persistent, 384 threads):
//   warp 0      TMA producer: per k-stage (64 columns) loads two 128-row corpus tiles (A0, A1) and
//               the 256-query tile (B) into a 3-deep ring (64 KB per stage), mbarrier expect_tx
//   warp 1      MMA issuer: per stage 4 x UMMA(128x256x16) into accumulator 0 (TMEM cols 0..255)
//               and 4 into accumulator 1 (cols 256..511), so every B tile read from L2 is used
//               for 256 corpus rows; tcgen05.commit frees the stage / signals the epilogue
//   warp 2      TMEM allocation (512 columns) and release
//   warps 4-11  epilogue: warps 4-7 own accumulator 0, warps 8-11 accumulator 1 (TMEM lane = row).
//               A thread scales its row's 256 scores by the row's inverse norm and compares each
//               with that query's current threshold (shared memory); the rare survivors are pushed
//               to a per-query pending list. After the super-tile, each warp folds the pending
//               entries of its 32 queries into sorted top-32 lists and raises the thresholds.
// After the grid finishes, dense_finalize_kernel merges the per-CTA lists per query, re-scores
// the 32 survivors in float64 against the float64 query (canonical routine), ranks them and
// writes the result — plus a certificate: the first pass used a bf16-rounded query, so a query is
// certified only if the worst kept first-pass score lies a safe margin below the k-th exact score
// (or fewer than 32 rows qualified). Uncertified queries are re-run by the exact fp32-query scan
// kernel in the same stream (device-side flag, no host round trip).
#include <cuda.h>
#include <stdlib.h>

#include "fcb_internal.cuh"
#include "fcb_topk.cuh"

#define DN_THREADS 384
#define DN_BLOCK_M 128
#define DN_SUPER_M 256                      // rows per super-tile (two accumulators)
#define DN_N 256                            // queries per pass (padded with zero rows)
#define DN_BLOCK_K 64                       // 64 bf16 = 128 bytes = one swizzle atom
#define DN_STAGES 3
#define DN_A_BYTES (DN_BLOCK_M * DN_BLOCK_K * 2)      // 16 KB
#define DN_B_BYTES (DN_N * DN_BLOCK_K * 2)            // 32 KB
#define DN_STAGE_BYTES (2 * DN_A_BYTES + DN_B_BYTES)  // 64 KB
#define DN_TMEM_COLS 512
#define DN_PEND_CAP 256                     // a query can receive at most 256 pushes per super-tile
#define DN_SPIN_LIMIT (1u << 28)

struct DenseParams {
    const float *inv_norm;
    const uint32_t *id_rank;
    const int32_t *dates;
    const uint8_t *cats;
    fcb_filter filter;                      // shared by the whole batch
    int64_t n_rows, row_base;
    int dim;
    int n_queries;                          // real queries (columns beyond are zero padding and are ignored)
    int n_eff;                              // UMMA N = n_queries rounded up to 32
    FcbCand *cta_top;                       // [grid][DN_N][32] sorted per-CTA lists
    FcbCand *pend;                          // [grid][DN_N][DN_PEND_CAP]
    int *error_flag;
    float *debug_out;                       // diagnostic: raw accumulators [debug_rows, DN_N] (NULL in production)
    int64_t debug_rows;
    long long *prof;                        // diagnostic: per-role cycle counters of CTA 0 (NULL in production)
    const float *slack;                     // [DN_N] how far below the cross-CTA bound a row may lie and still be listed (first-pass units)
    float *thr_out;                         // [DN_N][gt_stride] final pruning threshold of every CTA and query: no unlisted row scores above it
    int n_stages;                           // cta_group::2 kernel: depth of the shared-memory ring (3 or 4)
    unsigned *gtop;                         // [DN_N][gt_stride] best score of every CTA's list, order-preserving uint (cta_group::2 kernel)
    int gt_stride;
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// bounded wait: a protocol bug must fault the launch, never hang the GPU
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity, int *error_flag, int code) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > DN_SPIN_LIMIT) {
            *reinterpret_cast<volatile int *>(error_flag) = code;   // mapped pinned memory: the host reads it after the trap
            __threadfence_system();
            asm volatile("trap;");
        }
    }
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar, uint64_t hint) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
                 "[%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(hint) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}

// shared-memory matrix descriptor, K-major, 128-byte swizzle (cute::UMMA::SmemDescriptor layout):
// start>>4 [0,14) | LBO>>4 [16,30) = 1 (unused) | SBO>>4 [32,46) = 1024 B between 8-row groups |
// version [46,48) = 1 | layout_type [61,64) = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 (1) [4,6), a/b format BF16 (1)
// [7,10) / [10,13), both K-major, N>>3 [17,23), M>>4 [24,29). (A bf16 A operand with a float16 B operand — 8x finer
// rounding of the query tile — was tried: kind::f16 traps with "illegal instruction" when the two formats differ.)
__device__ __forceinline__ constexpr uint32_t make_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// order-preserving float <-> uint (larger float = larger uint), for atomicMax on scores of either sign
__device__ __forceinline__ unsigned dn_ord(float f) {
    const unsigned u = __float_as_uint(f);
    return u ^ ((u >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float dn_unord(unsigned o) {
    return __uint_as_float(o ^ ((o >> 31) ? 0x80000000u : 0xffffffffu));
}

__global__ void __launch_bounds__(DN_THREADS, 1)
dense_topk_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_q,
                  const __grid_constant__ DenseParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    // [stage][A0 | A1 | B] then barriers / bookkeeping
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + DN_STAGES * DN_STAGE_BYTES);
    uint64_t *empty_bar = full_bar + DN_STAGES;
    uint64_t *tmem_full = empty_bar + DN_STAGES;
    uint64_t *tmem_empty = tmem_full + 1;
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(tmem_empty + 1);
    float *thr_s = reinterpret_cast<float *>(tmem_ptr + 2);          // [DN_N]
    int *cnt_s = reinterpret_cast<int *>(thr_s + DN_N);              // [DN_N]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const fcb_filter f = p.filter;
    int64_t lo = f.row_begin - p.row_base, hi = f.row_end - p.row_base;
    if (lo < 0) lo = 0;
    if (hi > p.n_rows) hi = p.n_rows;
    const int64_t nrows = hi > lo ? hi - lo : 0;
    const int n_super = (int)((nrows + DN_SUPER_M - 1) / DN_SUPER_M);
    const int k_stages = p.dim / DN_BLOCK_K;

    FcbCand *my_top = p.cta_top + (size_t)blockIdx.x * DN_N * FCB_FUSED_CAND;
    FcbCand *my_pend = p.pend + (size_t)blockIdx.x * DN_N * DN_PEND_CAP;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < DN_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full, 1);
        mbar_init(tmem_empty, 256);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(DN_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < DN_N; i += DN_THREADS) { thr_s[i] = -INFINITY; cnt_s[i] = 0; }
    for (int i = tid; i < DN_N * FCB_FUSED_CAND; i += DN_THREADS) my_top[i] = FcbCand{-INFINITY, FCB_NO_ROW};
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            const uint64_t hint_stream = 0x12F0000000000000ull;      // evict-first: corpus is read once
            const uint64_t hint_keep = 0x14F0000000000000ull;        // evict-last: queries are re-read by every tile
            int stage = 0;
            uint32_t phase = 0;
            for (int st = blockIdx.x; st < n_super; st += gridDim.x) {
                const int row0 = (int)(lo + (int64_t)st * DN_SUPER_M);
                for (int ks = 0; ks < k_stages; ++ks) {
                    mbar_wait(&empty_bar[stage], phase ^ 1, p.error_flag, 1);
                    unsigned char *sp = smem + stage * DN_STAGE_BYTES;
                    mbar_expect_tx(&full_bar[stage], 2 * DN_A_BYTES + (uint32_t)p.n_eff * DN_BLOCK_K * 2);
                    tma_load_2d(sp, &map_x, ks * DN_BLOCK_K, row0, &full_bar[stage], hint_stream);
                    tma_load_2d(sp + DN_A_BYTES, &map_x, ks * DN_BLOCK_K, row0 + DN_BLOCK_M, &full_bar[stage], hint_stream);
                    tma_load_2d(sp + 2 * DN_A_BYTES, &map_q, ks * DN_BLOCK_K, 0, &full_bar[stage], hint_keep);
                    if (++stage == DN_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            const uint32_t idesc = make_idesc_bf16(DN_BLOCK_M, p.n_eff);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int st = blockIdx.x; st < n_super; st += gridDim.x, ++it) {
                mbar_wait(tmem_empty, (it & 1) ^ 1, p.error_flag, 2);             // epilogue drained the accumulators
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int ks = 0; ks < k_stages; ++ks) {
                    mbar_wait(&full_bar[stage], phase, p.error_flag, 3);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = smem_u32(smem + stage * DN_STAGE_BYTES);
                    const uint64_t a0 = make_sw128_desc(sa), a1 = make_sw128_desc(sa + DN_A_BYTES),
                                   b = make_sw128_desc(sa + 2 * DN_A_BYTES);
#pragma unroll
                    for (int kk = 0; kk < DN_BLOCK_K / 16; ++kk) {                // +32 bytes per UMMA_K: +2 in the >>4 field
                        umma_bf16(tmem_base, a0 + 2 * kk, b + 2 * kk, idesc, (ks | kk) != 0);
                        umma_bf16(tmem_base + DN_N, a1 + 2 * kk, b + 2 * kk, idesc, (ks | kk) != 0);
                    }
                    umma_commit(&empty_bar[stage]);                              // stage reusable once these MMAs retire
                    if (++stage == DN_STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(tmem_full);                                          // accumulators complete
            }
        }
    } else if (warp >= 4) {
        // ================= epilogue =================
        const int ew = warp - 4;                           // 0..7
        const int acc = ew >> 2;                           // accumulator 0 / 1
        const int lane_row = (warp & 3) * 32 + lane;       // TMEM lane this thread may read
        const bool has_date = f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX;
        const bool has_cat = f.category_mask != FCB_ALL_CATEGORIES;
        int it = 0;
        for (int st = blockIdx.x; st < n_super; st += gridDim.x, ++it) {
            const int64_t row = lo + (int64_t)st * DN_SUPER_M + acc * DN_BLOCK_M + lane_row;
            bool valid = row < hi;
            if (valid && has_date) { const int64_t d = p.dates[row]; valid = d >= f.date_lo && d <= f.date_hi; }
            if (valid && has_cat) { const unsigned c = p.cats[row]; valid = c < 64 && ((f.category_mask >> c) & 1ull); }
            const float inv = valid ? p.inv_norm[row] : 0.f;

            mbar_wait(tmem_full, it & 1, p.error_flag, 4);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + acc * DN_N;
            // ids ascending with the row (no id_rank): a later row never wins a tie, so ties need not be pushed
            const bool push_ties = p.id_rank != nullptr;
#pragma unroll 1
            for (int c = 0; c < p.n_eff; c += 32) {
                uint32_t v[32];
                tmem_ld32(taddr + c, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (p.debug_out != nullptr && row - lo < p.debug_rows) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) p.debug_out[(row - lo) * DN_N + c + j] = __uint_as_float(v[j]);
                }
                if (valid) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float s = __uint_as_float(v[j]) * inv;
                        const float t = thr_s[c + j];
                        if ((s > t || (push_ties && s == t)) && c + j < p.n_queries) {
                            const int slot = atomicAdd(&cnt_s[c + j], 1);
                            my_pend[(size_t)(c + j) * DN_PEND_CAP + slot] = FcbCand{s, (uint32_t)row};
                        }
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(tmem_empty);                                             // MMA may overwrite TMEM
            __threadfence_block();
            asm volatile("bar.sync 1, 256;" ::: "memory");
            // fold pending entries into the sorted per-query lists; warp ew owns queries ew, ew+8, ...
            for (int q = ew; q < p.n_queries; q += 8) {
                const int n = cnt_s[q];
                if (n == 0) continue;
                FcbCand e = my_top[(size_t)q * FCB_FUSED_CAND + lane];
                float thr = __shfl_sync(FCB_FULL, e.score, 31);
                for (int base = 0; base < n; base += 32) {
                    FcbCand pe{-INFINITY, FCB_NO_ROW};
                    if (base + lane < n) pe = my_pend[(size_t)q * DN_PEND_CAP + base + lane];
                    const int m = n - base < 32 ? n - base : 32;
                    for (int j = 0; j < m; ++j) {
                        const float s = __shfl_sync(FCB_FULL, pe.score, j);
                        const uint32_t r = __shfl_sync(FCB_FULL, pe.row, j);
                        if (s >= thr) {
                            fcb_warp_insert(e.score, e.row, s, r, p.id_rank, lane);
                            thr = __shfl_sync(FCB_FULL, e.score, 31);
                        }
                    }
                }
                my_top[(size_t)q * FCB_FUSED_CAND + lane] = e;
                if (lane == 31) thr_s[q] = e.score;
                if (lane == 0) cnt_s[q] = 0;
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
        }
    }
    __syncthreads();
    for (int q = tid; q < p.n_queries; q += DN_THREADS) p.thr_out[(size_t)q * p.gt_stride + blockIdx.x] = thr_s[q];
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(DN_TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------------
// cta_group::2 variant: two CTAs of a cluster (an SM pair) run ONE 256-row UMMA per instruction. Each
// CTA stages its own 128-row corpus tiles but only HALF of the query tile (the pair's MMA reads both
// halves), and keeps two accumulators, so a query tile read from L2 now serves 512 corpus rows: the
// L2->SM traffic per corpus byte drops from 1.0 to 0.5 byte — the bound of the 1-CTA kernel
// (profiles/r01_summary.md). 3 stages of 48 KB; the per-query sorted top-32 lists live in shared memory
// (64 KB) so that folding the pending entries after a super-tile costs no global round trips. The leader CTA (cluster rank 0) issues every MMA; TMA
// loads of both CTAs complete on the leader's full barrier; tcgen05.commit multicasts the stage-free
// and accumulators-ready arrivals to both CTAs; both epilogues arrive remotely on the leader's
// accumulators-drained barrier.
#define DN2_THREADS 640                    // 4 service warps + 16 epilogue warps (two per TMEM lane quarter and accumulator)
#define DN2_EPI_THREADS 512
#define DN2_MAX_STAGES 4                   // 4 x 48 KB of stages when the per-query lists are small (batch <= 96), else 3
#define DN2_STAGE_BYTES (2 * DN_A_BYTES + DN_B_BYTES / 2)   // 48 KB: A0, A1, half of B (max N = 256)
#define DN2_SUPER_M 512

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar, uint64_t hint) {
    // executed by both CTAs; the peer bit of the barrier address is cleared so the bytes complete on CTA 0's barrier
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
                 "[%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1), "l"(hint) : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t *bar) {       // arrive on this barrier in BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cta0(uint64_t *bar) {     // arrive on CTA 0's copy of the barrier
    asm volatile("{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\t"
                 "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
                 ::"r"(smem_u32(bar)), "r"(0) : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(DN2_THREADS, 1)
dense_topk2_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_q,
                   const __grid_constant__ DenseParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int n_stages = p.n_stages;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + n_stages * DN2_STAGE_BYTES);
    uint64_t *empty_bar = full_bar + DN2_MAX_STAGES;
    uint64_t *tmem_full = empty_bar + DN2_MAX_STAGES;                 // [2]: one per accumulator set
    uint64_t *tmem_empty = tmem_full + 2;                             // [2]
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(tmem_empty + 2);
    float *thr_s = reinterpret_cast<float *>(smem + n_stages * DN2_STAGE_BYTES + 256);   // [DN_N], 16-byte aligned
    int *cnt_s = reinterpret_cast<int *>(thr_s + DN_N);              // [DN_N]
    float *gb_s = reinterpret_cast<float *>(cnt_s + DN_N);           // [DN_N] cross-CTA bound of each query (written by the gossip warps)
    volatile int *gb_seq = reinterpret_cast<volatile int *>(gb_s + DN_N);   // super-tiles this CTA has published; -1 = finished
    FcbCand *top_s = reinterpret_cast<FcbCand *>(smem + n_stages * DN2_STAGE_BYTES + 4096);   // [DN_N][32] sorted lists

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t cta_rank = cluster_ctarank();
    const bool leader = cta_rank == 0;
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const fcb_filter f = p.filter;
    int64_t lo = f.row_begin - p.row_base, hi = f.row_end - p.row_base;
    if (lo < 0) lo = 0;
    if (hi > p.n_rows) hi = p.n_rows;
    const int64_t nrows = hi > lo ? hi - lo : 0;
    const int n_super = (int)((nrows + DN2_SUPER_M - 1) / DN2_SUPER_M);
    const int k_stages = p.dim / DN_BLOCK_K;
    const int n_half = p.n_eff / 2;                                    // query rows this CTA stages
    // TMEM: 512 columns. Up to 128 queries the two accumulators of a super-tile take 256 of them, so there are two
    // SETS and the MMAs of super-tile t + 1 run while the epilogue still scans super-tile t; above, one set.
    const bool dbuf = p.n_eff <= DN_N / 2;
    const uint32_t acc_stride = dbuf ? DN_N / 2 : DN_N;

    FcbCand *my_top = p.cta_top + (size_t)blockIdx.x * DN_N * FCB_FUSED_CAND;
    FcbCand *my_pend = p.pend + (size_t)blockIdx.x * DN_N * DN_PEND_CAP;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < n_stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&tmem_full[a], 1);
            mbar_init(&tmem_empty[a], 2 * DN2_EPI_THREADS);            // the epilogue threads of both CTAs
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(DN_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < DN_N; i += DN2_THREADS) { thr_s[i] = i < p.n_queries ? -INFINITY : INFINITY; cnt_s[i] = 0; }   // padding never qualifies
    for (int i = tid; i < DN_N; i += DN2_THREADS) gb_s[i] = -INFINITY;
    if (tid == 0) *gb_seq = 0;
    for (int i = tid; i < p.n_eff * FCB_FUSED_CAND; i += DN2_THREADS) top_s[i] = FcbCand{-INFINITY, FCB_NO_ROW};
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();                                                // barriers + TMEM of both CTAs are ready
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer (both CTAs) =================
        if (lane == 0) {
            const uint64_t hint_stream = 0x12F0000000000000ull, hint_keep = 0x14F0000000000000ull;
            const uint32_t stage_tx = 2u * (2u * DN_A_BYTES + (uint32_t)n_half * DN_BLOCK_K * 2u);   // both CTAs' bytes
            int stage = 0;
            uint32_t phase = 0;
            long long w_empty = 0, t_begin = clock64();
            for (int st = pair; st < n_super; st += n_pairs) {
                const int row0 = (int)(lo + (int64_t)st * DN2_SUPER_M) + (int)cta_rank * DN_BLOCK_M;
                for (int ks = 0; ks < k_stages; ++ks) {
                    const long long t0 = clock64();
                    mbar_wait(&empty_bar[stage], phase ^ 1, p.error_flag, 11);
                    w_empty += clock64() - t0;
                    unsigned char *sp = smem + stage * DN2_STAGE_BYTES;
                    if (leader) mbar_expect_tx(&full_bar[stage], stage_tx);
                    tma_load_2d_2sm(sp, &map_x, ks * DN_BLOCK_K, row0, &full_bar[stage], hint_stream);
                    tma_load_2d_2sm(sp + DN_A_BYTES, &map_x, ks * DN_BLOCK_K, row0 + 256, &full_bar[stage], hint_stream);
                    tma_load_2d_2sm(sp + 2 * DN_A_BYTES, &map_q, ks * DN_BLOCK_K, (int)cta_rank * n_half, &full_bar[stage], hint_keep);
                    if (++stage == n_stages) { stage = 0; phase ^= 1; }
                }
            }
            if (p.prof && blockIdx.x == 0) { p.prof[0] = clock64() - t_begin; p.prof[1] = w_empty; }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (leader CTA only) =================
        if (leader && lane == 0) {
            const uint32_t idesc = make_idesc_bf16(2 * DN_BLOCK_M, p.n_eff);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            long long w_tmem = 0, w_full = 0, t_begin = clock64();
            for (int st = pair; st < n_super; st += n_pairs, ++it) {
                long long t0 = clock64();
                const int set = dbuf ? (it & 1) : 0, use = dbuf ? (it >> 1) : it;     // accumulator set and how often it was used
                const uint32_t acc0 = tmem_base + set * DN_N, acc1 = acc0 + acc_stride;
                mbar_wait(&tmem_empty[set], (use & 1) ^ 1, p.error_flag, 12);
                w_tmem += clock64() - t0;
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int ks = 0; ks < k_stages; ++ks) {
                    t0 = clock64();
                    mbar_wait(&full_bar[stage], phase, p.error_flag, 13);
                    w_full += clock64() - t0;
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = smem_u32(smem + stage * DN2_STAGE_BYTES);
                    const uint64_t a0 = make_sw128_desc(sa), a1 = make_sw128_desc(sa + DN_A_BYTES),
                                   b = make_sw128_desc(sa + 2 * DN_A_BYTES);
#pragma unroll
                    for (int kk = 0; kk < DN_BLOCK_K / 16; ++kk) {
                        umma_bf16_2sm(acc0, a0 + 2 * kk, b + 2 * kk, idesc, (ks | kk) != 0);
                        umma_bf16_2sm(acc1, a1 + 2 * kk, b + 2 * kk, idesc, (ks | kk) != 0);
                    }
                    umma_commit_2sm(&empty_bar[stage]);
                    if (++stage == n_stages) { stage = 0; phase ^= 1; }
                }
                umma_commit_2sm(&tmem_full[set]);
            }
            if (p.prof && blockIdx.x == 0) { p.prof[2] = clock64() - t_begin; p.prof[3] = w_tmem; p.prof[4] = w_full; }
        }
    } else if (warp == 2 || warp == 3) {
        // ================= gossip warps: cross-CTA pruning bound, off everybody's critical path =================
        // Every CTA publishes the best score of each of its lists (p.gtop[query][cta]). The CTAs' best rows are distinct
        // rows, so any value that at least 32 of them reach bounds the final global 32nd best from below. Lane l takes
        // the maximum over CTAs l, l + 32, ...; the minimum of the 32 lane maxima is reached by one CTA in every lane's
        // group, i.e. by >= 32 rows. With ~148 CTAs this lets through only ~4x the rows an exact global threshold would
        // (a CTA's own 32nd best: ~150x), which keeps the epilogue's slow path and the fold rare. Lowered one ulp so
        // that a row tying with it still gets through (it may win the tie on id). Stale reads are harmless: every
        // value ever published is a valid bound.
        int seen = 0;
        for (;;) {
            const int seq = *gb_seq;
            if (seq < 0) break;
            if (seq < seen + (seen < 4 ? 1 : 4)) { __nanosleep(500); continue; }   // early super-tiles move the bound most
            seen = seq;
#pragma unroll 1
            for (int q0 = (warp - 2) * 8; q0 < p.n_queries; q0 += 16) {      // 8 queries per round trip, the two warps alternate
                unsigned m[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {                                // 40 independent L2 loads in flight per lane
                    const int qq = q0 + u < p.n_queries ? q0 + u : q0;
                    const unsigned *g = p.gtop + (size_t)qq * p.gt_stride;
                    unsigned mm = 0;
#pragma unroll
                    for (int t = 0; t < 5; ++t) {
                        const int c = lane + 32 * t;
                        const unsigned v = c < (int)gridDim.x ? __ldcg(g + c) : 0u;
                        mm = max(mm, v);
                    }
                    m[u] = mm;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    unsigned v = m[u];
#pragma unroll
                    for (int off = 16; off >= 1; off >>= 1) v = min(v, __shfl_xor_sync(FCB_FULL, v, off));
                    // ... and lowered by the query's slack (twice the worst-case first-pass error, dense_pack_queries_kernel):
                    // every row that could still belong to the exact top k then stays listed, and dense_finalize_kernel
                    // re-scores all of those instead of giving the query up
                    if (lane == 0 && q0 + u < p.n_queries && v > dn_ord(-INFINITY)) gb_s[q0 + u] = dn_unord(v - 1) - p.slack[q0 + u];
                }
            }
        }
    } else if (warp >= 4) {
        // ================= epilogue (both CTAs; TMEM lane = row of this CTA's half) =================
        const int ew = warp - 4;                           // 0..15
        const int acc = (ew >> 2) & 1;                     // accumulator 0 / 1
        const int col_half = ew >> 3;                      // which half of the query columns this warp scans
        const int lane_row = (warp & 3) * 32 + lane;
        const bool has_date = f.date_lo != FCB_DATE_MIN || f.date_hi != FCB_DATE_MAX;
        const bool has_cat = f.category_mask != FCB_ALL_CATEGORIES;
        const bool push_ties = p.id_rank != nullptr;
        int it = 0;
        long long w_acc = 0, t_a = 0, t_b = 0, t_begin = clock64();
        for (int st = pair; st < n_super; st += n_pairs, ++it) {
            const int64_t row = lo + (int64_t)st * DN2_SUPER_M + acc * 256 + (int)cta_rank * DN_BLOCK_M + lane_row;
            bool valid = row < hi;
            if (valid && has_date) { const int64_t d = p.dates[row]; valid = d >= f.date_lo && d <= f.date_hi; }
            if (valid && has_cat) { const unsigned c = p.cats[row]; valid = c < 64 && ((f.category_mask >> c) & 1ull); }
            const float inv = valid ? p.inv_norm[row] : 0.f;

            long long t0 = clock64();
            const int set = dbuf ? (it & 1) : 0, use = dbuf ? (it >> 1) : it;
            mbar_wait(&tmem_full[set], use & 1, p.error_flag, 14);
            long long t1 = clock64();
            w_acc += t1 - t0;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + set * DN_N + acc * acc_stride;
            const int c_half = dbuf ? DN_N / 4 : DN_N / 2;                 // columns per warp: half of an accumulator's width
            const int c_begin = col_half * c_half, c_end = min(p.n_eff, c_begin + c_half);
            // one 16-column chunk of this row: branch-free pass mask (thresholds: 4 broadcast 128-bit shared loads),
            // then the rare survivors are pushed from an unrolled bit scan (no dynamic register indexing)
            auto process = [&](const uint32_t (&v)[16], const int c) {
                if (p.debug_out != nullptr && row - lo < p.debug_rows) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) p.debug_out[(row - lo) * DN_N + c + j] = __uint_as_float(v[j]);
                }
                // fast path: largest margin (score - threshold) of the 16 columns, one FFMA + max per element;
                // only a row that has a survivor somewhere in the chunk computes the exact pass mask
                const float4 *t4 = reinterpret_cast<const float4 *>(thr_s + c);
                float dmax = -INFINITY;
#pragma unroll
                for (int j4 = 0; j4 < 4; ++j4) {
                    const float4 t = t4[j4];
                    const float d0 = fmaf(__uint_as_float(v[4 * j4 + 0]), inv, -t.x), d1 = fmaf(__uint_as_float(v[4 * j4 + 1]), inv, -t.y),
                                d2 = fmaf(__uint_as_float(v[4 * j4 + 2]), inv, -t.z), d3 = fmaf(__uint_as_float(v[4 * j4 + 3]), inv, -t.w);
                    dmax = fmaxf(fmaxf(dmax, fmaxf(d0, d1)), fmaxf(d2, d3));
                }
                if (valid && dmax >= 0.f) {
                    uint32_t mask = 0;
#pragma unroll
                    for (int j4 = 0; j4 < 4; ++j4) {
                        const float4 t = t4[j4];
                        const float s0 = __uint_as_float(v[4 * j4 + 0]) * inv, s1 = __uint_as_float(v[4 * j4 + 1]) * inv,
                                    s2 = __uint_as_float(v[4 * j4 + 2]) * inv, s3 = __uint_as_float(v[4 * j4 + 3]) * inv;
                        const bool p0 = push_ties ? s0 >= t.x : s0 > t.x, p1 = push_ties ? s1 >= t.y : s1 > t.y,
                                   p2 = push_ties ? s2 >= t.z : s2 > t.z, p3 = push_ties ? s3 >= t.w : s3 > t.w;
                        mask |= ((uint32_t)p0 | ((uint32_t)p1 << 1) | ((uint32_t)p2 << 2) | ((uint32_t)p3 << 3)) << (4 * j4);
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        if (mask & (1u << j)) {
                            const int slot = atomicAdd(&cnt_s[c + j], 1);
                            my_pend[(size_t)(c + j) * DN_PEND_CAP + slot] = FcbCand{__uint_as_float(v[j]) * inv, (uint32_t)row};
                        }
                    }
                }
            };
            // double-buffered TMEM reads: the next chunk's tcgen05.ld is in flight while this chunk is scanned
            uint32_t va[16], vb[16];
            if (c_begin < c_end) {
                tmem_ld16(taddr + c_begin, va);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            }
#pragma unroll 1
            for (int c = c_begin; c < c_end; c += 32) {                 // c_end - c_begin is a multiple of 32
                tmem_ld16(taddr + c + 16, vb);
                process(va, c);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (c + 32 < c_end) tmem_ld16(taddr + c + 32, va);
                process(vb, c + 16);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive_cta0(&tmem_empty[set]);                        // the leader's MMA may overwrite this set in both TMEMs
            __threadfence_block();
            asm volatile("bar.sync 1, 512;" ::: "memory");
            t0 = clock64();
            t_a += t0 - t1;
            // fold the pending entries into the sorted lists (shared memory). Lane i looks after query
            // ew + 8 i; queries with pending entries are handled four at a time so that their (global,
            // L2-resident) pending entries are fetched with one round trip instead of four.
            {
                const int my_q = ew + 16 * lane;
                const int my_n = my_q < p.n_queries ? cnt_s[my_q] : 0;
                unsigned todo = __ballot_sync(FCB_FULL, my_n > 0);
                while (todo) {
                    int qs[4], ns[4];
                    FcbCand pe[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (todo) {
                            const int bit = __ffs(todo) - 1;
                            todo &= todo - 1;
                            qs[u] = ew + 16 * bit;
                            ns[u] = __shfl_sync(FCB_FULL, my_n, bit);
                        } else { qs[u] = -1; ns[u] = 0; }
                        pe[u] = FcbCand{-INFINITY, FCB_NO_ROW};
                        if (lane < ns[u]) pe[u] = my_pend[(size_t)qs[u] * DN_PEND_CAP + lane];
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (qs[u] < 0) continue;
                        const int q = qs[u], n = ns[u];
                        FcbCand e = top_s[q * FCB_FUSED_CAND + lane];
                        float thr = __shfl_sync(FCB_FULL, e.score, 31);
                        for (int base = 0; base < n; base += 32) {
                            FcbCand cur = pe[u];
                            if (base > 0) {
                                cur = FcbCand{-INFINITY, FCB_NO_ROW};
                                if (base + lane < n) cur = my_pend[(size_t)q * DN_PEND_CAP + base + lane];
                            }
                            const int m = n - base < 32 ? n - base : 32;
                            for (int j = 0; j < m; ++j) {
                                const float s = __shfl_sync(FCB_FULL, cur.score, j);
                                const uint32_t r = __shfl_sync(FCB_FULL, cur.row, j);
                                if (s >= thr) {
                                    fcb_warp_insert(e.score, e.row, s, r, p.id_rank, lane);
                                    thr = __shfl_sync(FCB_FULL, e.score, 31);
                                }
                            }
                        }
                        top_s[q * FCB_FUSED_CAND + lane] = e;
                        if (lane == 31) thr_s[q] = fmaxf(thr_s[q], e.score);   // never below a cross-CTA bound adopted earlier
                        if (lane == 0) cnt_s[q] = 0;
                    }
                }
            }
            asm volatile("bar.sync 1, 512;" ::: "memory");
            // thresholds across CTAs: publish the BEST first-pass score of each of this CTA's lists, adopt the bound the
            // gossip warps derived from everybody's (see there). A bound is never lowered.
            {
                const int q = (int)threadIdx.x - 128;                        // epilogue threads 0..511 <-> queries
                if (q >= 0 && q < p.n_queries) {
                    __stcg(&p.gtop[(size_t)q * p.gt_stride + blockIdx.x], dn_ord(top_s[q * FCB_FUSED_CAND].score));
                    thr_s[q] = fmaxf(thr_s[q], *reinterpret_cast<volatile float *>(gb_s + q));
                }
                if (q == 0) *gb_seq = it + 1;
            }
            asm volatile("bar.sync 1, 512;" ::: "memory");
            t_b += clock64() - t0;
        }
        if (tid == 128) *gb_seq = -1;                                  // the gossip warps may stop
        if (p.prof && blockIdx.x == 0 && tid == 128) { p.prof[5] = clock64() - t_begin; p.prof[6] = w_acc; p.prof[7] = t_a; p.prof[8] = t_b; p.prof[9] = it; }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    for (int i = tid; i < p.n_eff * FCB_FUSED_CAND; i += DN2_THREADS) my_top[i] = top_s[i];   // hand the lists to dense_finalize_kernel
    for (int q = tid; q < p.n_queries; q += DN2_THREADS) p.thr_out[(size_t)q * p.gt_stride + blockIdx.x] = thr_s[q];   // ... and the final thresholds
    cluster_sync_all();                                                // nobody still targets the peer's barriers / TMEM
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(DN_TMEM_COLS) : "memory");
    }
}

// Worst-case distance, in cosine units, between a first-pass score of the tensor-core pass and the row's canonical float64
// score. First-pass score of a row x: a = fl32(acc * inv32), acc = the accumulator of x . h, h = bf16(q).
//   rho      = |h - q| / |q|             rounding of the query tile: |(h - q) . x| / (|x||q|) <= rho for ANY row (Cauchy-Schwarz);
//                                        measured from the tile itself, ~1.1e-3
//   eps_acc  = (dim / 16) * 17 * 2^-23   the accumulator: every UMMA K-step adds 16 exact products (8 x 8 significand bits) to it;
//                                        the model is one truncation to 24 significant bits, relative to the largest addend
//                                        (itself <= |x||h|), per addend and step: 2.0e-4 at dim 1536.
//                                        tests/test_gpu_dense.py measures the real error against this bound (~100x smaller).
//   eps_inv  = (n / 2 + 8) * 2^-24       the float32 inverse norm (sum of squares of n = dim/32 + 8 steps, rsqrtf) and the product
__host__ __device__ inline double dense_margin(double rho, int dim) {
    const double n_chain = (double)(dim / 32 + 8);
    const double eps_acc = (double)(dim / 16) * 17.0 * 1.1920928955078125e-7 * (1.0 + rho);
    const double eps_inv = (0.5 * n_chain + 8.0) * 5.9604644775390625e-8;
    return rho * (1.0 + 1e-9) + eps_acc + eps_inv + 1e-12;
}

// float64 queries -> the bf16 query tile [DN_N, dim] of the tensor-core pass (rows >= batch zero), one CTA per row; plus
// each query's SLACK in first-pass units (cosine x |q|): twice the worst-case first-pass error. The kernel lists every row
// whose first-pass score is within the slack of its cross-CTA pruning bound, so that dense_finalize_kernel finds every row
// that could still belong to the exact top k among the listed ones.
__global__ void __launch_bounds__(256) dense_pack_queries_kernel(const double *__restrict__ q, int batch, int dim,
                                                                 __nv_bfloat16 *__restrict__ out, float *__restrict__ slack) {
    __shared__ double s_qq[8], s_ee[8];
    const int b = blockIdx.x, tid = threadIdx.x;
    __nv_bfloat16 *row = out + (size_t)b * dim;
    if (b >= batch) {
        for (int j = tid; j < dim; j += 256) row[j] = __float2bfloat16_rn(0.f);
        if (tid == 0) slack[b] = 0.f;
        return;
    }
    const double *qb = q + (size_t)b * dim;
    double qq = 0.0, ee = 0.0;
    for (int j = tid; j < dim; j += 256) {
        const double qv = qb[j];
        const __nv_bfloat16 h = __float2bfloat16_rn((float)qv);
        row[j] = h;
        const double er = (double)__bfloat162float(h) - qv;
        qq = fma(qv, qv, qq);
        ee = fma(er, er, ee);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        qq += __shfl_xor_sync(FCB_FULL, qq, off);
        ee += __shfl_xor_sync(FCB_FULL, ee, off);
    }
    if ((tid & 31) == 0) { s_qq[tid >> 5] = qq; s_ee[tid >> 5] = ee; }
    __syncthreads();
    if (tid == 0) {
        double tq = 0.0, te = 0.0;
        for (int w = 0; w < 8; ++w) { tq += s_qq[w]; te += s_ee[w]; }
        const double nq = sqrt(tq);
        const double rho = nq > 0.0 ? sqrt(te) / nq : 0.0;
        slack[b] = __double2float_ru(2.0 * dense_margin(rho, dim) * nq * 1.001);
    }
}

struct DenseFinalParams {
    const void *vectors;
    const uint64_t *ids;
    const uint32_t *id_rank;
    const double *queries;
    const FcbCand *cta_top;
    const float *thr_out;                   // [DN_N][gt_stride] final pruning thresholds of the CTAs
    int gt_stride;
    int n_lists;
    int64_t row_base;
    int dim, k;
    uint64_t *out_ids;
    double *out_scores;
    int32_t *out_counts;
    int32_t *redo;                          // [batch]: 1 = could not be answered exactly from the lists: re-run by the exact scan
    int32_t *redo_list, *redo_count;        // compact list of those queries for the re-run launch
    int32_t *widened;                       // [batch]: 1 = answered by re-scoring more than the 32 best listed rows (measurement aid)
    int rigorous;                           // 1 (default): worst-case margin, 0: 8-sigma margin of the rounding model (A/B only)
};

#define DN_EXTRA_CAP 512                    // listed rows beyond the best 32 that one query may have re-scored

// One CTA per query. EXACT BY CONSTRUCTION, with m = dense_margin (the worst-case first-pass error):
//   1. the 32 best listed rows by first-pass score are re-scored canonically; E = the k-th best exact score among them.
//      No row outside the true top k... rather: every row of the true top k scores >= E exactly, hence >= E - m in the
//      first pass. Let W = E - m.
//   2. If the worst of the 32 lies below W, the 32 hold every such row: done (the common case).
//   3. Otherwise the candidates are ALL rows with a first-pass score >= W. A row that is not in any CTA's list was pruned
//      by a threshold that CTA exported (thr_out): its first-pass score is <= U = the largest of them. If U < W, every
//      candidate IS listed: re-score those beyond the first 32 too, rank all of them, write the k best. (The kernel keeps
//      its cross-CTA bound 2 m below the 32nd best score it can prove, so U < W holds unless some CTA's own 32-deep list
//      overflowed inside the window.)
//   4. Else (or with more than DN_EXTRA_CAP candidates): the flag is set and the exact float32-query scan, itself certified
//      and backed by the float64 kernel, answers the query.
__global__ void __launch_bounds__(256) dense_finalize_kernel(const __grid_constant__ DenseFinalParams p) {
    __shared__ FcbCand m_s[8][FCB_FUSED_CAND];
    __shared__ double f_score[FCB_FUSED_CAND];
    __shared__ uint64_t f_id[FCB_FUSED_CAND];
    __shared__ double x_score[DN_EXTRA_CAP];
    __shared__ uint64_t x_id[DN_EXTRA_CAP];
    __shared__ int s_mode;                  // 0: done, 1: widen (re-score every listed row >= s_window)
    __shared__ float s_window;
    __shared__ FcbCand s_worst;             // the worst of the 32 first-pass survivors
    __shared__ unsigned s_extra;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    const __nv_bfloat16 *vec = reinterpret_cast<const __nv_bfloat16 *>(p.vectors);
    const double *qd = p.queries + (size_t)b * p.dim;
    float my_s = -INFINITY;
    uint32_t my_r = FCB_NO_ROW;
    for (int L = warp; L < p.n_lists; L += 8)
        fcb_warp_merge_list(my_s, my_r, p.cta_top[((size_t)L * DN_N + b) * FCB_FUSED_CAND + lane], p.id_rank, lane);
    fcb_cta_merge_tree(my_s, my_r, m_s, p.id_rank, warp, lane);
    for (int c = warp; c < FCB_FUSED_CAND; c += 8) {
        const uint32_t row = m_s[0][c].row;
        double sc = -INFINITY;
        uint64_t id = UINT64_MAX;
        if (row != FCB_NO_ROW) {
            sc = fcb_canon_score_warp<__nv_bfloat16>(vec + (int64_t)row * p.dim, qd, p.dim, lane);
            id = p.ids ? p.ids[row] : (uint64_t)(p.row_base + row);
        }
        if (lane == 0) { f_score[c] = sc; f_id[c] = id; }
    }
    if (tid == 0) { s_mode = 0; s_extra = 0; }
    __syncthreads();
    if (warp == 0) {
        double kth;
        int nvalid;
        fcb_rank_and_write(m_s[0], f_score, f_id, p.k, lane, p.out_ids + (size_t)b * p.k, p.out_scores + (size_t)b * p.k,
                           p.out_counts + b, kth, nvalid);
        double qq = 0.0, ee = 0.0;
        for (int j = lane; j < p.dim; j += 32) {
            const double qv = qd[j];
            const double er = (double)__bfloat162float(__float2bfloat16_rn((float)qv)) - qv;
            qq = __fma_rn(qv, qv, qq);
            ee = __fma_rn(er, er, ee);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            qq += __shfl_xor_sync(FCB_FULL, qq, off);
            ee += __shfl_xor_sync(FCB_FULL, ee, off);
        }
        const double nq = sqrt(qq);
        const double rho = nq > 0.0 ? sqrt(ee) / nq : 1.0;
        const double margin = p.rigorous ? dense_margin(rho, p.dim) : 8.0 * rho * sqrt(3.0 / (double)p.dim) + 1e-6;
        const float worst32 = m_s[0][FCB_FUSED_CAND - 1].score;
        // the largest first-pass score an UNLISTED row can have
        float U = -INFINITY;
        for (int c = lane; c < p.n_lists; c += 32) U = fmaxf(U, p.thr_out[(size_t)b * p.gt_stride + c]);
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) U = fmaxf(U, __shfl_xor_sync(FCB_FULL, U, off));
        if (lane == 0) {
            int flag = 0, widened = 0;
            if (nvalid == FCB_FUSED_CAND) {
                const bool domain = nq > 8.9e-16 && nq < 1.1e15 && isfinite(worst32) && isfinite(kth);
                const float window = domain ? __double2float_rd((kth - margin) * nq) : 0.f;   // W in first-pass units, rounded down
                if (!domain) flag = 1;
                else if (worst32 < window) flag = 0;                  // the 32 hold every row that could matter
                else if (!p.rigorous || !(U < window)) flag = 1;      // an unlisted row could matter
                else { s_mode = 1; s_window = window; s_worst = m_s[0][FCB_FUSED_CAND - 1]; widened = 1; }
            }
            p.redo[b] = flag;
            p.widened[b] = widened;
            if (flag) p.redo_list[atomicAdd(p.redo_count, 1)] = b;
        }
    }
    __syncthreads();
    if (!s_mode) return;

    // ---- widen: every listed row with a first-pass score >= the window that is not one of the 32 already re-scored
    const float window = s_window;
    const FcbCand worst = s_worst;
    for (int L = warp; L < p.n_lists; L += 8) {
        const FcbCand e = p.cta_top[((size_t)L * DN_N + b) * FCB_FUSED_CAND + lane];
        const bool want = e.row != FCB_NO_ROW && e.score >= window && fcb_cand_better(worst.score, worst.row, e.score, e.row, p.id_rank);
        unsigned todo = __ballot_sync(FCB_FULL, want);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const uint32_t row = __shfl_sync(FCB_FULL, e.row, src);
            const double sc = fcb_canon_score_warp<__nv_bfloat16>(vec + (int64_t)row * p.dim, qd, p.dim, lane);
            if (lane == 0) {
                const unsigned slot = atomicAdd(&s_extra, 1u);
                if (slot < DN_EXTRA_CAP) { x_score[slot] = sc; x_id[slot] = p.ids ? p.ids[row] : (uint64_t)(p.row_base + row); }
            }
        }
    }
    __syncthreads();
    const unsigned n_extra = s_extra;
    if (n_extra > DN_EXTRA_CAP) {                                       // too many near-ties for this kernel: the exact scan answers
        if (tid == 0) { p.redo[b] = 1; p.widened[b] = 0; p.redo_list[atomicAdd(p.redo_count, 1)] = b; }
        return;
    }
    // rank the 32 + n_extra candidates by (score desc, id asc); ids are unique, so the ranks are distinct
    const unsigned n = FCB_FUSED_CAND + n_extra;
    for (unsigned e = tid; e < n; e += 256) {
        const double ms = e < FCB_FUSED_CAND ? f_score[e] : x_score[e - FCB_FUSED_CAND];
        const uint64_t mi = e < FCB_FUSED_CAND ? f_id[e] : x_id[e - FCB_FUSED_CAND];
        int rank = 0;
        for (unsigned j = 0; j < FCB_FUSED_CAND; ++j) rank += fcb_better64(f_score[j], f_id[j], ms, mi);
        for (unsigned j = 0; j < n_extra; ++j) rank += fcb_better64(x_score[j], x_id[j], ms, mi);
        if (rank < p.k) { p.out_ids[(size_t)b * p.k + rank] = mi; p.out_scores[(size_t)b * p.k + rank] = ms; }
    }
}

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int32_t encode_map(fcb_engine *e, CUtensorMap *map, const void *base, uint64_t rows, uint64_t dim, uint32_t box_rows) {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess || !sym)
            return fcb_fail(e, FCB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
        fn = (PFN_encodeTiled)sym;
    }
    const cuuint64_t gdim[2] = {dim, rows};
    const cuuint64_t gstride[1] = {dim * 2};
    const cuuint32_t box[2] = {DN_BLOCK_K, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fcb_fail(e, FCB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return FCB_OK;
}

struct FcbDense {
    CUtensorMap map_x;
    CUtensorMap map_q[DN_N / 32];           // query tile maps for N = 32, 64, ..., 256 (box rows = N)
    CUtensorMap map_qh[DN_N / 32];          // ... box rows = N / 2 for the cta_group::2 kernel
    int rigorous = 0;
    int two_cta = 1;
    __nv_bfloat16 *q_bf16 = nullptr;        // [DN_N, dim]
    float *slack = nullptr;                 // [DN_N] per-query slack of the cross-CTA pruning bound (first-pass units)
    float *thr_out = nullptr;               // [DN_N][gt_stride] final pruning thresholds
    int32_t *widened = nullptr;             // [DN_N]
    FcbCand *cta_top = nullptr, *pend = nullptr;
    unsigned *gtop = nullptr;               // [DN_N][gt_stride] per-CTA best scores of the cta_group::2 kernel
    int gt_stride = 0;
    int32_t *redo = nullptr;                // [DN_N] flags, then [DN_N] compact list, then the count
    int *error_flag = nullptr;              // device alias of h_error (mapped pinned memory: readable by the host after a trap)
    int *h_error = nullptr;
    int grid = 0;
};

// The bounded mbarrier waits of the tensor-core kernels write a code here and trap when the pipeline protocol is violated
// (a bug, never an input): 1/11 producer, 2/12 MMA issuer waiting for drained accumulators, 3/13 for a full stage, 4/14
// epilogue. The word lives in mapped pinned memory, so the host can still read it after the trap killed the context.
int fcb_dense_error_code(const fcb_engine *e) {
    const FcbDense *d = e ? reinterpret_cast<const FcbDense *>(e->dense) : nullptr;
    return d && d->h_error ? *reinterpret_cast<volatile int *>(d->h_error) : 0;
}

void fcb_dense_free(fcb_engine *e) {
    FcbDense *d = reinterpret_cast<FcbDense *>(e->dense);
    if (!d) return;
    cudaFree(d->q_bf16); cudaFree(d->slack); cudaFree(d->thr_out); cudaFree(d->widened); cudaFree(d->cta_top); cudaFree(d->pend); cudaFree(d->redo); cudaFreeHost(d->h_error); cudaFree(d->gtop);
    delete d;
    e->dense = nullptr;
}

bool fcb_dense_eligible(const fcb_engine *e) {
    const fcb_shard &s = e->shard;
    return s.dtype == FCB_DTYPE_BF16 && s.dim % DN_BLOCK_K == 0 && e->max_batch >= FCB_DENSE_MIN_BATCH &&
           s.n_rows < (int64_t)0x7fffff00 && ((uintptr_t)s.vectors & 127) == 0;
}

int32_t fcb_dense_alloc(fcb_engine *e) {
    const fcb_shard &s = e->shard;
    FcbDense *d = new FcbDense();
    e->dense = d;
    d->grid = e->num_sms;
    d->gt_stride = ((d->grid + 31) / 32) * 32;
    const size_t lists = (size_t)d->grid * DN_N;
    if (cudaMalloc(&d->q_bf16, (size_t)DN_N * s.dim * 2) != cudaSuccess ||
        cudaMalloc(&d->slack, (size_t)DN_N * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&d->thr_out, (size_t)DN_N * d->gt_stride * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&d->widened, (size_t)DN_N * sizeof(int32_t)) != cudaSuccess ||
        cudaMalloc(&d->cta_top, lists * FCB_FUSED_CAND * sizeof(FcbCand)) != cudaSuccess ||
        cudaMalloc(&d->pend, lists * DN_PEND_CAP * sizeof(FcbCand)) != cudaSuccess ||
        cudaMalloc(&d->redo, (size_t)(2 * DN_N + 1) * sizeof(int32_t)) != cudaSuccess ||
        cudaMalloc(&d->gtop, (size_t)DN_N * d->gt_stride * sizeof(unsigned)) != cudaSuccess ||
        cudaHostAlloc(&d->h_error, sizeof(int), cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer(&d->error_flag, d->h_error, 0) != cudaSuccess) {
        cudaGetLastError();
        fcb_dense_free(e);
        return fcb_fail(e, FCB_ERR_NOMEM, "fcb_load_shard: cannot allocate the batched-path scratch");
    }
    *d->h_error = 0;
    int32_t rc = encode_map(e, &d->map_x, s.vectors, (uint64_t)s.n_rows, (uint64_t)s.dim, DN_BLOCK_M);
    for (int i = 0; i < DN_N / 32 && rc == FCB_OK; ++i)
        rc = encode_map(e, &d->map_q[i], d->q_bf16, DN_N, (uint64_t)s.dim, 32 * (i + 1));
    for (int i = 0; i < DN_N / 32 && rc == FCB_OK; ++i)
        rc = encode_map(e, &d->map_qh[i], d->q_bf16, DN_N, (uint64_t)s.dim, 16 * (i + 1));
    const char *rig = getenv("FCB_DENSE_RIGOROUS");                 // "0": the statistical margin of round 1 (A/B runs only)
    d->rigorous = !(rig && rig[0] == '0');
    const char *one = getenv("FCB_DENSE_1CTA");                     // tuning / fallback switch
    d->two_cta = !(one && one[0] == '1') && (d->grid % 2 == 0);
    if (rc != FCB_OK) { fcb_dense_free(e); return rc; }
    const size_t smem = (size_t)DN_STAGES * DN_STAGE_BYTES + 4096;
    FCB_CUDA(e, cudaFuncSetAttribute(dense_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const size_t smem2 = (size_t)4 * DN2_STAGE_BYTES + 4096 + (size_t)96 * FCB_FUSED_CAND * sizeof(FcbCand);   // 4 stages + 96 lists >= 3 stages + 256 lists
    FCB_CUDA(e, cudaFuncSetAttribute(dense_topk2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    return FCB_OK;
}

// Runs the tensor-core pass + finalize for `batch` (<= DN_N) queries sharing filter `f`.
// `redo_out` receives the device pointer of the per-query "not certified" flags.
static float *g_debug_out = nullptr;
static int64_t g_debug_rows = 0;
static long long *g_dense_prof = nullptr;

// Diagnostic: per-role cycle counters of CTA 0 of the cta_group::2 kernel into `buf` (device int64[10]):
// producer total / waiting for a free stage; MMA issuer total / waiting for drained accumulators / waiting
// for a full stage; epilogue total / waiting for accumulators / scanning TMEM / folding candidates; super-tiles.
extern "C" int32_t fcb_debug_dense_prof(int64_t *buf) {
    g_dense_prof = reinterpret_cast<long long *>(buf);
    return FCB_OK;
}

// Diagnostic: make the next batched-path call also dump the raw tensor-core accumulators of the
// first `rows` rows of the segment into `out` (device float [rows, 256]); rows = 0 switches it off.
extern "C" int32_t fcb_debug_dense_dump(float *out, int64_t rows) {
    g_debug_out = rows > 0 ? out : nullptr;
    g_debug_rows = rows > 0 ? rows : 0;
    return FCB_OK;
}

int32_t fcb_search_dense(fcb_engine *e, const double *d_queries, const fcb_filter &f, int32_t batch, int32_t k,
                         uint64_t *d_out_ids, double *d_out_scores, int32_t *d_out_counts, cudaStream_t st,
                         const int32_t **redo_list_out, const int32_t **redo_count_out) {
    FcbDense *d = reinterpret_cast<FcbDense *>(e->dense);
    const fcb_shard &s = e->shard;
    if (!d) return fcb_fail(e, FCB_ERR_STATE, "fcb_search: batched path not initialised");
    dense_pack_queries_kernel<<<DN_N, 256, 0, st>>>(d_queries, batch, s.dim, d->q_bf16, d->slack);
    DenseParams p{};
    p.inv_norm = e->inv_norm; p.id_rank = s.id_rank; p.dates = s.dates; p.cats = s.categories; p.filter = f;
    p.n_rows = s.n_rows; p.row_base = s.row_base; p.dim = s.dim;
    p.n_queries = batch; p.n_eff = ((batch + 31) / 32) * 32;
    p.cta_top = d->cta_top; p.pend = d->pend; p.error_flag = d->error_flag;
    p.debug_out = g_debug_out; p.debug_rows = g_debug_rows; p.prof = g_dense_prof;
    p.gtop = d->gtop; p.gt_stride = d->gt_stride; p.slack = d->slack; p.thr_out = d->thr_out;
    FCB_CUDA(e, cudaMemsetAsync(d->gtop, 0, (size_t)DN_N * d->gt_stride * sizeof(unsigned), st));   // 0 orders below every float
    p.n_stages = p.n_eff <= 96 ? 4 : 3;                              // stages + 4 KB + lists (n_eff x 256 B) must fit 227 KB
    const size_t smem = (size_t)DN_STAGES * DN_STAGE_BYTES + 4096;
    if (d->two_cta) {
        const size_t smem2 = (size_t)p.n_stages * DN2_STAGE_BYTES + 4096 + (size_t)p.n_eff * FCB_FUSED_CAND * sizeof(FcbCand);
        dense_topk2_kernel<<<d->grid, DN2_THREADS, smem2, st>>>(d->map_x, d->map_qh[p.n_eff / 32 - 1], p);
    } else {
        dense_topk_kernel<<<d->grid, DN_THREADS, smem, st>>>(d->map_x, d->map_q[p.n_eff / 32 - 1], p);
    }
    FCB_CUDA(e, cudaGetLastError());
    DenseFinalParams fp{};
    fp.vectors = s.vectors; fp.ids = s.ids; fp.id_rank = s.id_rank; fp.queries = d_queries; fp.cta_top = d->cta_top;
    fp.n_lists = d->grid; fp.row_base = s.row_base; fp.dim = s.dim; fp.k = k;
    fp.out_ids = d_out_ids; fp.out_scores = d_out_scores; fp.out_counts = d_out_counts; fp.redo = d->redo;
    fp.rigorous = d->rigorous; fp.thr_out = d->thr_out; fp.gt_stride = d->gt_stride; fp.widened = d->widened;
    fp.redo_list = d->redo + DN_N; fp.redo_count = d->redo + 2 * DN_N;
    FCB_CUDA(e, cudaMemsetAsync(fp.redo_count, 0, sizeof(int32_t), st));
    dense_finalize_kernel<<<batch, 256, 0, st>>>(fp);
    FCB_CUDA(e, cudaGetLastError());
    e->last_launches += 3;
    *redo_list_out = d->redo + DN_N;
    *redo_count_out = d->redo + 2 * DN_N;
    return FCB_OK;
}

static int32_t count_flags(fcb_engine *e, const int32_t *d_flags, int32_t batch) {
    if (!e || !e->dense || batch < 1 || batch > DN_N) return -1;
    FcbDeviceGuard g(e->device);
    int32_t h[DN_N];
    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
    if (cudaMemcpy(h, d_flags, (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    int32_t n = 0;
    for (int i = 0; i < batch; ++i) n += h[i] != 0;
    return n;
}

// Measurement aid: how many queries of the last batched-path call could NOT be answered exactly from the tensor-core
// pass's lists and were therefore answered by the exact scan. Synchronises the device.
extern "C" int32_t fcb_last_dense_redo(fcb_engine *e, int32_t batch) {
    return e && e->dense ? count_flags(e, reinterpret_cast<FcbDense *>(e->dense)->redo, batch) : -1;
}

// Measurement aid: how many queries of the last batched-path call were answered by re-scoring MORE than the 32 best
// listed rows (every listed row inside the worst-case error window). Synchronises the device.
extern "C" int32_t fcb_last_dense_widened(fcb_engine *e, int32_t batch) {
    return e && e->dense ? count_flags(e, reinterpret_cast<FcbDense *>(e->dense)->widened, batch) : -1;
}
