// Segmented integer-cents group-by: the replacement for
//     transactions.groupby(plot_config.group_by)[plot_config.y_axis].sum()
// (reference tools/plot_tool.py:45 and :52) over device-resident columns, for the key
// (user, category, UTC calendar month) of BASELINE.json config 5. Exact int64 sums.
//
// HBM-bound target: 9 bytes per row (cents 4 + date 4 + category 1; the user is implicit in the
// CSR segment) + 9 bytes per output cell.
//
// Work = fixed tiles of AGG_TILE consecutive rows, grid-strided over a persistent grid. A tile is
// cut into pieces at user boundaries (binary search in seg_offsets). A piece accumulates into
// per-CTA shared-memory cells with native 32-bit atomics: each value is split as
// v = hi * 65536 + lo with lo in [-32768, 32767], so AGG_TILE <= 32768 rows can never overflow an
// int32 cell whatever the int32 input; `hi` is zero for |v| < 32768 and then costs nothing.
// Values with |v| < 2^18 (every realistic amount) take ONE atomic; only larger ones pay the split.
// Presence of a group (pandas lists observed groups only, even when their sum is 0) is a byte per
// cell written with plain stores (all writers store 1). Cells are flushed to the dense int64 output with one 64-bit global
// reduction per non-empty cell; very short pieces skip shared memory and reduce straight to global.
// The month of a row is a table lookup: day -> month index, built once per CTA in shared memory
// with the same civil-from-days arithmetic as the oracle.
#include <stdlib.h>

#include <type_traits>

#include "fcb_internal.cuh"

#define AGG_THREADS 256
#define AGG_TILE 4096
#define AGG_SMALL_PIECE 96
#ifndef AGG_BLOCK_TILES
#define AGG_BLOCK_TILES 16                 // consecutive tiles per binary search over the user offsets
#endif
#ifndef AGG_UNROLL
#define AGG_UNROLL 2
#endif
#ifndef AGG_CTAS
#define AGG_CTAS 4                         // resident CTAs per SM of aggregate_kernel (register cap 65536 / (256 * AGG_CTAS))
#endif
constexpr int kAggUnroll = AGG_UNROLL;     // (#pragma unroll takes a constant expression, not a macro)
#define AGG_MAX_MONTHS 240
#define AGG_MAX_CELLS 4096

struct AggParams {
    const int64_t *seg_offsets;
    const int32_t *dates;
    const uint8_t *cats;
    const int32_t *cents;
    int64_t seg_begin, seg_end;
    int64_t row_base, n_rows;
    int32_t month_lo, n_months, n_categories;
    int32_t day0, n_days;                  // table covers days [day0, day0 + n_days)
    int64_t *sums;                         // [segs, n_categories, n_months]
    uint8_t *present;                      // same shape
    unsigned long long *dropped;           // rows outside the category / month window
    int64_t small_piece;                   // pieces shorter than this reduce straight to the global cells
    // the window in 32-bit seconds, for the date form of the warp-table kernel (dates are int32): a date d is inside iff
    // (uint32)d - (uint32)base32 < window32 — the window clipped to the int32 range, so the wrap-around cannot alias — and
    // its day slot is ((uint32)d - base32 + shift32) / 86400 (shift32 != 0 only for a window that starts before 1901)
    int32_t base32;
    uint32_t window32, shift32;
    const uint16_t *keys;                  // group keys of the rows (keyed form of aggregate_warp_kernel), see "group keys" below
    uint32_t key_mo_off;                   // month_lo - key base: month slot of a row = (key's month field) - key_mo_off
    uint32_t key_stride;                   // words per category in a warp's pos / neg tables: n_months | 1 (odd, see agk_row)
    uint32_t key_neg_off, key_hi_off;      // byte offsets of the neg table and of the hi words in a warp's table
    uint32_t key_step_mo, key_step_w4, key_wrap_w4;   // the flush loop's walk from cell c to c + 32: month slot += key_step_mo, table
                                           //   byte offset += key_step_w4, and on a month wrap -= n_months, += key_wrap_w4
    int split;                             // any-order data, two launches: 1 = this launch skips users < AGS_TABLE_USER rows (the
                                           // atomic kernel), 2 = it handles ONLY them (the warp-table path of the ordered kernel)
};

// months since 1970-01 of a day number (days since 1970-01-01), proleptic Gregorian
__host__ __device__ inline int32_t month_of_day(int32_t days) {
    const int32_t z = days + 719468;
    const int32_t era = (z >= 0 ? z : z - 146096) / 146097;
    const int32_t doe = z - era * 146097;
    const int32_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
    int32_t y = yoe + era * 400;
    const int32_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
    const int32_t mp = (5 * doy + 2) / 153;
    const int32_t m = mp < 10 ? mp + 3 : mp - 9;
    if (m <= 2) ++y;
    return (y - 1970) * 12 + (m - 1);
}

// first day (days since epoch) of month index mi
static int32_t first_day_of_month(int32_t mi) {
    int32_t y = 1970 + (mi >= 0 ? mi / 12 : -((-mi + 11) / 12));
    int32_t m = mi - (y - 1970) * 12 + 1;                  // 1..12
    // days-from-civil
    y -= m <= 2;
    const int32_t era = (y >= 0 ? y : y - 399) / 400;
    const int32_t yoe = y - era * 400;
    const int32_t doy = (153 * (m > 2 ? m - 3 : m + 9) + 2) / 5;
    const int32_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + doe - 719468;
}

__device__ __forceinline__ int32_t floor_day(int32_t ts) {
    int32_t d = ts / 86400;
    if (ts - d * 86400 < 0) --d;
    return d;
}

// month slot of a date inside the window, or 0xffff: one 64-bit subtract + compare against the window, an unsigned
// division by a constant and a table lookup (cheaper than flooring a signed day number first)
__device__ __forceinline__ unsigned agg_month(const AggParams &p, int32_t date, const uint16_t *month_tab) {
    const uint64_t off = (uint64_t)((int64_t)date - (int64_t)p.day0 * 86400);
    return off < (uint64_t)p.n_days * 86400u ? month_tab[(uint32_t)off / 86400u] : 0xffffu;
}

// predicated shared-memory reduction / byte store on 32-bit shared-window addresses (no branch, no generic -> shared
// conversion per row)
__device__ __forceinline__ void agg_red_shared(uint32_t addr, int32_t v, bool pred) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q red.shared.add.s32 [%0], %1;\n\t}"
                 ::"r"(addr), "r"(v), "r"((int)pred) : "memory");
}
__device__ __forceinline__ void agg_st_shared_u8(uint32_t addr, bool pred) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %1, 0;\n\t@q st.shared.u8 [%0], 1;\n\t}" ::"r"(addr), "r"((int)pred) : "memory");
}

// shared-window addresses of one set of cells: lo [C] int32, hi [C] int32, present [C] uint8
struct AggCells { uint32_t lo, hi, present; };
__device__ __forceinline__ AggCells agg_cells(const int32_t *c_lo, const int32_t *c_hi, const uint8_t *c_present) {
    return AggCells{(uint32_t)__cvta_generic_to_shared(c_lo), (uint32_t)__cvta_generic_to_shared(c_hi),
                    (uint32_t)__cvta_generic_to_shared(c_present)};
}

// one row into shared cells: |v| < 2^18 -> ONE native 32-bit shared reduction (4096 such values cannot overflow an int32
// cell); larger values are split v = hi * 65536 + lo16 so that no int32 input can overflow either cell (rare second reduction)
__device__ __forceinline__ void agg_cell_add(const AggCells &cells, bool valid, int cell, int32_t v) {
    const bool large = !((uint32_t)v + ((1u << 18) - 1u) < (1u << 19) - 1u);   // not in (-2^18, 2^18)
    const int32_t lo16 = (int32_t)(int16_t)(v & 0xffff);
    agg_red_shared(cells.lo + 4u * (uint32_t)cell, large ? lo16 : v, valid);
    if (valid && large) agg_red_shared(cells.hi + 4u * (uint32_t)cell, (int32_t)(((int64_t)v - lo16) >> 16), true);   // rare
    agg_st_shared_u8(cells.present + (uint32_t)cell, valid);          // plain byte store: every writer stores the same value
}

// one row into the CTA's shared cells (big pieces: SMALL = false) or straight to the global cells (tiny pieces)
template <bool SMALL>
__device__ __forceinline__ void agg_row(const AggParams &p, int32_t v, unsigned cat, int32_t date, const AggCells &cells,
                                        const uint16_t *month_tab, int64_t *out_s, uint8_t *out_p, unsigned &dropped) {
    const unsigned mo = agg_month(p, date, month_tab);
    const bool valid = cat < (unsigned)p.n_categories && mo < (unsigned)p.n_months;
    dropped += valid ? 0u : 1u;
    const int cell = valid ? (int)(cat * p.n_months + mo) : 0;
    if (SMALL) {
        if (valid) {
            atomicAdd(reinterpret_cast<unsigned long long *>(out_s + cell), (unsigned long long)(long long)v);
            out_p[cell] = 1;
        }
        return;
    }
    agg_cell_add(cells, valid, cell, v);
}

#ifndef AGS_ROUND
#define AGS_ROUND 8                        // rows per lane and round of the small-user path
#endif
#ifndef AGS_TABLE_USER
#define AGS_TABLE_USER 8192                // users shorter than this: per-warp [category, month] table
#endif

// One small user's piece [la, lb) (local rows) on ONE warp: a private [category, month] table in shared memory (t_lo, t_hi,
// t_p: all zero on entry and on exit) fed with 32-bit shared atomics — one per row for |v| < 2^18, a lo/hi split otherwise; a
// piece has at most 4096 rows, so no cell can overflow. (This path keeps the plain form of the row update — signed day number,
// generic-pointer atomics: the leaner form of aggregate_kernel's agg_row costs aggregate_sorted_kernel, which inlines this
// function at its 64-register cap, spills and 8 % of its time.) AGS_ROUND x 32 rows per round, all column loads of a lane in flight
// before the first is used (one row per lane at a time streams ~1 TB/s across the GPU, latency-bound). The present cells go
// out once: plain stores when the piece is the whole user (`whole`), 64-bit reductions otherwise.
__device__ __forceinline__ void agg_warp_table(const AggParams &p, int64_t la, int64_t lb, bool whole, int64_t user, int32_t *t_lo,
                                               int32_t *t_hi, uint8_t *t_p, const uint16_t *month_tab, int lane, unsigned &dropped) {
    const int C = p.n_categories * p.n_months;
    for (int64_t r0 = la + lane; r0 < lb; r0 += 32 * AGS_ROUND) {
        int32_t vv[AGS_ROUND], dd[AGS_ROUND];
        unsigned cc[AGS_ROUND];
#pragma unroll
        for (int j = 0; j < AGS_ROUND; ++j) {
            const int64_t r = r0 + 32 * j;
            const bool in = r < lb;
            vv[j] = in ? p.cents[r] : 0; dd[j] = in ? p.dates[r] : 0; cc[j] = in ? (unsigned)p.cats[r] : 0x100u;
        }
#pragma unroll
        for (int j = 0; j < AGS_ROUND; ++j) {
            if (cc[j] == 0x100u) continue;                       // past the piece
            const int32_t di = floor_day(dd[j]) - p.day0;
            const unsigned mo = (unsigned)di < (unsigned)p.n_days ? month_tab[di] : 0xffffu;
            if (cc[j] >= (unsigned)p.n_categories || mo >= (unsigned)p.n_months) { ++dropped; continue; }
            const int cell = cc[j] * p.n_months + mo;
            const int32_t v = vv[j];
            if (v > -(1 << 18) && v < (1 << 18)) {
                atomicAdd(&t_lo[cell], v);
            } else {
                const int32_t lo16 = (int32_t)(int16_t)(v & 0xffff);
                atomicAdd(&t_lo[cell], lo16);
                atomicAdd(&t_hi[cell], (int32_t)(((int64_t)v - lo16) >> 16));
            }
#ifdef AGG_PRESENT_CHECK
            if (!t_p[cell]) t_p[cell] = 1;
#else
            t_p[cell] = 1;
#endif
        }
    }
    __syncwarp();
    int64_t *out_s = p.sums + (user - p.seg_begin) * (int64_t)C;
    uint8_t *out_p = p.present + (user - p.seg_begin) * (int64_t)C;
    for (int c = lane; c < C; c += 32) {
        if (t_p[c]) {
            const long long total = (long long)t_hi[c] * 65536 + t_lo[c];
            if (whole) out_s[c] = total;
            else atomicAdd(reinterpret_cast<unsigned long long *>(out_s + c), (unsigned long long)total);
            out_p[c] = 1;
            t_lo[c] = 0; t_hi[c] = 0; t_p[c] = 0;
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(AGG_THREADS, AGG_CTAS) aggregate_kernel(const __grid_constant__ AggParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int C = p.n_categories * p.n_months;
    // TWO sets of cells, used by alternate pieces: a piece is accumulated into one set, ONE barrier, then every thread flushes
    // its cells of that set and goes straight on to the next piece in the other set — by the time a set is accumulated into
    // again, the barrier of the piece in between has seen every thread finish flushing it.
    int32_t *c_lo = reinterpret_cast<int32_t *>(smem_raw);             // [2][C]
    int32_t *c_hi = c_lo + 2 * C;                                      // [2][C]
    uint16_t *month_tab = reinterpret_cast<uint16_t *>(c_hi + 2 * C);  // [n_days]
    uint8_t *c_present = reinterpret_cast<uint8_t *>(month_tab + ((p.n_days + 1) & ~1));   // [2][C]
    __shared__ int64_t s_piece[2];

    const int tid = threadIdx.x;
    const AggCells cells_0 = agg_cells(c_lo, c_hi, c_present);
    int cur = 0;                                                       // the set the next large piece accumulates into
    for (int i = tid; i < p.n_days; i += AGG_THREADS) month_tab[i] = (uint16_t)(month_of_day(p.day0 + i) - p.month_lo);
    for (int i = tid; i < 2 * C; i += AGG_THREADS) { c_lo[i] = 0; c_hi[i] = 0; c_present[i] = 0; }
    __syncthreads();

    int64_t R0 = p.seg_offsets[p.seg_begin], R1 = p.seg_offsets[p.seg_end];
    if (R0 < p.row_base) R0 = p.row_base;
    if (R1 > p.row_base + p.n_rows) R1 = p.row_base + p.n_rows;
    const int64_t n_tiles = R1 > R0 ? (R1 - R0 + AGG_TILE - 1) / AGG_TILE : 0;
    unsigned dropped = 0;
    const int4 *cents4 = reinterpret_cast<const int4 *>(p.cents);
    const int4 *dates4 = reinterpret_cast<const int4 *>(p.dates);
    const uchar4 *cats4 = reinterpret_cast<const uchar4 *>(p.cats);

    // Tiles are dealt in BLOCKS of AGG_BLOCK_TILES consecutive tiles: the user of a block's first row costs one binary search
    // over seg_offsets (17 dependent global loads, ~8 us during which the whole CTA waits at a barrier — the round-2 ncu
    // capture showed the shared-atomic pipe 11 % busy and the warps stalled on exactly this: barrier + long scoreboard);
    // inside the block the user index only moves forward.
    const int64_t n_blocks = (n_tiles + AGG_BLOCK_TILES - 1) / AGG_BLOCK_TILES;
    for (int64_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        const int64_t t0 = blk * AGG_BLOCK_TILES, t1 = t0 + AGG_BLOCK_TILES < n_tiles ? t0 + AGG_BLOCK_TILES : n_tiles;
        if (tid == 0) {                                    // first user whose segment ends after the block's first row
            const int64_t a0 = R0 + t0 * AGG_TILE;
            int64_t lo = p.seg_begin, hi = p.seg_end - 1;
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if (p.seg_offsets[mid + 1] > a0) hi = mid; else lo = mid + 1;
            }
            s_piece[0] = lo;
        }
        __syncthreads();
        int64_t u = s_piece[0];
        __syncthreads();
      for (int64_t t = t0; t < t1; ++t) {
        const int64_t a = R0 + t * AGG_TILE;
        const int64_t b = a + AGG_TILE < R1 ? a + AGG_TILE : R1;
        int64_t pa = a;
        while (pa < b) {
            int64_t uend = p.seg_offsets[u + 1];
            while (uend <= pa) { ++u; uend = p.seg_offsets[u + 1]; }   // skip empty users
            const int64_t pb = uend < b ? uend : b;
            if (p.split == 1 && uend - p.seg_offsets[u] < AGS_TABLE_USER) { pa = pb; continue; }   // the other launch's user
            int64_t *out_s = p.sums + (u - p.seg_begin) * (int64_t)C;
            uint8_t *out_p = p.present + (u - p.seg_begin) * (int64_t)C;
            const bool small = pb - pa < p.small_piece;
            // local row range of the piece; quads aligned on the LOCAL row index so the 128-bit loads are aligned
            const int64_t la = pa - p.row_base, lb = pb - p.row_base;
            const int64_t q0 = la >> 2, q1 = (lb + 3) >> 2;            // quads [q0, q1) cover [la, lb)
            // one quad loop per kind of piece, so that the row path carries no per-row "small piece?" test
            const AggCells cells{cells_0.lo + (uint32_t)(cur * C) * 4u, cells_0.hi + (uint32_t)(cur * C) * 4u,
                                 cells_0.present + (uint32_t)(cur * C)};   // (computed: an indexed array would live in local memory)
            auto quads = [&](auto small_tag) {
                constexpr bool SMALL = decltype(small_tag)::value;
#pragma unroll kAggUnroll
                for (int64_t q = q0 + tid; q < q1; q += AGG_THREADS) {
                    const int64_t r = q << 2;
                    if (r >= la && r + 4 <= lb) {                       // interior quad: three vector loads
                        const int4 v = __ldg(cents4 + q);
                        const int4 d = __ldg(dates4 + q);
                        const uchar4 c = __ldg(cats4 + q);
                        agg_row<SMALL>(p, v.x, c.x, d.x, cells, month_tab, out_s, out_p, dropped);
                        agg_row<SMALL>(p, v.y, c.y, d.y, cells, month_tab, out_s, out_p, dropped);
                        agg_row<SMALL>(p, v.z, c.z, d.z, cells, month_tab, out_s, out_p, dropped);
                        agg_row<SMALL>(p, v.w, c.w, d.w, cells, month_tab, out_s, out_p, dropped);
                    } else {                                            // ragged head / tail of the piece
                        for (int64_t rr = r > la ? r : la; rr < r + 4 && rr < lb; ++rr)
                            agg_row<SMALL>(p, p.cents[rr], p.cats[rr], p.dates[rr], cells, month_tab, out_s, out_p, dropped);
                    }
                }
            };
            if (small) quads(std::true_type{}); else quads(std::false_type{});
            if (!small) {
                __syncthreads();
                int32_t *f_lo = c_lo + cur * C, *f_hi = c_hi + cur * C;
                uint8_t *f_p = c_present + cur * C;
                for (int c = tid; c < C; c += AGG_THREADS) {
                    if (f_p[c]) {
                        const long long s = (long long)f_hi[c] * 65536 + f_lo[c];
                        atomicAdd(reinterpret_cast<unsigned long long *>(out_s + c), (unsigned long long)s);
                        out_p[c] = 1;
                        f_lo[c] = 0; f_hi[c] = 0; f_p[c] = 0;
                    }
                }
                cur ^= 1;                                              // no second barrier: see the note at the cell sets
            }
            pa = pb;
        }
      }
    }
    if (dropped) atomicAdd(p.dropped, (unsigned long long)dropped);
}

// ---- the warp-table kernel (default for every table) and group keys: the month of every row resolved ONCE, at load -------
// plot_tool groups by (category, calendar month of the date). The month of a row never changes, so fcb_load_shard derives a
// 16-bit GROUP KEY per row and keeps it next to the columns (2 bytes per row, library-owned):
//     bit 15      the row's amount is "rare": zero, or |v| >= 2^18 (see below)
//     bits 14..9  category (0..63)
//     bits 8..0   month - key_base (0..510); 511 = none
// 0x7fff stands for a category >= 64 or a month outside [key_base, key_base + 511). A query whose month window lies inside
// the key range then reads cents + key (6 bytes per row instead of 9) and spends no arithmetic on calendars. Rows with key
// 0x7fff are outside every such window, so they count as dropped exactly as the date path counts them. Windows that reach
// outside the key range, and columns that come without keys (fcb_aggregate_columns), take the DATE form of the same kernel
// (aggregate_warp_kernel<false>: month through a day -> month table, rare test from the value). The two kernels above are
// round 1's; they stay behind FCB_AGG_KERNEL as independent implementations for the tests.
#define AGK_MONTH_BITS 9
#define AGK_MONTHS ((1 << AGK_MONTH_BITS) - 1)          // representable months; the all-ones month field means "none"
#define AGK_NONE 0x7fffu
#ifndef AGK_TILE
#define AGK_TILE 8192                      // rows per piece at most: (2^18 - 1) * 8192 < 2^31, a 32-bit cell cannot overflow
#endif
#ifndef AGK_BLOCK_TILES
#define AGK_BLOCK_TILES 4                  // consecutive tiles one warp takes per binary search over the users
#endif
#ifndef AGK_QUADS
#define AGK_QUADS 4                        // quads (of 4 rows) per lane whose loads are in flight together
#endif
#ifndef AGK_CTAS
#define AGK_CTAS 4
#endif

// Every WARP owns a private table in shared memory and a stream of pieces (tile x user); nothing in the kernel is CTA-wide:
// no barrier, no shared piece descriptor. Per cell: a (pos, neg) pair of 32-bit sums of the positive / negative values —
// for values below 2^18 in magnitude neither can cancel to zero nor overflow within a piece, so "pos | neg != 0" IS the
// presence of the cell and the common row costs ONE shared reduction and no presence store — plus a hi word and a presence
// byte for the rare rows. The row update is branch-free and unconditional (a predicated reduction compiles to a branch
// around it): a row outside the window adds 1 to its lane's own drop counter instead of v to a cell. Rare rows — zero
// (presence only) and |v| >= 2^18 (split v = hi * 65536 + lo16) — carry a flag in their key (date form: tested from the value) and are handled a second time,
// quad by quad: the large ones add lo16 - v to the word the common path added v to (modular arithmetic: the word ends up
// holding lo16) and their hi part to the hi word.
__device__ __forceinline__ bool agk_is_rare(int32_t v) { return !((uint32_t)v + ((1u << 18) - 1u) < (1u << 19) - 1u && v != 0); }

// table of one warp: pos [n_categories][key_stride] int32 | neg, same shape, at key_neg_off | hi [C] int32 at key_hi_off |
// present [C] bytes | drop counters [32] uint32. Bank layout: in a time-ordered log the 128 rows of a warp round share
// their month, so their words differ by category and sign only. key_stride is ODD — 16 categories land in 16 different
// banks (the natural stride, 24 words for two years, would give 4) — and the neg table starts 16 banks (mod 32) after the
// pos table, so the negative values use the other 16.
// `mo`: month slot inside the window (anything >= n_months: outside)
__device__ __forceinline__ void agk_row(const AggParams &p, int32_t v, uint32_t mo, uint32_t cat, uint32_t tab, uint32_t drop_word) {
    const bool valid = mo < (uint32_t)p.n_months && cat < (uint32_t)p.n_categories;
    const uint32_t addr = tab + 4u * (cat * p.key_stride + mo) + ((uint32_t)v >> 31) * p.key_neg_off;
    asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(valid ? addr : drop_word), "r"(valid ? v : 1) : "memory");
}

// second visit of a row that may be rare (see above): nothing to do for rows outside the window or with a common amount
__device__ __noinline__ void agk_rare_cell(const AggParams &p, int32_t v, uint32_t mo, uint32_t cat, uint32_t tab, uint32_t C) {
    if (!agk_is_rare(v) || !(mo < (uint32_t)p.n_months && cat < (uint32_t)p.n_categories)) return;
    const uint32_t cell = cat * (uint32_t)p.n_months + mo;
    if (v != 0) {
        const int32_t lo16 = (int32_t)(int16_t)(v & 0xffff);
        agg_red_shared(tab + 4u * (cat * p.key_stride + mo) + ((uint32_t)v >> 31) * p.key_neg_off,
                       (int32_t)((uint32_t)lo16 - (uint32_t)v), true);
        agg_red_shared(tab + p.key_hi_off + 4u * cell, (int32_t)(((int64_t)v - lo16) >> 16), true);
    }
    agg_st_shared_u8(tab + p.key_hi_off + 4u * C + cell, true);
}

// keyed form: month slot and category of a 16-bit key
__device__ __forceinline__ uint32_t agk_key_mo(const AggParams &p, uint32_t key) { return (key & (uint32_t)AGK_MONTHS) - p.key_mo_off; }
__device__ __forceinline__ uint32_t agk_key_cat(uint32_t key) { return (key >> AGK_MONTH_BITS) & 63u; }
// date form: month slot of a date — one 32-bit subtract + compare against the window (see AggParams), an unsigned division by a
// constant and a lookup in the CTA's day -> month table, whose entry [n_days] is 0xffff (outside)
__device__ __forceinline__ uint32_t agk_date_mo(const AggParams &p, int32_t date, uint32_t mtab) {
    const uint32_t off = (uint32_t)date - (uint32_t)p.base32;
    const uint32_t day = off < p.window32 ? (off + p.shift32) / 86400u : (uint32_t)p.n_days;
    unsigned short mo;
    asm("ld.shared.u16 %0, [%1];" : "=h"(mo) : "r"(mtab + 2u * day));
    return mo;
}

#ifndef AGD_QUADS
#define AGD_QUADS 2                        // date form: quads per lane and round (9 registers of columns each; 3: +1 %, 4: spills)
#endif

// the four rows of a quad, then — rarely — a second visit of the rare ones. Keyed form: v = amounts, k = two packed pairs of
// keys (bit 15 of a key = rare amount). Date form: v, d = dates, c = four category bytes.
template <bool KEYED>
__device__ __forceinline__ void agk_decode(const AggParams &p, int i, const uint2 &k, const int4 &d, uint32_t c, uint32_t mtab,
                                           uint32_t &mo, uint32_t &cat) {
    if (KEYED) {
        const uint32_t key = (i & 2 ? k.y : k.x) >> (i & 1 ? 16 : 0);
        mo = agk_key_mo(p, key); cat = agk_key_cat(key);
    } else {
        mo = agk_date_mo(p, i == 0 ? d.x : i == 1 ? d.y : i == 2 ? d.z : d.w, mtab);
        cat = (c >> (8 * i)) & 0xffu;
    }
}
template <bool KEYED>
__device__ __forceinline__ void agk_quad(const AggParams &p, const int4 &v, const uint2 &k, const int4 &d, uint32_t c, uint32_t tab,
                                         uint32_t mtab, uint32_t drop_word, uint32_t C) {
    const int32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        uint32_t mo, cat;
        agk_decode<KEYED>(p, i, k, d, c, mtab, mo, cat);
        agk_row(p, vv[i], mo, cat, tab, drop_word);
    }
    const bool rare = KEYED ? ((k.x | k.y) & 0x80008000u) != 0
                            : (agk_is_rare(v.x) || agk_is_rare(v.y) || agk_is_rare(v.z) || agk_is_rare(v.w));
    if (rare) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            uint32_t mo, cat;
            agk_decode<KEYED>(p, i, k, d, c, mtab, mo, cat);
            agk_rare_cell(p, vv[i], mo, cat, tab, C);
        }
    }
}

// KEYED: cents + group keys (6 B/row). !KEYED: cents + dates + categories (9 B/row), the month of a row through the CTA's
// day -> month table: the same warp-private tables, the same row update, for columns that come without keys and for windows
// outside the keys' range.
template <bool KEYED>
__global__ void __launch_bounds__(AGG_THREADS, AGK_CTAS) aggregate_warp_kernel(const __grid_constant__ AggParams p) {
    constexpr int QUADS = KEYED ? AGK_QUADS : AGD_QUADS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t C = (uint32_t)(p.n_categories * p.n_months);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t mtab_bytes = KEYED ? 0u : ((((uint32_t)p.n_days + 1u) * 2u + 15u) & ~15u);   // [n_days + 1] uint16, last = 0xffff
    const uint32_t mtab = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t tab_bytes = ((p.key_hi_off + 5u * C + 15u) & ~15u) + 128u;   // see agk_row
    unsigned char *tab_g = smem_raw + mtab_bytes + (size_t)warp * tab_bytes;
    const uint32_t tab = (uint32_t)__cvta_generic_to_shared(tab_g);
    const uint32_t drop_word = tab + tab_bytes - 128u + 4u * lane;      // this lane's count of rows outside the window
    for (uint32_t i = lane; i < tab_bytes / 4; i += 32) reinterpret_cast<uint32_t *>(tab_g)[i] = 0;
    if (!KEYED) {
        uint16_t *mt = reinterpret_cast<uint16_t *>(smem_raw);
        for (int i = threadIdx.x; i <= p.n_days; i += blockDim.x)
            mt[i] = i < p.n_days ? (uint16_t)(month_of_day(p.day0 + i) - p.month_lo) : (uint16_t)0xffff;
        __syncthreads();                                               // the only CTA-wide step of the kernel
    }
    __syncwarp();

    int64_t R0 = p.seg_offsets[p.seg_begin], R1 = p.seg_offsets[p.seg_end];
    if (R0 < p.row_base) R0 = p.row_base;
    if (R1 > p.row_base + p.n_rows) R1 = p.row_base + p.n_rows;
    const int64_t n_tiles = R1 > R0 ? (R1 - R0 + AGK_TILE - 1) / AGK_TILE : 0;
    const int64_t n_blocks = (n_tiles + AGK_BLOCK_TILES - 1) / AGK_BLOCK_TILES;
    const int warps = blockDim.x >> 5;                                  // (fewer than 8 when the tables are large)
    const int64_t n_warps = (int64_t)gridDim.x * warps;

    for (int64_t blk = (int64_t)blockIdx.x * warps + warp; blk < n_blocks; blk += n_warps) {
        const int64_t a0 = R0 + blk * AGK_BLOCK_TILES * AGK_TILE;
        const int64_t b0 = a0 + (int64_t)AGK_BLOCK_TILES * AGK_TILE < R1 ? a0 + (int64_t)AGK_BLOCK_TILES * AGK_TILE : R1;
        int64_t u;                                                  // first user whose segment ends after a0 (every lane the same)
        {
            int64_t lo = p.seg_begin, hi = p.seg_end - 1;
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if (p.seg_offsets[mid + 1] > a0) hi = mid; else lo = mid + 1;
            }
            u = lo;
        }
        int64_t pa = a0;
        int64_t ubeg = p.seg_offsets[u], uend = p.seg_offsets[u + 1];
        while (pa < b0) {
            while (uend <= pa) { ++u; ubeg = uend; uend = p.seg_offsets[u + 1]; }      // skip empty users
            const int64_t tile_end = R0 + ((pa - R0) / AGK_TILE + 1) * AGK_TILE;
            int64_t pb = uend < b0 ? uend : b0;
            if (tile_end < pb) pb = tile_end;
            const int64_t la = pa - p.row_base, lb = pb - p.row_base;           // local rows of the piece
            // rows that do not fill an aligned quad (at most 3 at either end; quads are aligned on the LOCAL row index, where
            // the 128-bit loads are) go first, one per lane
            const int64_t qa = (la + 3) >> 2, qb = lb >> 2;
            const int64_t head_end = qa * 4 < lb ? qa * 4 : lb;
            const int n_head = (int)(head_end - la), n_tail = qb >= qa ? (int)(lb - qb * 4) : 0;
            if (lane < n_head + n_tail) {
                const int64_t r = lane < n_head ? la + lane : qb * 4 + (lane - n_head);
                const int32_t v = p.cents[r];
                uint32_t mo, cat;
                if (KEYED) { const uint32_t key = p.keys[r]; mo = agk_key_mo(p, key); cat = agk_key_cat(key); }
                else { mo = agk_date_mo(p, p.dates[r], mtab); cat = p.cats[r]; }
                agk_row(p, v, mo, cat, tab, drop_word);
                agk_rare_cell(p, v, mo, cat, tab, C);
            }
            if (qb > qa) {
                const int4 *cv = reinterpret_cast<const int4 *>(p.cents) + qa;
                const uint2 *kv = reinterpret_cast<const uint2 *>(p.keys) + qa;
                const int4 *dv = reinterpret_cast<const int4 *>(p.dates) + qa;
                const uint32_t *tv = reinterpret_cast<const uint32_t *>(p.cats) + qa;
                const int nq = (int)(qb - qa);
                // QUADS quads per lane and round, all their loads in flight before the first row is used (issuing the next
                // round's loads ahead of the current round's rows, from a second register set, measured 8 % SLOWER)
                for (int q = lane; q < nq; q += 32 * QUADS) {
                    int4 v[QUADS], d[QUADS];
                    uint2 k[QUADS];
                    uint32_t c[QUADS];
                    const bool full = q + 32 * (QUADS - 1) < nq;            // every round but a lane's last: no per-quad tests
                    if (full) {
#pragma unroll
                        for (int j = 0; j < QUADS; ++j) {
                            v[j] = __ldg(cv + q + 32 * j);
                            if (KEYED) k[j] = __ldg(kv + q + 32 * j);
                            else { d[j] = __ldg(dv + q + 32 * j); c[j] = __ldg(tv + q + 32 * j); }
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < QUADS; ++j) {
                            const bool in = q + 32 * j < nq;
                            v[j] = in ? __ldg(cv + q + 32 * j) : make_int4(0, 0, 0, 0);
                            if (KEYED) k[j] = in ? __ldg(kv + q + 32 * j) : make_uint2(0u, 0u);
                            else { d[j] = in ? __ldg(dv + q + 32 * j) : make_int4(0, 0, 0, 0); c[j] = in ? __ldg(tv + q + 32 * j) : 0u; }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < QUADS; ++j) {
                        if (!full && q + 32 * j >= nq) break;
                        agk_quad<KEYED>(p, v[j], k[j], d[j], c[j], tab, mtab, drop_word, C);
                    }
                }
            }
            __syncwarp();
            // the piece's cells go out once: plain stores when the piece is the whole user (nobody else touches its cells,
            // which the launch zeroed), 64-bit reductions otherwise; the table is left all zero
            {
                const bool whole = pa == ubeg && pb == uend;
                // this lane's cells lane, lane + 32, ...: pointers and table slots advance by constant steps (no per-cell
                // multiplications or divisions)
                int64_t *os = p.sums + (u - p.seg_begin) * (int64_t)C + lane;
                uint8_t *op = p.present + (u - p.seg_begin) * (int64_t)C + lane;
                uint32_t mo = (uint32_t)lane % (uint32_t)p.n_months;
                uint32_t w4 = tab + 4u * (((uint32_t)lane / (uint32_t)p.n_months) * p.key_stride + mo);   // pos word of the cell
                uint32_t h4 = tab + p.key_hi_off + 4u * lane, pb1 = tab + p.key_hi_off + 4u * C + lane;
                for (uint32_t c = lane; c < C; c += 32) {
                    int32_t ps, ng, hi;
                    uint32_t pr;
                    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(ps) : "r"(w4));
                    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(ng) : "r"(w4 + p.key_neg_off));
                    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(hi) : "r"(h4));
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(pr) : "r"(pb1));
                    if ((ps | ng | hi | (int32_t)pr) != 0) {
                        const long long s = (long long)hi * 65536 + ps + ng;
                        if (whole) *os = s;
                        else atomicAdd(reinterpret_cast<unsigned long long *>(os), (unsigned long long)s);
                        *op = 1;
                        asm volatile("st.shared.b32 [%0], %1;" ::"r"(w4), "r"(0) : "memory");
                        asm volatile("st.shared.b32 [%0], %1;" ::"r"(w4 + p.key_neg_off), "r"(0) : "memory");
                        asm volatile("st.shared.b32 [%0], %1;" ::"r"(h4), "r"(0) : "memory");
                        asm volatile("st.shared.u8 [%0], %1;" ::"r"(pb1), "r"(0) : "memory");
                    }
                    os += 32; op += 32; h4 += 128u; pb1 += 32u;
                    mo += p.key_step_mo; w4 += p.key_step_w4;                    // cell c + 32
                    if (mo >= (uint32_t)p.n_months) { mo -= (uint32_t)p.n_months; w4 += p.key_wrap_w4; }
                }
            }
            __syncwarp();
            pa = pb;
        }
    }
    __syncwarp();
    unsigned dropped = reinterpret_cast<const uint32_t *>(tab_g + tab_bytes - 128u)[lane];
#pragma unroll
    for (int o = 16; o; o >>= 1) dropped += __shfl_xor_sync(0xffffffffu, dropped, o);
    if (lane == 0 && dropped) atomicAdd(p.dropped, (unsigned long long)dropped);
}

// month range of a shard's dates: [0] = min, [1] = max month index (months since 1970-01)
__global__ void __launch_bounds__(256) agk_month_range_kernel(const int32_t *__restrict__ dates, int64_t n_rows, int *range) {
    int lo = INT32_MAX, hi = INT32_MIN;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
        const int m = month_of_day(floor_day(dates[r]));
        lo = m < lo ? m : lo; hi = m > hi ? m : hi;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const int l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
        lo = l2 < lo ? l2 : lo; hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0 && lo <= hi) { atomicMin(range, lo); atomicMax(range + 1, hi); }
}

__global__ void __launch_bounds__(256) agk_build_keys_kernel(const int32_t *__restrict__ dates, const uint8_t *__restrict__ cats,
                                                             const int32_t *__restrict__ cents, int64_t n_rows, int32_t key_base,
                                                             uint16_t *__restrict__ keys) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
        const int64_t m = (int64_t)month_of_day(floor_day(dates[r])) - key_base;
        const unsigned c = cats[r];
        keys[r] = (m >= 0 && m < AGK_MONTHS && c < 64u)
                      ? (uint16_t)((agk_is_rare(cents[r]) ? 0x8000u : 0u) | (c << AGK_MONTH_BITS) | (unsigned)m) : (uint16_t)AGK_NONE;
    }
}

// Derives the key column of [n_rows] rows into `keys` (device, 16-byte aligned, n_rows * 2 bytes rounded up to 16). The key
// range ends at the newest month of the data: rows older than 510 months before it get the "none" key, and a window that
// reaches back there takes the date kernels. Waits for `st` (the month range of the data comes back to the host).
static int32_t agk_build(fcb_engine *e, int num_sms, const int32_t *dates, const uint8_t *cats, const int32_t *cents, int64_t n_rows,
                         uint16_t *keys, int32_t *key_base, cudaStream_t st) {
    int *d_range = nullptr;
    FCB_CUDA(e, cudaMalloc(&d_range, 2 * sizeof(int)));
    const int init[2] = {INT32_MAX, INT32_MIN};
    int h_range[2] = {0, 0};
    cudaError_t rc = cudaMemcpyAsync(d_range, init, sizeof init, cudaMemcpyHostToDevice, st);
    if (rc == cudaSuccess) {
        agk_month_range_kernel<<<num_sms * 8, 256, 0, st>>>(dates, n_rows, d_range);
        rc = cudaMemcpyAsync(h_range, d_range, sizeof h_range, cudaMemcpyDeviceToHost, st);
    }
    if (rc == cudaSuccess) rc = cudaStreamSynchronize(st);
    cudaFree(d_range);
    FCB_CUDA(e, rc);
    const int64_t span = (int64_t)h_range[1] - h_range[0] + 1;
    *key_base = h_range[0] > h_range[1] ? 0 : (span <= AGK_MONTHS ? h_range[0] : h_range[1] - (AGK_MONTHS - 1));
    agk_build_keys_kernel<<<num_sms * 8, 256, 0, st>>>(dates, cats, cents, n_rows, *key_base, keys);
    FCB_CUDA(e, cudaGetLastError());
    return FCB_OK;
}

// Called by fcb_load_shard when the shard has dates, categories and amounts. Not being able to allocate the column is not
// an error: the date kernels serve.
int32_t fcb_build_group_keys(fcb_engine *e) {
    const fcb_shard &s = e->shard;
    cudaFree(e->group_keys); e->group_keys = nullptr; e->key_base = 0;
    const char *knob = getenv("FCB_AGG_KEYS");
    if ((knob && knob[0] == '0') || !s.dates || !s.categories || !s.cents || !s.seg_offsets || s.n_rows <= 0) return FCB_OK;
    if ((reinterpret_cast<uintptr_t>(s.cents) & 15u) != 0) return FCB_OK;          // the key kernel loads amounts 16 bytes at a time
    uint16_t *keys = nullptr;
    if (cudaMalloc(&keys, ((size_t)s.n_rows * 2 + 15) & ~(size_t)15) != cudaSuccess) { cudaGetLastError(); return FCB_OK; }
    const int32_t rc = agk_build(e, e->num_sms, s.dates, s.categories, s.cents, s.n_rows, keys, &e->key_base, e->stream);
    if (rc != FCB_OK) { cudaFree(keys); return rc; }
    e->group_keys = keys;
    return FCB_OK;
}

// ---- date-ordered segments: no atomics on the hot path ---------------------------------------------
// Transaction logs are appended in time order, so within a user the rows of one calendar month form
// one contiguous run. A warp then only has to tell CATEGORIES apart: every lane owns a private
// column of bins in shared memory (bins[category][lane], one bank per lane: conflict-free plain
// load/add/store, no atomics), a run ends with a warp-shuffle reduction of the 16 (<= 32) bins and one
// 64-bit global reduction per non-empty cell. Correct for ANY row order (a step that mixes months is
// processed month by month, flushing in between) — only the speed relies on the order, so the host
// selects this kernel when the load-time check found the dates non-decreasing within every user.
#define AGS_WARP_ROWS 4096                 // consecutive rows one warp takes per task

#ifndef AGS_AHEAD
#define AGS_AHEAD 1                        // steps of columns each warp keeps in flight (ring of AGS_AHEAD + 1 slots: power of two)
#endif
#ifndef AGS_CTAS
#define AGS_CTAS 4
#endif
#define AGS_SLOT_BYTES 1152                // one step of one warp: 32 x (16 B cents + 16 B dates + 4 B categories)

__device__ __forceinline__ void ags_cp16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void ags_cp4(void *dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
#ifndef AGS_SMALL_PIECE
#define AGS_SMALL_PIECE 1024               // shorter pieces (tiny users): per-row 64-bit global reductions
#endif

__global__ void __launch_bounds__(AGG_THREADS, AGS_CTAS) aggregate_sorted_kernel(const __grid_constant__ AggParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int C = p.n_categories * p.n_months;
    uint16_t *month_tab = reinterpret_cast<uint16_t *>(smem_raw);                                   // [n_days] day -> month
    uint32_t *mstart = reinterpret_cast<uint32_t *>(smem_raw + (((size_t)p.n_days * 2 + 15) & ~(size_t)15));   // [n_months + 1]
    long long *bins_all = reinterpret_cast<long long *>(reinterpret_cast<unsigned char *>(mstart) +
                                                        (((size_t)(p.n_months + 1) * 4 + 15) & ~(size_t)15));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    long long *bins = bins_all + (size_t)warp * p.n_categories * 32;                                // [n_categories][32]
    const bool table_ok = C * 9 <= p.n_categories * 32 * 8;        // the table (lo, hi, presence) fits the warp's bin matrix
    unsigned char *ring = reinterpret_cast<unsigned char *>(bins_all + (size_t)(AGG_THREADS / 32) * p.n_categories * 32);   // per-warp cp.async rings
    for (int i = tid; i < p.n_days; i += AGG_THREADS) month_tab[i] = (uint16_t)(month_of_day(p.day0 + i) - p.month_lo);
    for (int i = lane; i < p.n_categories * 32; i += 32) bins[i] = 0;
    __syncthreads();
    // seconds (relative to the window start) at which each month of the window begins
    for (int i = tid; i < p.n_days; i += AGG_THREADS)
        if (i == 0 || month_tab[i] != month_tab[i - 1]) mstart[month_tab[i]] = (uint32_t)i * 86400u;
    if (tid == 0) mstart[p.n_months] = (uint32_t)p.n_days * 86400u;
    __syncthreads();

    int64_t R0 = p.seg_offsets[p.seg_begin], R1 = p.seg_offsets[p.seg_end];
    if (R0 < p.row_base) R0 = p.row_base;
    if (R1 > p.row_base + p.n_rows) R1 = p.row_base + p.n_rows;
    const int64_t n_tasks = R1 > R0 ? (R1 - R0 + AGS_WARP_ROWS - 1) / AGS_WARP_ROWS : 0;
    const int4 *cents4 = reinterpret_cast<const int4 *>(p.cents);
    const int4 *dates4 = reinterpret_cast<const int4 *>(p.dates);
    const uchar4 *cats4 = reinterpret_cast<const uchar4 *>(p.cats);
    const int64_t t_window = (int64_t)p.day0 * 86400;
    const uint32_t window_len = (uint32_t)p.n_days * 86400u;
    unsigned dropped = 0;

    int64_t cur_user = -1;
    int cur_month = -1;
    uint32_t cur_base = 0, cur_len = 0;                             // the current run's month as [base, base + len) unix seconds (mod 2^32)
    uint32_t seen = 0;                                              // categories this lane added to since the last flush

    // end of a (user, month) run: lane c sums row c of the warp's bin matrix (skewed column order: no bank
    // conflicts), clears it, and issues the one 64-bit global reduction of that cell
    auto flush = [&]() {
        uint32_t any = seen;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) any |= __shfl_xor_sync(FCB_FULL, any, off);
        if (cur_month >= 0 && any != 0) {
            __syncwarp();
            if (lane < p.n_categories && ((any >> lane) & 1u)) {
                long long total = 0;
#pragma unroll 8
                for (int i = 0; i < 32; ++i) {
                    const int col = (i + lane) & 31;
                    total += bins[lane * 32 + col];
                    bins[lane * 32 + col] = 0;
                }
                const int cell = lane * p.n_months + cur_month;
                int64_t *out_s = p.sums + (cur_user - p.seg_begin) * (int64_t)C;
                atomicAdd(reinterpret_cast<unsigned long long *>(out_s + cell), (unsigned long long)total);
                p.present[(cur_user - p.seg_begin) * (int64_t)C + cell] = 1;
            }
            __syncwarp();
        }
        seen = 0;
    };
    auto set_month = [&](int mo) {
        cur_month = mo;
        cur_base = mo >= 0 ? (uint32_t)t_window + mstart[mo] : 0;
        cur_len = mo >= 0 ? mstart[mo + 1] - mstart[mo] : 0;
    };

    struct Quad { int4 v, d; uchar4 c; };
    auto load_quad = [&](int64_t r, int64_t lb) {
        Quad q;
        q.v = make_int4(0, 0, 0, 0); q.d = make_int4(0, 0, 0, 0); q.c = make_uchar4(0, 0, 0, 0);
        if (r < lb) {
            if (r + 4 <= p.n_rows) {
                q.v = __ldg(cents4 + (r >> 2)); q.d = __ldg(dates4 + (r >> 2)); q.c = __ldg(cats4 + (r >> 2));
            } else {
                int v[4] = {0, 0, 0, 0}, d[4] = {0, 0, 0, 0};
                unsigned char c[4] = {0, 0, 0, 0};
                for (int i = 0; i < 4 && r + i < p.n_rows; ++i) { v[i] = p.cents[r + i]; d[i] = p.dates[r + i]; c[i] = p.cats[r + i]; }
                q.v = make_int4(v[0], v[1], v[2], v[3]); q.d = make_int4(d[0], d[1], d[2], d[3]);
                q.c = make_uchar4(c[0], c[1], c[2], c[3]);
            }
        }
        return q;
    };

    const int64_t n_warps = (int64_t)gridDim.x * (AGG_THREADS / 32);
    for (int64_t t = (int64_t)blockIdx.x * (AGG_THREADS / 32) + warp; t < n_tasks; t += n_warps) {
        const int64_t a = R0 + t * AGS_WARP_ROWS;
        const int64_t b = a + AGS_WARP_ROWS < R1 ? a + AGS_WARP_ROWS : R1;
        int64_t u;                                                  // first user whose segment ends after a
        {
            int64_t lo = p.seg_begin, hi = p.seg_end - 1;
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if (p.seg_offsets[mid + 1] > a) hi = mid; else lo = mid + 1;
            }
            u = lo;
        }
        int64_t pa = a;
        while (pa < b) {
            int64_t uend = p.seg_offsets[u + 1];
            while (uend <= pa) { ++u; uend = p.seg_offsets[u + 1]; }
            const int64_t pb = uend < b ? uend : b;
            const int64_t la = pa - p.row_base, lb = pb - p.row_base;
            if (table_ok && uend - p.seg_offsets[u] < AGS_TABLE_USER) {
                // small user (a month is only a few dozen rows, runs are not worth setting up): the warp's bin matrix,
                // empty between runs, serves as a private [category, month] table — 32-bit shared atomics (one per row for
                // |v| < 2^18, a lo/hi split otherwise; a piece has at most 4096 rows, so no cell can overflow), then the
                // present cells go out once: plain stores when the piece is the whole user, 64-bit reductions otherwise
                if (cur_month >= 0) flush();
                cur_user = -1; set_month(-1);
                __syncwarp();
                int32_t *t_lo = reinterpret_cast<int32_t *>(bins), *t_hi = t_lo + C;
                agg_warp_table(p, la, lb, pa == p.seg_offsets[u] && pb == uend, u, t_lo, t_hi, reinterpret_cast<uint8_t *>(t_hi + C),
                               month_tab, lane, dropped);
                pa = pb;
                continue;
            }
            if (p.split == 2) { pa = pb; continue; }               // large users belong to the other launch
            if (pb - pa < AGS_SMALL_PIECE) {
                // tiny piece of a large user (a task boundary): months change every few rows, a run is not worth setting up
                int64_t *out_s = p.sums + (u - p.seg_begin) * (int64_t)C;
                uint8_t *out_p = p.present + (u - p.seg_begin) * (int64_t)C;
                for (int64_t r = la + lane; r < lb; r += 32) {
                    const unsigned cat = p.cats[r];
                    const uint64_t off_s = (uint64_t)((int64_t)p.dates[r] - t_window);
                    const unsigned mo = off_s < window_len ? month_tab[(uint32_t)off_s / 86400u] : 0xffffu;
                    if (cat >= (unsigned)p.n_categories || mo >= (unsigned)p.n_months) { ++dropped; continue; }
                    const int cell = cat * p.n_months + mo;
                    atomicAdd(reinterpret_cast<unsigned long long *>(out_s + cell), (unsigned long long)(long long)p.cents[r]);
                    out_p[cell] = 1;
                }
                pa = pb;
                continue;
            }
            if (u != cur_user) { flush(); cur_user = u; set_month(-1); }

            // one step of <= 128 rows whose rows `in_mask` belong to the piece: fast path if they all continue
            // the current run, else resolve the months one at a time, flushing in between
            auto step = [&](const Quad &q, unsigned in_mask) {
                const int v[4] = {q.v.x, q.v.y, q.v.z, q.v.w};
                const int dts[4] = {q.d.x, q.d.y, q.d.z, q.d.w};
                const unsigned cat[4] = {q.c.x, q.c.y, q.c.z, q.c.w};
                bool same = true;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const bool hit = (uint32_t)dts[i] - cur_base < cur_len && cat[i] < (unsigned)p.n_categories;
                    same = same && (hit || !((in_mask >> i) & 1u));
                }
                if (__all_sync(FCB_FULL, same)) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if ((in_mask >> i) & 1u) { bins[cat[i] * 32 + lane] += (long long)v[i]; seen |= 1u << cat[i]; }
                    }
                    return;
                }
                int m[4];
                unsigned todo = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const bool in = (in_mask >> i) & 1u;
                    const uint64_t off_s = (uint64_t)((int64_t)dts[i] - t_window);
                    const unsigned mo = off_s < window_len ? month_tab[(uint32_t)off_s / 86400u] : 0xffffu;
                    const bool ok = in && cat[i] < (unsigned)p.n_categories && mo < (unsigned)p.n_months;
                    if (in && !ok) ++dropped;
                    m[i] = ok ? (int)mo : 0x7fffffff;
                    if (ok) todo |= 1u << i;
                }
                while (__any_sync(FCB_FULL, todo != 0)) {
                    int mn = 0x7fffffff;
#pragma unroll
                    for (int i = 0; i < 4; ++i) if ((todo >> i) & 1u) mn = min(mn, m[i]);
#pragma unroll
                    for (int off = 16; off >= 1; off >>= 1) mn = min(mn, __shfl_xor_sync(FCB_FULL, mn, off));
                    if (mn != cur_month) { flush(); set_month(mn); }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (((todo >> i) & 1u) && m[i] == cur_month) {
                            bins[cat[i] * 32 + lane] += (long long)v[i];
                            seen |= 1u << cat[i];
                            todo &= ~(1u << i);
                        }
                    }
                }
            };
            auto ragged = [&](int64_t base) {                      // a step that may stick out of [la, lb) / the arrays
                const int64_t r = base + 4 * lane;
                unsigned in_mask = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) if (r + i >= la && r + i < lb) in_mask |= 1u << i;
                step(load_quad(r, lb), in_mask);
            };

            int64_t base = la & ~(int64_t)3;
            if (base < la) { ragged(base); base += 128; }           // head: first quad not aligned with the piece
            // full steps: all 128 rows inside the piece; plain pointer bumps, next step's columns already in flight
            const int64_t full_end = lb - 127;                      // base < full_end  <=>  base + 128 <= lb
            if (base < full_end) {
                // 32-bit step counter and 32-bit offsets from the task's first quad (one IMAD.WIDE per load instead of
                // predicated 64-bit pointer bumps); the next step's columns are always in flight (the last step re-reads
                // itself, so the prefetch needs no predicate); categories stay ONE raw 32-bit register until they are
                // used, so nothing waits for the load inside the step that issued it
                const int n_full = (int)((full_end - base + 127) >> 7);
                const int4 *tv = cents4 + (base >> 2) + lane, *td = dates4 + (base >> 2) + lane;
                const uint32_t *tc = reinterpret_cast<const uint32_t *>(p.cats) + (base >> 2) + lane;
                char *bl = reinterpret_cast<char *>(bins + lane);   // this lane's column; a category's bin is 256 bytes further per code
                // Columns travel global -> shared with cp.async, AGS_AHEAD steps ahead, into this warp's ring of
                // AGS_AHEAD + 1 slots (a lane only ever reads what it copied itself, so no warp barrier is involved): data in
                // flight costs shared memory instead of registers.
                unsigned char *my_ring = ring + (size_t)warp * (AGS_AHEAD + 1) * AGS_SLOT_BYTES;
                auto issue = [&](int j) {
                    const int nx = min(j, n_full - 1) * 32;          // past the end: re-read the last step (harmless, no predicate)
                    unsigned char *slot = my_ring + (size_t)(j & AGS_AHEAD) * AGS_SLOT_BYTES;
                    ags_cp16(slot + lane * 16, tv + nx);
                    ags_cp16(slot + 512 + lane * 16, td + nx);
                    ags_cp4(slot + 1024 + lane * 4, tc + nx);
                    asm volatile("cp.async.commit_group;" ::: "memory");
                };
#pragma unroll
                for (int j = 0; j < AGS_AHEAD; ++j) issue(j);
                for (int it = 0; it < n_full; ++it) {
                    asm volatile("cp.async.wait_group %0;" ::"n"(AGS_AHEAD - 1) : "memory");
                    const unsigned char *slot = my_ring + (size_t)(it & AGS_AHEAD) * AGS_SLOT_BYTES;
                    const int4 v = *reinterpret_cast<const int4 *>(slot + lane * 16);
                    const int4 d = *reinterpret_cast<const int4 *>(slot + 512 + lane * 16);
                    const uint32_t c = *reinterpret_cast<const uint32_t *>(slot + 1024 + lane * 4);
                    issue(it + AGS_AHEAD);                            // into the slot read one step ago
                    const unsigned c0 = c & 0xffu, c1 = (c >> 8) & 0xffu, c2 = (c >> 16) & 0xffu, c3 = c >> 24;
                    const bool same =
                        (uint32_t)d.x - cur_base < cur_len && (uint32_t)d.y - cur_base < cur_len &&
                        (uint32_t)d.z - cur_base < cur_len && (uint32_t)d.w - cur_base < cur_len &&
                        max(max(c0, c1), max(c2, c3)) < (unsigned)p.n_categories;
                    if (__all_sync(FCB_FULL, same)) {
                        *reinterpret_cast<long long *>(bl + (c0 << 8)) += (long long)v.x;
                        *reinterpret_cast<long long *>(bl + (c & 0xff00u)) += (long long)v.y;
                        *reinterpret_cast<long long *>(bl + ((c >> 8) & 0xff00u)) += (long long)v.z;
                        *reinterpret_cast<long long *>(bl + (c3 << 8)) += (long long)v.w;
                        seen |= (1u << c0) | (1u << c1) | (1u << c2) | (1u << c3);
                    } else {
                        Quad q;
                        q.v = v; q.d = d; q.c = make_uchar4((unsigned char)c0, (unsigned char)c1, (unsigned char)c2, (unsigned char)c3);
                        step(q, 0xfu);
                    }
                }
                asm volatile("cp.async.wait_group 0;" ::: "memory");  // the re-reads past the end must land before the ring is reused
                base += (int64_t)n_full * 128;
            }
            if (base < lb) ragged(base);                            // tail
            pa = pb;
        }
    }
    flush();
    if (dropped) atomicAdd(p.dropped, (unsigned long long)dropped);
}

// non-decreasing dates within every user of [seg_begin, seg_end)? (load-time check; writes 0 on a violation)
__global__ void __launch_bounds__(256) dates_sorted_kernel(const int64_t *__restrict__ seg_offsets, int64_t n_segs,
                                                           const int32_t *__restrict__ dates, int64_t row_base,
                                                           int64_t n_rows, int *__restrict__ sorted_flag) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r + 1 < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
        if (dates[r] > dates[r + 1]) {
            // a decrease is fine only across a user boundary: is global row (row_base + r + 1) a segment start?
            const int64_t g = row_base + r + 1;
            int64_t lo = 0, hi = n_segs;
            while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (seg_offsets[mid] < g) lo = mid + 1; else hi = mid; }
            if (!(lo <= n_segs && seg_offsets[lo] == g)) *sorted_flag = 0;
        }
    }
}

int32_t fcb_check_dates_sorted(fcb_engine *e) {
    const fcb_shard &s = e->shard;
    e->dates_sorted = 0;
    if (!s.dates || !s.seg_offsets || s.n_segs <= 0) return FCB_OK;
    int *flag = nullptr;
    FCB_CUDA(e, cudaMalloc(&flag, sizeof(int)));
    int one = 1;
    FCB_CUDA(e, cudaMemcpyAsync(flag, &one, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    dates_sorted_kernel<<<e->num_sms * 8, 256, 0, e->stream>>>(s.seg_offsets, s.n_segs, s.dates, s.row_base, s.n_rows, flag);
    int h = 0;
    FCB_CUDA(e, cudaMemcpyAsync(&h, flag, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    FCB_CUDA(e, cudaStreamSynchronize(e->stream));
    cudaFree(flag);
    e->dates_sorted = h;
    return FCB_OK;
}

// One side stream + two events per device, created on first use (a stream is a handle, not memory): lets the two launches
// of the any-order split overlap.
struct AggSide { cudaStream_t stream; cudaEvent_t fork, join; int state; };   // state: 0 = not tried, 1 = ready, -1 = unavailable
static AggSide g_agg_side[64];
static AggSide *agg_side_stream(int device) {
    if (device < 0 || device >= 64) return nullptr;
    AggSide &a = g_agg_side[device];
    if (a.state == 0) {
        a.state = (cudaStreamCreateWithFlags(&a.stream, cudaStreamNonBlocking) == cudaSuccess &&
                   cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming) == cudaSuccess &&
                   cudaEventCreateWithFlags(&a.join, cudaEventDisableTiming) == cudaSuccess) ? 1 : -1;
        if (a.state < 0) cudaGetLastError();
    }
    return a.state == 1 ? &a : nullptr;
}

// Validates the window, zeroes the outputs and launches the kernel(s) for one shard's columns. `e` may be NULL (the
// engine-less entry point): errors then land in the per-thread message fcb_last_error(NULL) returns.
static int32_t agg_launch(fcb_engine *e, int device, int num_sms, int dates_sorted, const int32_t *dates, const uint8_t *cats,
                          const int32_t *cents, const int64_t *seg_offsets, int64_t n_rows, int64_t n_segs, int64_t row_base,
                          int64_t seg_begin, int64_t seg_end, int32_t month_lo, int32_t n_months, int32_t n_categories,
                          int64_t *sums, uint8_t *present, uint64_t *dropped, cudaStream_t st, const uint16_t *keys = nullptr,
                          int32_t key_base = 0) {
    if (!cents || !seg_offsets || (!keys && (!dates || !cats)))
        return fcb_fail(e, FCB_ERR_STATE, "fcb_aggregate: needs the dates, categories, cents and seg_offsets columns");
    if (!sums || !present || !dropped) return fcb_fail(e, FCB_ERR_INVALID, "fcb_aggregate: NULL output");
    // the kernels read the columns 4 rows at a time (128-bit loads of cents / dates, 32-bit of categories, 64-bit of keys)
    if ((reinterpret_cast<uintptr_t>(cents) & 15u) || (reinterpret_cast<uintptr_t>(dates) & 15u) ||
        (reinterpret_cast<uintptr_t>(cats) & 3u) || (reinterpret_cast<uintptr_t>(keys) & 7u) ||
        (reinterpret_cast<uintptr_t>(seg_offsets) & 7u))
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_aggregate: columns must be aligned (cents / dates 16 bytes, categories 4, keys 8, "
                        "seg_offsets 8)");
    if (seg_begin < 0 || seg_end > n_segs || seg_begin > seg_end)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_aggregate: segment range [%lld, %lld) outside 0..%lld",
                        (long long)seg_begin, (long long)seg_end, (long long)n_segs);
    if (n_months < 1 || n_months > AGG_MAX_MONTHS || n_categories < 1 || n_categories > 64 ||
        (int64_t)n_months * n_categories > AGG_MAX_CELLS)
        return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_aggregate: need 1..%d months, 1..64 categories, <= %d cells per user",
                        AGG_MAX_MONTHS, AGG_MAX_CELLS);
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(e, FCB_ERR_CUDA, "fcb_aggregate: cannot select device %d", device);
    const int64_t segs = seg_end - seg_begin;
    const int C = n_months * n_categories;
    FCB_CUDA(e, cudaMemsetAsync(sums, 0, (size_t)segs * C * sizeof(int64_t), st));
    FCB_CUDA(e, cudaMemsetAsync(present, 0, (size_t)segs * C, st));
    FCB_CUDA(e, cudaMemsetAsync(dropped, 0, sizeof(uint64_t), st));
    if (segs == 0) return FCB_OK;

    AggParams p{};
    p.seg_offsets = seg_offsets; p.dates = dates; p.cats = cats; p.cents = cents;
    p.seg_begin = seg_begin; p.seg_end = seg_end; p.row_base = row_base; p.n_rows = n_rows;
    p.month_lo = month_lo; p.n_months = n_months; p.n_categories = n_categories;
    p.day0 = first_day_of_month(month_lo);
    p.n_days = first_day_of_month(month_lo + n_months) - p.day0;
    {
        const int64_t t0 = (int64_t)p.day0 * 86400, t1 = t0 + (int64_t)p.n_days * 86400;
        const int64_t lo = t0 < INT32_MIN ? (int64_t)INT32_MIN : t0, hi = t1 > (int64_t)INT32_MAX + 1 ? (int64_t)INT32_MAX + 1 : t1;
        p.base32 = (int32_t)(hi > lo ? lo : 0);
        p.window32 = hi > lo ? (uint32_t)(hi - lo) : 0u;
        p.shift32 = hi > lo ? (uint32_t)(lo - t0) : 0u;
    }
    p.sums = sums; p.present = present; p.dropped = (unsigned long long *)dropped;
    p.small_piece = AGG_SMALL_PIECE;
    if (const char *knob = getenv("FCB_AGG_SMALL_PIECE")) p.small_piece = atoll(knob);   // tuning knob
    const size_t smem = (size_t)C * 16 + (size_t)((p.n_days + 1) & ~1) * 2 + (size_t)C * 2 + 16;   // two sets of cells
    if (smem > 48 * 1024)
        FCB_CUDA(e, cudaFuncSetAttribute(aggregate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // FCB_AGG_KERNEL (tuning / tests): "key" and "warp" = the keyed / date form of the warp-table kernel, "date" = whatever the
    // date columns would get, "cta" = round 1's routing for any-order dates (CTA-wide atomic kernel + warp tables for small
    // users), "atomic" / "sorted" = that one kernel
    const char *force = getenv("FCB_AGG_KERNEL");
    // The warp-table kernel (aggregate_warp_kernel): any row order, one launch. Keyed form when the shard / caller has group
    // keys and the window lies inside their range; date form otherwise. 8 warps per CTA up to ~1900 cells, fewer above.
    auto launch_warp_tables = [&](bool keyed) -> int32_t {            // FCB_OK, an error, or 1 = the tables do not fit
        const char *lay = getenv("FCB_AGK_LAYOUT");                 // tuning: "n" = natural stride, neg table right after pos
        const bool natural = lay && lay[0] == 'n';
        p.key_stride = natural ? (uint32_t)n_months : (uint32_t)(n_months | 1);
        const uint32_t tab_words = (uint32_t)n_categories * p.key_stride;
        const uint32_t neg_word = natural ? tab_words
                                          : (tab_words <= 16 ? 16u : ((tab_words - 16u + 31u) / 32u) * 32u + 16u);   // >= tab_words, = 16 (mod 32)
        p.key_neg_off = 4u * neg_word;
        p.key_hi_off = 4u * (neg_word + tab_words);
        p.key_step_mo = 32u % (uint32_t)n_months;
        p.key_step_w4 = 4u * ((32u / (uint32_t)n_months) * p.key_stride + p.key_step_mo);
        p.key_wrap_w4 = 4u * (p.key_stride - (uint32_t)n_months);
        const size_t tab_bytes = (((size_t)p.key_hi_off + 5u * (size_t)C + 15u) & ~(size_t)15) + 128;
        const size_t mtab_bytes = keyed ? 0 : ((((size_t)p.n_days + 1) * 2 + 15) & ~(size_t)15);
        int warps = (int)((200u * 1024u - mtab_bytes) / tab_bytes);
        if (warps > AGG_THREADS / 32) warps = AGG_THREADS / 32;
        if (warps < 1) return 1;
        auto kern = keyed ? aggregate_warp_kernel<true> : aggregate_warp_kernel<false>;
        const size_t smem_k = mtab_bytes + tab_bytes * warps;
        if (smem_k > 48 * 1024) FCB_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_k));
        // shared-memory carve-out: what AGK_CTAS CTAs need and no more — the rest of the 256 KB stays L1, which is where
        // the loads in flight are tracked (with the maximum carve-out the kernel ran 14 % slower)
        int carve = (int)(((smem_k + 1024) * AGK_CTAS * 100 + 233471) / 233472);
        if (const char *knob = getenv("FCB_AGK_CARVEOUT")) carve = atoi(knob);      // tuning
        if (carve > 100) carve = 100;
        FCB_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
        p.keys = keyed ? keys : nullptr;
        p.key_mo_off = keyed ? (uint32_t)(month_lo - key_base) : 0u;
        int per_sm = 0;
        FCB_CUDA(e, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, warps * 32, smem_k));
        if (per_sm > AGK_CTAS) per_sm = AGK_CTAS;
        if (per_sm < 1) per_sm = 1;
        kern<<<num_sms * per_sm, warps * 32, smem_k, st>>>(p);
        FCB_CUDA(e, cudaGetLastError());
        return FCB_OK;
    };
    const bool key_range = keys && month_lo >= key_base && (int64_t)month_lo + n_months <= (int64_t)key_base + AGK_MONTHS;
    if (key_range && !(force && force[0] != 'k')) {
        const int32_t rc = launch_warp_tables(true);
        if (rc <= 0) return rc;
    }
    if (!dates || !cats)
        return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_aggregate: months [%d, %d) reach outside the group keys' range [%d, %d) and there "
                        "are no date / category columns to fall back on", month_lo, month_lo + n_months, key_base, key_base + AGK_MONTHS);
    if (force && (force[0] == 'k' || force[0] == 'd')) force = nullptr;   // ("date": no keys, the usual choice for date columns)
    if (!force || force[0] == 'w') {                                  // date columns, any row order: the date form of the warp-table
                                                                      // kernel (1.9 ms per 1B rows; round 1's kernels: 2.9 any order, 2.1 time-ordered)
        const int32_t rc = launch_warp_tables(false);
        if (rc <= 0) return rc;
    }
    if (force && (force[0] == 'w' || force[0] == 'c')) force = nullptr;   // "cta": round 1's routing below
    const bool sorted_ok = n_categories <= 32;
    const bool use_sorted = sorted_ok && (force ? force[0] == 's' : dates_sorted != 0);
    // any-order data whose table fits the warp's bin matrix: users below AGS_TABLE_USER rows go to the ordered kernel's
    // warp-table path (order-agnostic, 4-6 ns per 1000 rows; the atomic kernel handles them CTA-wide at 7-12 ns), the rest
    // to the atomic kernel; the two launches touch disjoint users
    const char *split_knob = getenv("FCB_AGG_SPLIT");
    const bool split = !use_sorted && sorted_ok && C * 9 <= n_categories * 32 * 8 && !(split_knob && split_knob[0] == '0') &&
                       !(force && force[0] == 'a');
    const size_t smem_s = (((size_t)p.n_days * 2 + 15) & ~(size_t)15) + (((size_t)(n_months + 1) * 4 + 15) & ~(size_t)15) +
                          (size_t)(AGG_THREADS / 32) * n_categories * 32 * 8 +
                          (size_t)(AGG_THREADS / 32) * (AGS_AHEAD + 1) * AGS_SLOT_BYTES;
    if ((split || use_sorted) && smem_s > 48 * 1024)
        FCB_CUDA(e, cudaFuncSetAttribute(aggregate_sorted_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_s));
    if (split) {
        // the two launches touch disjoint users: the small-user one runs on a side stream next to the large-user one
        // (fork / join by events), so its tail does not add to the total. FCB_AGG_CONCURRENT=0: one after the other.
        const char *conc = getenv("FCB_AGG_CONCURRENT");
        AggSide *side = (conc && conc[0] == '0') ? nullptr : agg_side_stream(device);
        p.split = 2;
        if (side) {
            FCB_CUDA(e, cudaEventRecord(side->fork, st));
            FCB_CUDA(e, cudaStreamWaitEvent(side->stream, side->fork, 0));
            aggregate_sorted_kernel<<<num_sms * AGS_CTAS, AGG_THREADS, smem_s, side->stream>>>(p);
            FCB_CUDA(e, cudaEventRecord(side->join, side->stream));
        } else {
            aggregate_sorted_kernel<<<num_sms * AGS_CTAS, AGG_THREADS, smem_s, st>>>(p);
        }
        p.split = 1;
        aggregate_kernel<<<num_sms * AGG_CTAS, AGG_THREADS, smem, st>>>(p);
        if (side) FCB_CUDA(e, cudaStreamWaitEvent(st, side->join, 0));
    } else if (use_sorted) {
        aggregate_sorted_kernel<<<num_sms * AGS_CTAS, AGG_THREADS, smem_s, st>>>(p);
    } else {
        aggregate_kernel<<<num_sms * AGG_CTAS, AGG_THREADS, smem, st>>>(p);
    }
    FCB_CUDA(e, cudaGetLastError());
    return FCB_OK;
}

extern "C" int32_t fcb_aggregate(fcb_engine *e, int64_t seg_begin, int64_t seg_end, int32_t month_lo, int32_t n_months,
                                 int32_t n_categories, int64_t *sums, uint8_t *present, uint64_t *dropped, void *stream) {
    if (!e) return FCB_ERR_INVALID;
    if (!e->loaded) return fcb_fail(e, FCB_ERR_STATE, "fcb_aggregate: no shard loaded");
    const fcb_shard &s = e->shard;
    return agg_launch(e, e->device, e->num_sms, e->dates_sorted, s.dates, s.categories, s.cents, s.seg_offsets, s.n_rows, s.n_segs,
                      s.row_base, seg_begin, seg_end, month_lo, n_months, n_categories, sums, present, dropped, (cudaStream_t)stream,
                      e->group_keys, e->key_base);
}

// The same group-by over caller-owned device columns, without an engine: no vectors, no inverse norms, nothing allocated.
extern "C" int32_t fcb_aggregate_columns(int32_t device, const int32_t *dates, const uint8_t *categories, const int32_t *cents,
                                         const int64_t *seg_offsets, int64_t n_rows, int64_t n_segs, int64_t row_base,
                                         int32_t dates_sorted, int64_t seg_begin, int64_t seg_end, int32_t month_lo,
                                         int32_t n_months, int32_t n_categories, int64_t *sums, uint8_t *present,
                                         uint64_t *dropped, void *stream) {
    if (n_rows < 0 || n_segs < 0 || row_base < 0)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_aggregate_columns: negative n_rows / n_segs / row_base");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_aggregate_columns: no CUDA device %d (this library has no CPU path)", device);
    static int sms[64];                                            // SM count per device, looked up once
    if (device < 64 && sms[device] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || v <= 0)
            return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_aggregate_columns: cannot query device %d", device);
        sms[device] = v;
    }
    return agg_launch(nullptr, device, device < 64 ? sms[device] : 148, dates_sorted, dates, categories, cents, seg_offsets, n_rows,
                      n_segs, row_base, seg_begin, seg_end, month_lo, n_months, n_categories, sums, present, dropped,
                      (cudaStream_t)stream);
}

static int agg_device_sms(int32_t device, const char *who) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        fcb_fail(nullptr, FCB_ERR_CUDA, "%s: no CUDA device %d (this library has no CPU path)", who, device);
        return -1;
    }
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || v <= 0) {
        fcb_fail(nullptr, FCB_ERR_CUDA, "%s: cannot query device %d", who, device);
        return -1;
    }
    return v;
}

// Group keys for caller-owned columns (what fcb_load_shard builds for an engine's shard): see "group keys" above.
extern "C" int32_t fcb_group_keys_build(int32_t device, const int32_t *dates, const uint8_t *categories, const int32_t *cents,
                                        int64_t n_rows, uint16_t *keys, int32_t *key_base, void *stream) {
    if (!dates || !categories || !cents || !keys || !key_base || n_rows < 0)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_group_keys_build: bad argument");
    if ((reinterpret_cast<uintptr_t>(keys) & 15u) != 0)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_group_keys_build: keys must be a 16-byte aligned device pointer");
    const int sms = agg_device_sms(device, "fcb_group_keys_build");
    if (sms < 0) return FCB_ERR_CUDA;
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_group_keys_build: cannot select device %d", device);
    *key_base = 0;
    if (n_rows == 0) return FCB_OK;
    return agk_build(nullptr, sms, dates, categories, cents, n_rows, keys, key_base, (cudaStream_t)stream);
}

extern "C" int32_t fcb_aggregate_columns_keyed(int32_t device, const uint16_t *keys, int32_t key_base, const int32_t *dates,
                                               const uint8_t *categories, const int32_t *cents, const int64_t *seg_offsets,
                                               int64_t n_rows, int64_t n_segs, int64_t row_base, int32_t dates_sorted,
                                               int64_t seg_begin, int64_t seg_end, int32_t month_lo, int32_t n_months,
                                               int32_t n_categories, int64_t *sums, uint8_t *present, uint64_t *dropped, void *stream) {
    if (n_rows < 0 || n_segs < 0 || row_base < 0)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_aggregate_columns_keyed: negative n_rows / n_segs / row_base");
    const int sms = agg_device_sms(device, "fcb_aggregate_columns_keyed");
    if (sms < 0) return FCB_ERR_CUDA;
    return agg_launch(nullptr, device, sms, dates_sorted, dates, categories, cents, seg_offsets, n_rows, n_segs, row_base, seg_begin,
                      seg_end, month_lo, n_months, n_categories, sums, present, dropped, (cudaStream_t)stream, keys, key_base);
}

// ---- group-by over the points a search just returned ------------------------------
// One CTA per query: each returned point id is mapped back to its row (ids absent: id - row_base; ids
// ascending: binary search; otherwise binary search through the inverse of id_rank), the row's cents are
// added to the CTA's shared [category, month] cells with 64-bit shared-memory atomics (at most 10000
// points per query: latency-bound, not a bandwidth problem), and the cells are written out densely.
struct AggHitsParams {
    const uint64_t *hit_ids;               // [batch, k] as written by fcb_search
    const int32_t *hit_counts;             // [batch]
    int k;
    const uint64_t *ids;                   // shard columns (device)
    const uint32_t *row_of_rank;           // inverse of id_rank, or NULL when the ids ascend with the rows
    const int32_t *dates;
    const uint8_t *cats;
    const int32_t *cents;
    int64_t row_base, n_rows;
    int32_t month_lo, n_months, n_categories;
    int64_t *sums;                         // [batch, n_categories, n_months]
    uint8_t *present;
    unsigned long long *stats;             // [0] rows outside the window, [1] points not resident in this shard
};

__global__ void __launch_bounds__(256) aggregate_hits_kernel(const __grid_constant__ AggHitsParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int C = p.n_categories * p.n_months;
    unsigned long long *cells = reinterpret_cast<unsigned long long *>(smem_raw);   // [C]
    uint8_t *seen = reinterpret_cast<uint8_t *>(cells + C);                          // [C]
    const int tid = threadIdx.x, b = blockIdx.x;
    for (int c = tid; c < C; c += blockDim.x) { cells[c] = 0; seen[c] = 0; }
    __syncthreads();
    int n = p.hit_counts[b];
    if (n > p.k) n = p.k;
    unsigned dropped = 0, absent = 0;
    for (int j = tid; j < n; j += blockDim.x) {
        const uint64_t id = p.hit_ids[(int64_t)b * p.k + j];
        int64_t row = -1;
        if (!p.ids) {
            const int64_t r = (int64_t)id - p.row_base;
            if (id <= (uint64_t)INT64_MAX && r >= 0 && r < p.n_rows) row = r;
        } else {
            int64_t lo = 0, hi = p.n_rows;                          // first position whose id is >= the point's
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                const uint64_t at = p.ids[p.row_of_rank ? (int64_t)p.row_of_rank[mid] : mid];
                if (at < id) lo = mid + 1; else hi = mid;
            }
            if (lo < p.n_rows) {
                const int64_t r = p.row_of_rank ? (int64_t)p.row_of_rank[lo] : lo;
                if (p.ids[r] == id) row = r;
            }
        }
        if (row < 0) { ++absent; continue; }
        const unsigned cat = p.cats[row];
        const int32_t mo = month_of_day(floor_day(p.dates[row])) - p.month_lo;
        if (cat >= (unsigned)p.n_categories || mo < 0 || mo >= p.n_months) { ++dropped; continue; }
        const int cell = cat * p.n_months + mo;
        atomicAdd(&cells[cell], (unsigned long long)(long long)p.cents[row]);
        seen[cell] = 1;
    }
    __syncthreads();
    for (int c = tid; c < C; c += blockDim.x) {
        p.sums[(int64_t)b * C + c] = (int64_t)cells[c];
        p.present[(int64_t)b * C + c] = seen[c];
    }
    if (dropped) atomicAdd(p.stats, (unsigned long long)dropped);
    if (absent) atomicAdd(p.stats + 1, (unsigned long long)absent);
}

__global__ void __launch_bounds__(256) invert_rank_kernel(const uint32_t *__restrict__ id_rank, int64_t n_rows,
                                                          uint32_t *__restrict__ row_of_rank, unsigned *__restrict__ bad) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t k = id_rank[r];
        if ((int64_t)k < n_rows) row_of_rank[k] = (uint32_t)r; else atomicAdd(bad, 1u);
    }
}

__global__ void __launch_bounds__(256) count_unset_kernel(const uint32_t *__restrict__ row_of_rank, int64_t n_rows,
                                                          unsigned *__restrict__ bad) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x)
        if (row_of_rank[r] == 0xffffffffu) atomicAdd(bad, 1u);
}

// the inverse of id_rank, built on first use (ids that do not ascend with the rows)
static int32_t build_row_of_rank(fcb_engine *e, cudaStream_t st) {
    const fcb_shard &s = e->shard;
    if (s.n_rows >= 0xffffffffLL) return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_aggregate_hits: too many rows for a 32-bit row map");
    uint32_t *map = nullptr;
    unsigned *bad = nullptr;
    if (cudaMalloc(&map, (size_t)s.n_rows * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&bad, sizeof(unsigned)) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(map);
        return fcb_fail(e, FCB_ERR_NOMEM, "fcb_aggregate_hits: cannot allocate the %zu-byte row map", (size_t)s.n_rows * 4);
    }
    cudaMemsetAsync(map, 0xff, (size_t)s.n_rows * sizeof(uint32_t), st);
    cudaMemsetAsync(bad, 0, sizeof(unsigned), st);
    invert_rank_kernel<<<e->num_sms * 8, 256, 0, st>>>(s.id_rank, s.n_rows, map, bad);
    count_unset_kernel<<<e->num_sms * 8, 256, 0, st>>>(map, s.n_rows, bad);
    unsigned h_bad = 0;
    cudaMemcpyAsync(&h_bad, bad, sizeof(unsigned), cudaMemcpyDeviceToHost, st);
    const cudaError_t err = cudaStreamSynchronize(st);
    cudaFree(bad);
    if (err != cudaSuccess || h_bad) {
        cudaFree(map);
        if (err != cudaSuccess) return fcb_fail(e, FCB_ERR_CUDA, "fcb_aggregate_hits: %s", cudaGetErrorString(err));
        return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_aggregate_hits: id_rank is not a permutation of this shard's rows");
    }
    e->row_of_rank = map;
    return FCB_OK;
}

extern "C" int32_t fcb_aggregate_hits(fcb_engine *e, const uint64_t *hit_ids, const int32_t *hit_counts, int32_t batch,
                                      int32_t k, int32_t month_lo, int32_t n_months, int32_t n_categories, int64_t *sums,
                                      uint8_t *present, uint64_t *stats, void *stream) {
    if (!e) return FCB_ERR_INVALID;
    if (!e->loaded) return fcb_fail(e, FCB_ERR_STATE, "fcb_aggregate_hits: no shard loaded");
    const fcb_shard &s = e->shard;
    if (!s.dates || !s.categories || !s.cents)
        return fcb_fail(e, FCB_ERR_STATE, "fcb_aggregate_hits: the shard needs dates, categories and cents");
    if (!hit_ids || !hit_counts || !sums || !present || !stats)
        return fcb_fail(e, FCB_ERR_INVALID, "fcb_aggregate_hits: NULL argument");
    if (batch < 0 || k < 1) return fcb_fail(e, FCB_ERR_INVALID, "fcb_aggregate_hits: need batch >= 0 and k >= 1");
    if (n_months < 1 || n_months > AGG_MAX_MONTHS || n_categories < 1 || n_categories > 64 ||
        (int64_t)n_months * n_categories > AGG_MAX_CELLS)
        return fcb_fail(e, FCB_ERR_UNSUPPORTED, "fcb_aggregate_hits: need 1..%d months, 1..64 categories, <= %d cells per query",
                        AGG_MAX_MONTHS, AGG_MAX_CELLS);
    FcbDeviceGuard g(e->device);
    if (!g.ok) return fcb_fail(e, FCB_ERR_CUDA, "fcb_aggregate_hits: cannot select device %d", e->device);
    cudaStream_t st = (cudaStream_t)stream;
    FCB_CUDA(e, cudaMemsetAsync(stats, 0, 2 * sizeof(uint64_t), st));
    if (batch == 0) return FCB_OK;
    if (s.id_rank && !e->row_of_rank) {
        const int32_t rc = build_row_of_rank(e, st);
        if (rc != FCB_OK) return rc;
    }
    AggHitsParams p{};
    p.hit_ids = hit_ids; p.hit_counts = hit_counts; p.k = k;
    p.ids = s.ids; p.row_of_rank = s.id_rank ? e->row_of_rank : nullptr;
    p.dates = s.dates; p.cats = s.categories; p.cents = s.cents;
    p.row_base = s.row_base; p.n_rows = s.n_rows;
    p.month_lo = month_lo; p.n_months = n_months; p.n_categories = n_categories;
    p.sums = sums; p.present = present; p.stats = (unsigned long long *)stats;
    const size_t smem = (size_t)n_months * n_categories * 9;
    aggregate_hits_kernel<<<batch, 256, smem, st>>>(p);
    FCB_CUDA(e, cudaGetLastError());
    return FCB_OK;
}
