/* ORACLE — test infrastructure only (never linked into, or called by, the product).
 *
 * Plain-C restatement of the hot path, with the floating-point summation ORDER
 * pinned so that "bit-exact ids, ties by id" is a decidable statement:
 *
 *   - canon_search_*  : the payload-filtered exact cosine top-k behind
 *       client.query_points(...).points   (reference tools/qdrant_tool.py:147-153)
 *       with the filter of tools/qdrant_tool.py:105-131 (user_id match, date range;
 *       category match is the north_star extension) evaluated over columns.
 *       The arithmetic is that of qdrant-client 1.14.2 local mode (requirements.txt:78;
 *       third-party, absent — PARITY UNPINNED, see oracle/local_mode.py): float32
 *       storage, float64 query, cosine = dot / (|row| |query|), zero norms guarded by
 *       float32 eps, descending score, first `limit` filtered-in rows. Local mode leaves
 *       the summation order to BLAS and the tie order to numpy's unstable argsort; this
 *       file FIXES both (32 interleaved partial sums + xor-butterfly; ties by ascending
 *       id) — every order is a valid instance of the reference's algorithm, and pinning
 *       one lets the CUDA path be compared bit for bit.
 *   - canon_groupby   : DataFrame.groupby(key)[y].sum()  (reference tools/plot_tool.py:45,52)
 *       for key = (user, category, UTC calendar month), int64 sums of integer cents,
 *       keys ascending, empty groups absent (pandas sort=True, observed groups only).
 *
 * Canonical score of one row x[0..d) (float32, or bfloat16 widened exactly) against a
 * float64 query q[0..d):
 *     for l in 0..31: dot[l] = ss[l] = qq[l] = 0
 *     for j in 0..d-1 (ascending), l = j mod 32:
 *         dot[l] = fma(x[j], q[j], dot[l]); ss[l] = fma(x[j], x[j], ss[l]); qq[l] = fma(q[j], q[j], qq[l])
 *     for off in 16,8,4,2,1: for l < off: p[l] = p[l] + p[l+off]      (each of dot, ss, qq)
 *     nr = sqrt(ss[0]); if nr == 0: nr = EPS;  nq = sqrt(qq[0]); if nq == 0: nq = EPS
 *     score = dot[0] / (nr * nq)
 * All operations IEEE-754 binary64, round-to-nearest; fma is the fused operation.
 * Build with -ffp-contract=off so nothing else is fused.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CANON_EPS 1.1920929e-7

typedef struct {
    int64_t row_begin, row_end;   /* CSR segment of the user (row_begin >= row_end: no rows)   */
    int64_t date_lo, date_hi;     /* inclusive bounds on metadata.date (unix seconds)            */
    uint64_t category_mask;       /* bit c set => category c passes; ~0 => no category filter   */
} canon_filter;

static inline double bf16_to_double(uint16_t b) {
    uint32_t u = ((uint32_t)b) << 16;
    float f;
    memcpy(&f, &u, 4);
    return (double)f;
}

static double reduce32(double *p) {
    for (int off = 16; off >= 1; off >>= 1)
        for (int l = 0; l < off; ++l) p[l] = p[l] + p[l + off];
    return p[0];
}

static double canon_finish(double *dot, double *ss, double *qq) {
    double d = reduce32(dot), s = reduce32(ss), q = reduce32(qq);
    double nr = sqrt(s), nq = sqrt(q);
    if (nr == 0.0) nr = CANON_EPS;
    if (nq == 0.0) nq = CANON_EPS;
    return d / (nr * nq);
}

double canon_score_f32(const float *x, const double *q, int d) {
    double dot[32] = {0}, ss[32] = {0}, qq[32] = {0};
    for (int j = 0; j < d; ++j) {
        int l = j & 31;
        double xv = (double)x[j];
        dot[l] = fma(xv, q[j], dot[l]);
        ss[l] = fma(xv, xv, ss[l]);
        qq[l] = fma(q[j], q[j], qq[l]);
    }
    return canon_finish(dot, ss, qq);
}

double canon_score_bf16(const uint16_t *x, const double *q, int d) {
    double dot[32] = {0}, ss[32] = {0}, qq[32] = {0};
    for (int j = 0; j < d; ++j) {
        int l = j & 31;
        double xv = bf16_to_double(x[j]);
        dot[l] = fma(xv, q[j], dot[l]);
        ss[l] = fma(xv, xv, ss[l]);
        qq[l] = fma(q[j], q[j], qq[l]);
    }
    return canon_finish(dot, ss, qq);
}

/* scores of rows [0,n) into out[n]; dtype 0 = float32, 1 = bfloat16 */
void canon_scores(const void *vectors, int dtype, int64_t n, int d, const double *q, double *out) {
    for (int64_t r = 0; r < n; ++r)
        out[r] = dtype == 0 ? canon_score_f32((const float *)vectors + r * (int64_t)d, q, d)
                            : canon_score_bf16((const uint16_t *)vectors + r * (int64_t)d, q, d);
}

typedef struct { double score; uint64_t id; } cand_t;

static int cand_cmp(const void *a, const void *b) {
    const cand_t *x = (const cand_t *)a, *y = (const cand_t *)b;
    if (x->score > y->score) return -1;
    if (x->score < y->score) return 1;
    if (x->id < y->id) return -1;
    if (x->id > y->id) return 1;
    return 0;
}

static int row_passes(const canon_filter *f, int64_t r, const int32_t *dates, const uint8_t *cats) {
    if (dates) {
        int64_t dt = dates[r];
        if (dt < f->date_lo || dt > f->date_hi) return 0;
    }
    if (cats && f->category_mask != ~(uint64_t)0) {
        uint8_t c = cats[r];
        if (c >= 64 || !((f->category_mask >> c) & 1)) return 0;
    }
    return 1;
}

/* Filtered exact top-k for one query. ids == NULL means id = id_base + row.
 * Returns the number of results written (<= k), or -1 on allocation failure. */
int64_t canon_search(const void *vectors, int dtype, int64_t n, int d, const uint64_t *ids,
                     uint64_t id_base, const int32_t *dates, const uint8_t *cats,
                     const canon_filter *f, const double *q, int64_t k,
                     uint64_t *out_ids, double *out_scores) {
    int64_t lo = f->row_begin < 0 ? 0 : f->row_begin;
    int64_t hi = f->row_end > n ? n : f->row_end;
    if (hi <= lo || k <= 0) return 0;
    cand_t *c = (cand_t *)malloc((size_t)(hi - lo) * sizeof(cand_t));
    if (!c) return -1;
    int64_t m = 0;
    for (int64_t r = lo; r < hi; ++r) {
        if (!row_passes(f, r, dates, cats)) continue;
        c[m].score = dtype == 0 ? canon_score_f32((const float *)vectors + r * (int64_t)d, q, d)
                                : canon_score_bf16((const uint16_t *)vectors + r * (int64_t)d, q, d);
        c[m].id = ids ? ids[r] : id_base + (uint64_t)r;
        ++m;
    }
    qsort(c, (size_t)m, sizeof(cand_t), cand_cmp);
    int64_t w = m < k ? m : k;
    for (int64_t i = 0; i < w; ++i) { out_ids[i] = c[i].id; out_scores[i] = c[i].score; }
    free(c);
    return w;
}

/* ---- aggregation ------------------------------------------------------------------ */

/* months since 1970-01 of a unix timestamp, UTC, proleptic Gregorian (floor semantics) */
int32_t canon_month_index(int64_t ts) {
    int64_t days = ts / 86400;
    if (ts % 86400 < 0) --days;
    /* civil-from-days, era-based */
    int64_t z = days + 719468;
    int64_t era = (z >= 0 ? z : z - 146096) / 146097;
    int64_t doe = z - era * 146097;
    int64_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
    int64_t y = yoe + era * 400;
    int64_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
    int64_t mp = (5 * doy + 2) / 153;
    int64_t mth = mp < 10 ? mp + 3 : mp - 9;      /* 1..12 */
    if (mth <= 2) ++y;
    return (int32_t)((y - 1970) * 12 + (mth - 1));
}

typedef struct { int64_t user; int32_t category; int32_t month; int64_t sum; int64_t rows; } cell_t;

static int cell_cmp(const void *a, const void *b) {
    const cell_t *x = (const cell_t *)a, *y = (const cell_t *)b;
    if (x->user != y->user) return x->user < y->user ? -1 : 1;
    if (x->category != y->category) return x->category < y->category ? -1 : 1;
    if (x->month != y->month) return x->month < y->month ? -1 : 1;
    return 0;
}

/* Segmented sum of cents grouped by (user segment, category, month) over the users
 * [seg_lo, seg_hi). Output sorted ascending by (user, category, month); only groups that
 * have at least one row. Returns number of groups, -1 on allocation failure, or
 * -2 if out_cap is too small. */
int64_t canon_groupby(const int64_t *seg_offsets, int64_t seg_lo, int64_t seg_hi,
                      const int32_t *dates, const uint8_t *cats, const int32_t *cents,
                      int64_t out_cap, int64_t *out_user, int32_t *out_category,
                      int32_t *out_month, int64_t *out_sum, int64_t *out_rows) {
    int64_t n_out = 0;
    for (int64_t u = seg_lo; u < seg_hi; ++u) {
        int64_t lo = seg_offsets[u], hi = seg_offsets[u + 1];
        if (hi <= lo) continue;
        cell_t *c = (cell_t *)malloc((size_t)(hi - lo) * sizeof(cell_t));
        if (!c) return -1;
        for (int64_t r = lo; r < hi; ++r) {
            cell_t *e = &c[r - lo];
            e->user = u; e->category = cats[r]; e->month = canon_month_index(dates[r]);
            e->sum = cents[r]; e->rows = 1;
        }
        qsort(c, (size_t)(hi - lo), sizeof(cell_t), cell_cmp);
        int64_t i = 0;
        while (i < hi - lo) {
            int64_t j = i, s = 0;
            while (j < hi - lo && cell_cmp(&c[i], &c[j]) == 0) { s += c[j].sum; ++j; }
            if (n_out >= out_cap) { free(c); return -2; }
            out_user[n_out] = u; out_category[n_out] = c[i].category; out_month[n_out] = c[i].month;
            out_sum[n_out] = s; out_rows[n_out] = j - i;
            ++n_out; i = j;
        }
        free(c);
    }
    return n_out;
}
