/* fcb200.h — C ABI of the B200-native transaction retrieval / aggregation library.
 *
 * This is the drop-in boundary for the ONE hot path of kyshu11027/financial-chatbot-llm:
 * the payload-filtered top-k lookup of tools/qdrant_tool.py and the group-by sum of
 * tools/plot_tool.py. Every entry point names the reference interface it replaces
 * (file:line in the reference tree). The reference binds that path through the
 * third-party `qdrant-client` Python package; the binding a maintainer adds is the ctypes
 * stub in INTEGRATION.md (financial-chatbot-llm_b200/_lib.py is that stub, shipped).
 *
 * Conventions
 *   - plain C types only; every function returns FCB_OK (0) or a negative FCB_ERR_* code and
 *     never throws across the ABI; fcb_last_error() returns the message for the last failure
 *     on that engine (or, with a NULL engine, for the last failed fcb_open on this thread).
 *   - the CALLER owns every corpus/query/result buffer. "device" pointers are CUDA device
 *     pointers on the engine's device (e.g. torch.Tensor.data_ptr()); they must stay alive
 *     while the shard is loaded. The library allocates only its own scratch, in fcb_open and
 *     fcb_load_shard, and frees it in fcb_close.
 *   - one call in flight per engine (not re-entrant); distinct engines may run concurrently.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream). The
 *     *_host variants take host buffers, run on the engine's own stream and return after
 *     the results are in the caller's memory.
 *   - there is no CPU fallback: without a CUDA device every call fails with FCB_ERR_CUDA.
 */
#ifndef FCB200_H
#define FCB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FCB_ABI_VERSION 1

#define FCB_OK 0
#define FCB_ERR_INVALID (-1)     /* bad argument                                   */
#define FCB_ERR_CUDA (-2)        /* a CUDA runtime call or kernel failed            */
#define FCB_ERR_NOMEM (-3)       /* scratch allocation failed                        */
#define FCB_ERR_STATE (-4)       /* no shard loaded / wrong order of calls           */
#define FCB_ERR_UNSUPPORTED (-5) /* valid request outside what this build implements */

#define FCB_DTYPE_F32 0
#define FCB_DTYPE_BF16 1

#define FCB_MAX_K 10000          /* = the tool's default and maximum limit, tools/qdrant_tool.py:46-51,145 */
#define FCB_DATE_MIN INT64_MIN
#define FCB_DATE_MAX INT64_MAX
#define FCB_ALL_CATEGORIES UINT64_MAX

typedef struct fcb_engine fcb_engine; /* opaque */

typedef struct fcb_config {
    int32_t device;     /* CUDA device ordinal                                              */
    int32_t max_batch;  /* largest `batch` fcb_search will be given (>=1)                   */
    int32_t max_k;      /* largest `k` fcb_search will be given (1..FCB_MAX_K)              */
    int32_t reserved;   /* must be 0                                                        */
} fcb_config;

/* One contiguous point range of the collection, resident on the engine's device.
 * Rows are sorted by user; a user's rows are the CSR segment [seg_offsets[u], seg_offsets[u+1])
 * in GLOBAL row numbers; this shard holds global rows [row_base, row_base + n_rows).
 * Replaces: the collection a Qdrant server holds for QDRANT_COLLECTION_NAME
 * (reference config.py:47), with the payload fields the tool filters on stored as columns:
 * metadata.user_id -> segment, metadata.date -> dates (tools/qdrant_tool.py:107,120). */
typedef struct fcb_shard {
    const void *vectors;        /* device [n_rows, dim] row-major, float32 or bfloat16; 16-byte aligned */
    const uint64_t *ids;        /* device [n_rows] point ids, or NULL: id = row_base + row              */
    const uint32_t *id_rank;    /* device [n_rows] rank of ids[row] among the shard's ids, or NULL when
                                   ids ascend with the row number (always true when ids == NULL)       */
    const int32_t *dates;       /* device [n_rows] metadata.date, unix seconds; NULL: date filters reject */
    const uint8_t *categories;  /* device [n_rows] category code 0..63; NULL: category filters reject    */
    const int32_t *cents;       /* device [n_rows] signed integer cents (aggregation only), or NULL       */
    const int64_t *seg_offsets; /* device [n_segs + 1] global CSR offsets (aggregation only), or NULL     */
    int64_t n_rows;
    int64_t n_segs;
    int64_t row_base;
    int32_t dim;                /* multiple of 8 (bf16) / 4 (f32), <= 8192                               */
    int32_t dtype;              /* FCB_DTYPE_*                                                           */
} fcb_shard;

/* The filter the tool builds at tools/qdrant_tool.py:105-131, already resolved by the host:
 *   FieldCondition(key="metadata.user_id", match=MatchValue(user_id))  -> [row_begin, row_end)
 *   FieldCondition(key="metadata.date", range=Range(gte/gt/lte/lt))    -> [date_lo, date_hi] inclusive
 *   FieldCondition(key="metadata.category", match=MatchValue/MatchAny) -> category_mask (north_star
 *   extension; the reference has no category filter). */
typedef struct fcb_filter {
    int64_t row_begin, row_end; /* global rows; empty when row_begin >= row_end (unknown user) */
    int64_t date_lo, date_hi;   /* FCB_DATE_MIN / FCB_DATE_MAX = unbounded                      */
    uint64_t category_mask;     /* bit c set = category c passes; FCB_ALL_CATEGORIES = no filter */
} fcb_filter;

int32_t fcb_abi_version(void);

/* Replaces QdrantHttpClient(url=..., api_key=...) at tools/qdrant_tool.py:24-27 (one backend
 * object per process). Allocates the engine's scratch on cfg->device. */
int32_t fcb_open(const fcb_config *cfg, fcb_engine **out);
int32_t fcb_close(fcb_engine *e);
const char *fcb_last_error(const fcb_engine *e);

/* Attach a resident shard (see fcb_shard) and compute the library-owned derived columns: float32
 * inverse row norms (4 bytes/row) and, when the shard has dates, categories and cents, the 16-bit
 * group keys of fcb_aggregate (2 bytes/row; see fcb_group_keys_build). Both are functions of the
 * caller's columns AS THEY ARE NOW: the columns must not be modified while the shard is loaded
 * (load again after a change). Replaces the server-side collection load; nothing in the reference
 * writes points. Synchronous. */
int32_t fcb_load_shard(fcb_engine *e, const fcb_shard *shard);

/* Replaces client.query_points(collection_name, query, limit, search_params,
 * query_filter).points at tools/qdrant_tool.py:147-153, for `batch` independent queries.
 *   queries      device [batch, dim] float64 (the list[float] embed_query returns, :137)
 *   filters      HOST  [batch]
 *   k            limit, 1..cfg.max_k
 *   out_ids      device [batch, k]   point ids, best first
 *   out_scores   device [batch, k]   float64 cosine scores (canonical order of operations,
 *                                    see DESIGN.md "canonical score")
 *   out_counts   device [batch]      number of valid results (<= k)
 * Exact (every filtered-in row is scored); search_params (hnsw_ef / exact, :98-101) have no
 * counterpart because nothing is approximate. Order: score descending, ties by ascending id.
 * Exact BY CONSTRUCTION: the fast kernels rank rows by a float32 (or tensor-core) first pass, re-score a bounded
 * number of candidates in float64 and then CERTIFY the answer — the exact k-th score must clear the best score any
 * dropped row could have, worst-case first-pass error included. A query that cannot be certified (more near-ties at
 * the k-th score than the candidate margin holds, e.g. many duplicated transactions) is answered again, on the same
 * stream and without a host round trip, by an all-float64 kernel (score every passing row canonically, radix-select
 * the k-th, resolve ties by id). fcb_exact_reruns() counts those. Asynchronous on `stream`. */
int32_t fcb_search(fcb_engine *e, const double *queries, const fcb_filter *filters, int32_t batch,
                   int32_t k, uint64_t *out_ids, double *out_scores, int32_t *out_counts,
                   void *stream);

/* Same call with HOST buffers for queries and results (the shape the Python tool has in hand:
 * a list of floats in, scored points out). Copies in, searches, copies out, synchronises. */
int32_t fcb_search_host(fcb_engine *e, const double *queries, const fcb_filter *filters,
                        int32_t batch, int32_t k, uint64_t *out_ids, double *out_scores,
                        int32_t *out_counts);

/* No counterpart in the reference (nothing there allocates GPU memory at all). The same
 * filtered top-k as fcb_search, for ONE query, for callers that may not allocate: a worker
 * sharing a GPU whose memory a resident corpus owner has already filled, or a request thread that
 * must not meet the device-wide synchronisation of cudaMalloc/cudaFree. Needs no engine: the
 * caller passes the shard description (HOST struct, device pointers inside), the query
 * (device, float64 [dim]), the filter (HOST) and the three output buffers (device: k ids,
 * k float64 scores, 1 count). Makes no allocation on host or device; its scratch is a
 * 262,400-byte array inside the library image that every GPU loading the library keeps until
 * the process exits. Limits: k <= 16, dim <= 8192, one query per call — larger requests are
 * refused with FCB_ERR_UNSUPPORTED. All calls on a GPU share the scratch: calls that can
 * overlap on the device race (serialise them on one stream). The first call on each device
 * resolves the scratch address and should happen before any stream capture begins. Errors are
 * described in the caller's `err` buffer (may be NULL). Asynchronous on `stream`. */
int32_t fcb_search_noalloc(int32_t device, const fcb_shard *shard, const double *query,
                           const fcb_filter *filter, int32_t k, uint64_t *out_ids,
                           double *out_scores, int32_t *out_count, void *stream, char *err,
                           int32_t err_len);

/* Duration in milliseconds of the device work of the last fcb_search_host call (CUDA events on
 * the engine's stream around the kernels, excluding the copies); < 0 if none. Measurement aid. */
float fcb_last_search_ms(const fcb_engine *e);

/* Batched tensor-core path only (bf16 shard, batch >= 32, identical filters): number of queries of
 * the last call that could not be answered exactly from the first pass's candidate lists and that the exact scan
 * answered instead; -1 if the path is not active. Synchronises the device. Measurement aid. */
int32_t fcb_last_dense_redo(fcb_engine *e, int32_t batch);

/* Same path: number of queries of the last call that were answered by re-scoring more than the 32 best listed rows —
 * every listed row inside the worst-case error window of the bf16 first pass (csrc/fcb_dense.cu). Measurement aid. */
int32_t fcb_last_dense_widened(fcb_engine *e, int32_t batch);

/* Diagnostic for the batched tensor-core path: the next calls also dump the raw accumulators
 * (un-normalised bf16 dot products) of the first `rows` rows of the segment into `out`
 * (device float32 [rows, 256]); rows = 0 switches the dump off. Not for production use. */
int32_t fcb_debug_dense_dump(float *out, int64_t rows);

/* Diagnostic for the batched tensor-core path (cta_group::2 kernel): while `buf` (device int64[10]) is
 * set, CTA 0 writes per-role cycle counters: [0] producer total, [1] waiting for a free stage, [2] MMA
 * issuer total, [3] waiting for drained accumulators, [4] waiting for a full stage, [5] epilogue total,
 * [6] waiting for accumulators, [7] scanning TMEM, [8] folding candidates, [9] super-tiles. NULL = off. */
int32_t fcb_debug_dense_prof(int64_t *buf);

/* Diagnostic for the fused scan: while `buf` (device uint64[9]) is set, every launch writes the GPU
 * global timer (ns) at nine phase boundaries of query 0 (kernel start, query staged, scan done,
 * CTA list written, last CTA starts, lists merged, re-score starts / ends, results written).
 * NULL switches it off. Not for production use. */
int32_t fcb_debug_scan_timing(uint64_t *buf);

/* The loop shape of the fused scan kernel in use for this engine's shard: rows per warp step (*r) and unroll of the
 * column loop (*u); fcb_load_shard measures the candidates on shards of at least 1M rows and keeps the fastest (every
 * shape computes bit-identical results). Both 0: not measured (small shard, FCB_SCAN_TUNE=0 or FCB_SCAN_CFG set) — the
 * built-in table decides per call. Measurement aid; no counterpart in the reference. */
int32_t fcb_scan_config(const fcb_engine *e, int32_t *r, int32_t *u);

/* How many queries since fcb_load_shard the certificate of a fast path refused and the exact float64 kernel
 * answered instead (see fcb_search); -1 without a shard; -2 (and fcb_last_error) if that kernel ever failed to assemble its
 * grid because other work saturated the GPU — its waits are bounded, the answers of that call are then not valid.
 * Synchronises the device. Measurement aid; no counterpart in
 * the reference, whose local mode is brute force in float64 throughout. */
int64_t fcb_exact_reruns(fcb_engine *e);

/* Number of kernels the last fcb_search / fcb_search_host call launched. Measurement aid. */
int32_t fcb_last_search_launches(const fcb_engine *e);

/* Replaces transactions.groupby(plot_config.group_by)[plot_config.y_axis].sum() at
 * tools/plot_tool.py:45 (bar) and :52 (pie), over the shard's device columns, for the key
 * (user, category, UTC calendar month) of BASELINE.json config 5. Exact int64 sums of int32 cents.
 *   seg_begin/end   user (CSR segment) range; rows outside this shard's row range are skipped
 *   month_lo        first month of the dense window, in months since 1970-01 (UTC)
 *   n_months        1..240; n_categories 1..64; n_months * n_categories <= 4096
 *   sums            device int64  [seg_end-seg_begin, n_categories, n_months]   (zeroed by the call)
 *   present         device uint8  same shape: 1 where the group has at least one row — pandas
 *                   lists observed groups only, in ascending key order = this dense order
 *   dropped         device uint64 [1]: rows whose category/month fall outside the window (must
 *                   read 0 for the result to be the complete group-by)
 * Asynchronous on `stream`. */
int32_t fcb_aggregate(fcb_engine *e, int64_t seg_begin, int64_t seg_end, int32_t month_lo,
                      int32_t n_months, int32_t n_categories, int64_t *sums, uint8_t *present,
                      uint64_t *dropped, void *stream);

/* The same group-by over caller-owned device columns, without an engine (nothing is allocated, no vectors are needed):
 * what the tool-shaped path uses for the rows of ONE transactions_json (tools/plot_tool.py:29,45,52 — one key column that
 * becomes the CSR segments, one value column) and what a caller holding only the payload columns uses at scale.
 *   dates / categories / cents   device [n_rows] as in fcb_shard; seg_offsets device [n_segs + 1] global CSR offsets;
 *                                aligned: cents / dates 16 bytes, categories 4 (FCB_ERR_INVALID otherwise)
 *   row_base                     global row number of the columns' first row
 *   dates_sorted                 1 if the dates are non-decreasing within every user, else 0. A hint only: the result is the
 *                                same either way, and since round 2 the default kernel is the same too (the ordered-log
 *                                kernel it used to select is kept behind FCB_AGG_KERNEL=sorted)
 * Other arguments and outputs as for fcb_aggregate. Errors: fcb_last_error(NULL). Asynchronous on `stream`. */
int32_t fcb_aggregate_columns(int32_t device, const int32_t *dates, const uint8_t *categories, const int32_t *cents,
                              const int64_t *seg_offsets, int64_t n_rows, int64_t n_segs, int64_t row_base,
                              int32_t dates_sorted, int64_t seg_begin, int64_t seg_end, int32_t month_lo, int32_t n_months,
                              int32_t n_categories, int64_t *sums, uint8_t *present, uint64_t *dropped, void *stream);

/* Group keys: the (category, month) of every row resolved once, so that the group-by reads 6 bytes per row (amount + key)
 * instead of 9 and spends no arithmetic on calendars. plot_tool.py:45 derives the month of every row on every call
 * (`dt.to_period("M")`); a resident table does it once. fcb_load_shard builds the keys of an engine's shard by itself
 * (fcb_aggregate then uses them); these two entry points do the same for caller-owned columns.
 *   keys       device [n_rows] uint16, 16-byte aligned, caller-allocated (n_rows * 2 bytes rounded up to 16)
 *   key_base   out: first month (months since 1970-01) the keys can express; they cover [key_base, key_base + 511),
 *              ending at the newest month of the data
 * fcb_group_keys_build waits for `stream` (the month range of the data comes back to the host). */
int32_t fcb_group_keys_build(int32_t device, const int32_t *dates, const uint8_t *categories, const int32_t *cents,
                             int64_t n_rows, uint16_t *keys, int32_t *key_base, void *stream);

/* fcb_aggregate_columns with the rows' group keys. Windows inside [key_base, key_base + 511) are aggregated from cents +
 * keys (any row order, one launch); other windows fall back on dates / categories, which may be NULL if the caller never
 * asks for such a window (FCB_ERR_UNSUPPORTED otherwise). keys == NULL: exactly fcb_aggregate_columns. Same outputs,
 * bit for bit, whichever way the rows are read. */
int32_t fcb_aggregate_columns_keyed(int32_t device, const uint16_t *keys, int32_t key_base, const int32_t *dates,
                                    const uint8_t *categories, const int32_t *cents, const int64_t *seg_offsets,
                                    int64_t n_rows, int64_t n_segs, int64_t row_base, int32_t dates_sorted,
                                    int64_t seg_begin, int64_t seg_end, int32_t month_lo, int32_t n_months,
                                    int32_t n_categories, int64_t *sums, uint8_t *present, uint64_t *dropped, void *stream);

/* Group-by over the points a search just returned: the reference's agent feeds the
 * output of retrieve_transactions (tools/qdrant_tool.py:147-170) back through JSON into
 * transactions.groupby(...)[...].sum() (tools/plot_tool.py:29,45,52); here the ids fcb_search wrote stay on
 * the device and their rows' cents are summed per (category, UTC calendar month) directly.
 *   hit_ids / hit_counts   device outputs of fcb_search for `batch` queries with this `k` (same stream, or ordered
 *                          after it)
 *   sums / present         device int64 / uint8 [batch, n_categories, n_months], written densely (see fcb_aggregate)
 *   stats                  device uint64 [2]: [0] returned points whose category / month fall outside the window,
 *                          [1] returned points that are not rows of this shard (a sharded search returns points of
 *                          other shards too: each rank sums its own, the per-rank sums add up)
 * Window limits as for fcb_aggregate. When the shard's ids do not ascend with the rows (id_rank given), the first
 * call builds a 4-byte-per-row inverse map: it allocates and synchronises once. Asynchronous on `stream`. */
int32_t fcb_aggregate_hits(fcb_engine *e, const uint64_t *hit_ids, const int32_t *hit_counts, int32_t batch,
                           int32_t k, int32_t month_lo, int32_t n_months, int32_t n_categories, int64_t *sums,
                           uint8_t *present, uint64_t *stats, void *stream);

/* The exchange step of the sharded search (north_star: "each shard computes a local top-k and the
 * k x shards candidates are merged after a small NCCL allgather"). No counterpart in the
 * reference, whose Qdrant server merges its own segments internally. Inputs are the
 * all-gathered per-shard results of fcb_search for the same batch: shard g's ids / scores /
 * counts start `shard_stride_bytes` * g bytes after shard 0's (device pointers). Each list is
 * sorted (score desc, id asc) with `counts[b]` valid entries; ids are unique across shards.
 * Writes the global top-k in the same order and format as fcb_search. Needs no engine and no
 * scratch. Asynchronous on `stream`. */
int32_t fcb_merge_topk(int32_t device, const uint64_t *cand_ids, const double *cand_scores,
                       const int32_t *cand_counts, int64_t shard_stride_bytes, int32_t n_shards,
                       int32_t batch, int32_t k, uint64_t *out_ids, double *out_scores,
                       int32_t *out_counts, void *stream);

/* ---- Owner-computes exchange of the sharded search (north_star's sharded corpus; no counterpart in the reference, whose
 * Qdrant server routes and merges internally). With the shard cuts on user boundaries a tenant lives on ONE GPU, so its
 * query needs no merge: the owner computes the final top-k and hands it to every other rank. fcb_search_routed does both in
 * one call per rank and batch: the fused scan over the queries THIS rank owns, then one kernel that writes their results
 * straight into the exchange buffer of every rank over NVLink peer mappings (plain stores + a stamp) and waits until its
 * own buffer holds every query's answer. No NCCL call, no merge launch, no host round trip on the data path.
 *
 * Exchange buffers: every rank allocates one with fcb_peer_alloc (fcb_exchange_bytes(slots, k) bytes), the 64-byte
 * handles travel by any side channel (torch.distributed all_gather_object in sharded.py), and every rank opens the
 * others' with fcb_peer_open. Layout of one buffer, two parities p = seq & 1 at byte p * fcb_exchange_bytes / 2:
 *   ids uint64 [slots][k] | scores float64 [slots][k] | counts int32 [slots] | stamps uint32 [slots]. */
#define FCB_MAX_WORLD 8
#define FCB_PEER_HANDLE_BYTES 64
typedef struct fcb_exchange {
    int32_t world, rank;                /* ranks of the box (<= FCB_MAX_WORLD) and this process's                     */
    void *bufs[FCB_MAX_WORLD];          /* bufs[g]: rank g's exchange buffer as mapped in THIS process (device pointer) */
    const int64_t *cuts;                /* HOST [world + 1]: rank g holds global rows [cuts[g], cuts[g + 1])           */
} fcb_exchange;

int64_t fcb_exchange_bytes(int32_t slots, int32_t k);
int32_t fcb_peer_alloc(int32_t device, int64_t bytes, void **ptr, uint8_t *handle /* [FCB_PEER_HANDLE_BYTES] */);
int32_t fcb_peer_open(int32_t device, const uint8_t *handle, void **ptr);
int32_t fcb_peer_close(int32_t device, void *ptr, int32_t owned /* 1: from fcb_peer_alloc, 0: from fcb_peer_open */);

/* The same call on every rank, with the same queries (device [batch, dim] float64, resident on every rank), filters (HOST
 * [batch]), k (<= 16) and sequence number `seq` (1, 2, 3, ... per exchange; it selects the parity and is the stamp). A query
 * whose row range lies in rank g's shard is scanned by rank g only; a range that spans two shards is refused with
 * FCB_ERR_UNSUPPORTED (answer those with fcb_search + all-gather + fcb_merge_topk). When the work enqueued by this call has
 * finished on `stream`, parity seq & 1 of bufs[rank] holds all `batch` answers; *status (device int32, zeroed by the caller)
 * reads 1 if a peer never delivered (the wait is bounded: a few seconds). Consume the results on the same stream before the
 * next call (that is what keeps a fast peer from overwriting them, see csrc/fcb_exchange.cu). Asynchronous on `stream`. */
int32_t fcb_search_routed(fcb_engine *e, const fcb_exchange *x, const double *queries, const fcb_filter *filters,
                          int32_t batch, int32_t k, uint32_t seq, int32_t *status, void *stream);

/* Test hook for the exchange of fcb_search_routed: bit 0 = push this rank's results, bit 1 = wait for every rank's (3 = both,
 * the production setting). Kernels of different ranks that wait on one another must each have their own GPU; a single-GPU
 * test plays several ranks with push-only exchanges (mask 1) and inspects the buffers. Not for production use. */
int32_t fcb_debug_exchange_phases(int32_t mask);

/* 1 if fcb_load_shard found metadata.date non-decreasing within every user segment (a time-ordered
 * transaction log), 0 if not, -1 without a shard. fcb_aggregate then uses its atomic-free kernel; the
 * result is the same either way. */
int32_t fcb_shard_dates_sorted(const fcb_engine *e);

/* Synthetic corpus (no counterpart in the reference, which has no ingestion code or data —
 * SURVEY.md §2 row 13, §8d): counter-based integer generator, bit-identical to
 * financial-chatbot-llm_b200/synth.py, writing straight into caller-owned device buffers. */
int32_t fcb_synth_vectors(int32_t device, void *out, int32_t dtype, int64_t row_lo, int64_t n_rows,
                          int32_t dim, uint64_t seed, int32_t n_clusters, void *stream);
int32_t fcb_synth_columns(int32_t device, int32_t *dates, uint8_t *categories, int32_t *cents,
                          int64_t row_lo, int64_t n_rows, uint64_t seed, int32_t n_categories,
                          int64_t date_span_seconds, void *stream);

/* Variant of the date column for time-ordered logs: row r of user segment [a, b) gets
 * DATE_BASE + floor((r - a) * span / (b - a)) (same integer recipe as synth.dates_sorted). */
int32_t fcb_synth_dates_sorted(int32_t device, int32_t *dates, const int64_t *seg_offsets, int64_t n_segs,
                               int64_t row_lo, int64_t n_rows, int64_t date_span_seconds, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* FCB200_H */
