// k-way merge of per-shard top-k candidate lists (the step after the NCCL all-gather of the
// sharded search). Every shard contributes, per query, a list sorted by
// (score desc, id asc); point ids are unique across shards, so the union has a strict total order
// and the global rank of an element is
//     its index in its own list + sum over the other lists of (#elements better than it),
// each term a binary search. One thread per candidate, no scratch, no sorting network:
// G * k * G * log2(k) comparisons per query (G <= 8 shards, k <= 10000).
#include "fcb_internal.cuh"

struct MergeParams {
    const unsigned char *ids, *scores, *counts;   // shard 0 base pointers; shard g at + g * stride
    int64_t stride;                               // bytes between consecutive shards' buffers
    int G, B, k;
    uint64_t *out_ids;
    double *out_scores;
    int32_t *out_counts;
};

__global__ void __launch_bounds__(256) merge_topk_kernel(const __grid_constant__ MergeParams p) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;      // candidate (g, j) = (t / k, t % k)
    int total = 0;
    for (int g = 0; g < p.G; ++g) total += reinterpret_cast<const int32_t *>(p.counts + g * p.stride)[b];
    const int count = total < p.k ? total : p.k;
    if (t == 0) p.out_counts[b] = count;
    if (t >= count && t < p.k) { p.out_ids[(size_t)b * p.k + t] = UINT64_MAX; p.out_scores[(size_t)b * p.k + t] = -INFINITY; }
    if (t >= p.G * p.k) return;
    const int g = t / p.k, j = t - g * p.k;
    const int my_n = reinterpret_cast<const int32_t *>(p.counts + g * p.stride)[b];
    if (j >= my_n) return;
    const uint64_t mi = reinterpret_cast<const uint64_t *>(p.ids + g * p.stride)[(size_t)b * p.k + j];
    const double ms = reinterpret_cast<const double *>(p.scores + g * p.stride)[(size_t)b * p.k + j];
    int rank = j;
    for (int o = 0; o < p.G && rank < p.k; ++o) {
        if (o == g) continue;
        const int n = reinterpret_cast<const int32_t *>(p.counts + o * p.stride)[b];
        const uint64_t *oi = reinterpret_cast<const uint64_t *>(p.ids + o * p.stride) + (size_t)b * p.k;
        const double *os = reinterpret_cast<const double *>(p.scores + o * p.stride) + (size_t)b * p.k;
        int lo = 0, hi = n;                                   // first index whose element is NOT better than mine
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (fcb_better64(os[mid], oi[mid], ms, mi)) lo = mid + 1; else hi = mid;
        }
        rank += lo;
    }
    if (rank < p.k) { p.out_ids[(size_t)b * p.k + rank] = mi; p.out_scores[(size_t)b * p.k + rank] = ms; }
}

extern "C" int32_t fcb_merge_topk(int32_t device, const uint64_t *cand_ids, const double *cand_scores,
                                  const int32_t *cand_counts, int64_t shard_stride_bytes, int32_t n_shards,
                                  int32_t batch, int32_t k, uint64_t *out_ids, double *out_scores,
                                  int32_t *out_counts, void *stream) {
    if (!cand_ids || !cand_scores || !cand_counts || !out_ids || !out_scores || !out_counts || n_shards < 1 ||
        n_shards > 64 || batch < 1 || batch > 65535 || k < 1 || k > FCB_MAX_K || shard_stride_bytes < 0)
        return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_merge_topk: bad argument");
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_merge_topk: cannot select device %d", device);
    MergeParams p{};
    p.ids = (const unsigned char *)cand_ids; p.scores = (const unsigned char *)cand_scores;
    p.counts = (const unsigned char *)cand_counts; p.stride = shard_stride_bytes;
    p.G = n_shards; p.B = batch; p.k = k;
    p.out_ids = out_ids; p.out_scores = out_scores; p.out_counts = out_counts;
    const int threads = n_shards * k;
    merge_topk_kernel<<<dim3((threads + 255) / 256, batch), 256, 0, (cudaStream_t)stream>>>(p);
    cudaError_t st = cudaGetLastError();
    if (st != cudaSuccess) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_merge_topk: %s", cudaGetErrorString(st));
    return FCB_OK;
}
