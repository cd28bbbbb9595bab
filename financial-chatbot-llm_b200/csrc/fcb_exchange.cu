// Owner-computes exchange of the sharded search (BASELINE config 4: tenants sharded by point range over the GPUs of one
// box). With shard cuts on user boundaries every tenant lives on ONE GPU, so a query needs no merge: its owner computes
// the final top-k and only has to hand it to the other ranks. NCCL would do that with an all-gather (~15-20 us of launch
// and protocol latency per batch) followed by a merge launch; here ONE kernel per rank does it over peer memory:
//   push   every rank writes the results of the queries it owns straight into the exchange buffer of EVERY rank
//          (plain 8-byte stores through NVLink peer mappings), fences, then stamps each slot with the call's sequence number;
//   wait   the same kernel then waits (bounded) until every slot of its OWN buffer carries that stamp.
// When the kernel retires, the rank holds the answers of the whole batch. No collective call, no merge, no host round
// trip. Peer buffers are plain cudaMalloc allocations opened in the other processes through CUDA IPC; only ordinary
// loads and stores touch them (TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load
// from another GPU's memory hung the GPU).
//
// Buffer of one rank, two parities (sequence number & 1) so that the pushes of call s + 1 cannot land on data the owner
// is still reading from call s:   [parity][ ids u64 [slots][k] | scores f64 [slots][k] | counts i32 [slots] | stamps u32 [slots] ]
// Flow control: a rank's pushes of call s + 2 (same parity as s) are issued by its exchange kernel s + 2, which is stream-
// ordered after its kernel s + 1, which waited for every peer's pushes s + 1, which those peers issued after whatever they
// enqueued to consume results s. So: consume the results on the SAME stream, before the next call.
#include <string.h>

#include "fcb_internal.cuh"

#define XCH_THREADS 256
#define XCH_SPIN_LIMIT (1u << 24)           // x ~0.2 us of back-off: a missing peer gives up after a few seconds

struct ExchangeParams {
    unsigned char *bufs[FCB_MAX_WORLD];    // rank g's exchange buffer as mapped in this process
    int world, rank, slots, k;
    unsigned seq;
    const int32_t *owner;                  // device [slots]
    const uint64_t *ids;                   // local results of the search (only the owned slots are valid)
    const double *scores;
    const int32_t *counts;
    int32_t *status;                       // device: set to 1 if a peer never delivered
    int phases;                            // bit 0: push, bit 1: wait (both in production; see fcb_debug_exchange_phases)
};

__global__ void __launch_bounds__(XCH_THREADS) exchange_kernel(const __grid_constant__ ExchangeParams p) {
    const int lane = threadIdx.x & 31;
    const int slot = blockIdx.x * (XCH_THREADS / 32) + (threadIdx.x >> 5);           // one warp per slot
    if (slot >= p.slots) return;
    const size_t par = (size_t)(p.seq & 1u) * fcb_exchange_parity_bytes(p.slots, p.k);
    const size_t off_scores = (size_t)p.slots * p.k * 8, off_counts = 2 * off_scores, off_stamps = off_counts + (size_t)p.slots * 4;
    if ((p.phases & 1) && p.owner[slot] == p.rank) {
        // ---- push: k ids + k scores (+ the count) of this slot to every rank, then the stamp
        const int n = 2 * p.k;
        for (int g = 0; g < p.world; ++g) {
            unsigned char *dst = p.bufs[g] + par;
            uint64_t *d_ids = reinterpret_cast<uint64_t *>(dst) + (size_t)slot * p.k;
            uint64_t *d_sc = reinterpret_cast<uint64_t *>(dst + off_scores) + (size_t)slot * p.k;
            for (int i = lane; i < n; i += 32) {
                if (i < p.k) d_ids[i] = p.ids[(size_t)slot * p.k + i];
                else d_sc[i - p.k] = reinterpret_cast<const uint64_t *>(p.scores)[(size_t)slot * p.k + i - p.k];
            }
            if (lane == 0) reinterpret_cast<int32_t *>(dst + off_counts)[slot] = p.counts[slot];
        }
        __threadfence_system();                                                     // this lane's stores, system-wide ...
        __syncwarp();                                                               // ... of every lane, before any stamp
        if (lane < p.world) {
            volatile unsigned *stamp = reinterpret_cast<volatile unsigned *>(p.bufs[lane] + par + off_stamps) + slot;
            *stamp = p.seq;
        }
    }
    // ---- wait: this slot of MY buffer must carry the stamp (its owner may be me)
    if ((p.phases & 2) && lane == 0) {
        volatile unsigned *mine = reinterpret_cast<volatile unsigned *>(p.bufs[p.rank] + par + off_stamps) + slot;
        unsigned spins = 0;
        while (*mine != p.seq) {
            if (++spins > XCH_SPIN_LIMIT) { *p.status = 1; break; }
            __nanosleep(200);
        }
        __threadfence_system();                                                     // acquire: the data stores precede the stamp
    }
}

static int g_exchange_phases = 3;

// Test hook: which halves of the exchange kernel run (bit 0: push this rank's results to every rank, bit 1: wait for
// everybody's). Production = 3. Kernels that wait on one another must not be run as separate launches on ONE GPU (nothing
// guarantees that they are resident together), so the single-GPU test plays two ranks with push-only exchanges and then
// inspects both buffers; the wait half runs wherever every rank has its own GPU.
extern "C" int32_t fcb_debug_exchange_phases(int32_t mask) {
    g_exchange_phases = mask & 3;
    return FCB_OK;
}

int32_t fcb_exchange_launch(fcb_engine *e, const fcb_exchange *x, int32_t slots, int32_t k, uint32_t seq, const int32_t *d_owner,
                            const uint64_t *d_ids, const double *d_scores, const int32_t *d_counts, int32_t *d_status,
                            cudaStream_t st) {
    ExchangeParams p{};
    for (int g = 0; g < x->world; ++g) p.bufs[g] = reinterpret_cast<unsigned char *>(x->bufs[g]);
    p.world = x->world; p.rank = x->rank; p.slots = slots; p.k = k; p.seq = seq;
    p.owner = d_owner; p.ids = d_ids; p.scores = d_scores; p.counts = d_counts; p.status = d_status;
    p.phases = g_exchange_phases;
    const int warps = XCH_THREADS / 32;
    exchange_kernel<<<(slots + warps - 1) / warps, XCH_THREADS, 0, st>>>(p);
    FCB_CUDA(e, cudaGetLastError());
    return FCB_OK;
}

// ---- peer buffers (CUDA IPC) ---------------------------------------------------------------------

extern "C" int64_t fcb_exchange_bytes(int32_t slots, int32_t k) {
    if (slots < 1 || k < 1) return -1;
    return (int64_t)(2 * fcb_exchange_parity_bytes(slots, k));
}

extern "C" int32_t fcb_peer_alloc(int32_t device, int64_t bytes, void **ptr, uint8_t *handle) {
    if (!ptr || !handle || bytes <= 0) return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_peer_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == FCB_PEER_HANDLE_BYTES, "handle size");
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_peer_alloc: cannot select device %d", device);
    void *p = nullptr;
    if (cudaMalloc(&p, (size_t)bytes) != cudaSuccess) {
        cudaGetLastError();
        return fcb_fail(nullptr, FCB_ERR_NOMEM, "fcb_peer_alloc: cannot allocate %lld bytes", (long long)bytes);
    }
    cudaError_t st = cudaMemset(p, 0, (size_t)bytes);               // stamps start at 0: sequence numbers start at 1
    cudaIpcMemHandle_t h;
    if (st == cudaSuccess) st = cudaIpcGetMemHandle(&h, p);
    if (st != cudaSuccess) {
        cudaFree(p);
        return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_peer_alloc: %s", cudaGetErrorString(st));
    }
    memcpy(handle, &h, sizeof h);
    *ptr = p;
    return FCB_OK;
}

extern "C" int32_t fcb_peer_open(int32_t device, const uint8_t *handle, void **ptr) {
    if (!ptr || !handle) return fcb_fail(nullptr, FCB_ERR_INVALID, "fcb_peer_open: bad argument");
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_peer_open: cannot select device %d", device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    void *p = nullptr;
    const cudaError_t st = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (st != cudaSuccess) {
        cudaGetLastError();
        return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_peer_open: cudaIpcOpenMemHandle: %s", cudaGetErrorString(st));
    }
    *ptr = p;
    return FCB_OK;
}

extern "C" int32_t fcb_peer_close(int32_t device, void *ptr, int32_t owned) {
    if (!ptr) return FCB_OK;
    FcbDeviceGuard g(device);
    if (!g.ok) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_peer_close: cannot select device %d", device);
    const cudaError_t st = owned ? cudaFree(ptr) : cudaIpcCloseMemHandle(ptr);
    if (st != cudaSuccess) return fcb_fail(nullptr, FCB_ERR_CUDA, "fcb_peer_close: %s", cudaGetErrorString(st));
    return FCB_OK;
}
