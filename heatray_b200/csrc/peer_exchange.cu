// peer_exchange.cu — the one exchange step of the multi-GPU path as ONE kernel over NVLink peer memory.
//
// One process per GPU. Every rank owns a snapshot buffer of its fp32 accumulation buffer and a small block of flags; both are
// exported with CUDA IPC (the handles travel through the caller's process group: plumbing) and mapped by every peer. The exchange
// of a frame is then a single launch per rank of k_peer_exchange:
//   1. arrive: the rank tells every peer that its snapshot is in place (release store, system scope) and waits for all of them;
//   2. work, sample-index sharding (SUM): rank r sums slice r of the frame over all ranks' snapshots, read straight out of the
//      peers' memory in rank order 0..N-1 (a fixed order: the result does not depend on who computes it or when), and stores the
//      sums into rank 0's result buffer — the reduce-scatter and the gather onto the root of the usual two-collective recipe, fused;
//      screen-band sharding (BANDS): the rank stores the rows it rendered into rank 0's result buffer (disjoint rows, no arithmetic,
//      bit-identical to the single-GPU frame);
//   3. done: the rank's last block tells every peer that it has finished reading and writing, and leaves when all peers have.
// At kernel end on rank 0 the combined frame is complete in its result buffer, in stream order; at kernel end on any rank its
// snapshot buffer is free for the next exchange. No NCCL call, no host synchronisation, N x (N-1) / N frame bytes over NVLink.
//
// Spins are bounded (kSpinLimitCycles): a peer that never arrives sets a sticky error instead of hanging the GPU.
#include "../../include/heatray_cuda.h"

#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace {

constexpr int kMaxRanks = HRC_EXCHANGE_MAX_RANKS;
constexpr long long kSpinLimitCycles = 6000000000ll;      // ~3 s at 2 GHz

struct Flags {                       // one per rank, in that rank's memory, written by its peers
    uint32_t arrive[kMaxRanks];      // arrive[p]: last epoch for which rank p's snapshot is in place
    uint32_t done[kMaxRanks];        // done[p]: last epoch rank p has finished
    uint32_t blocksDone;             // local: blocks of this rank's kernel that finished their work
    uint32_t error;                  // local, sticky: 1 = a peer did not arrive in time, 2 = a peer did not finish in time
};

struct PeerArgs {
    const float4* snap[kMaxRanks];   // every rank's snapshot buffer (snap[rank] is local)
    Flags* flags[kMaxRanks];         // every rank's flag block (flags[rank] is local)
    float4* result;                  // rank 0's result buffer
    uint64_t count;                  // float4 per frame
    uint32_t epoch;
    int rank, world;
    int width, bandHeight;           // BANDS only
};

__device__ __forceinline__ void storeRelease(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t loadAcquire(const uint32_t* p) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ float4 loadPeer(const float4* p) {      // peer data is read once, after the arrival barrier (acquire, system scope)
    float4 v; asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ bool reached(uint32_t seen, uint32_t epoch) { return (int32_t)(seen - epoch) >= 0; }

// Waits until word[p] has reached `epoch` for every rank p; false on timeout.
__device__ bool waitAll(const uint32_t* word, int world, uint32_t epoch) {
    const long long t0 = clock64();
    for (int p = 0; p < world; ++p)
        while (!reached(loadAcquire(word + p), epoch)) {
            if (clock64() - t0 > kSpinLimitCycles) return false;
            __nanosleep(64);
        }
    return true;
}

template <bool BANDS>
__global__ void __launch_bounds__(256) k_peer_exchange(const PeerArgs a) {
    Flags* mine = a.flags[a.rank];
    __shared__ int ok;
    if (threadIdx.x == 0) {
        // the grid is small enough to be resident as a whole (runExchange), so block 0 is running: it alone signals, every block waits
        if (blockIdx.x == 0) for (int p = 0; p < a.world; ++p) storeRelease(&a.flags[p]->arrive[a.rank], a.epoch);
        ok = waitAll(mine->arrive, a.world, a.epoch) ? 1 : 0;
        if (!ok) mine->error = 1u;
    }
    __syncthreads();
    if (ok) {
        if (BANDS) {
            // the rows this rank rendered, y with (y / bandHeight) % world == rank, straight from the caller's buffer into rank 0's result:
            // own row k (k-th row of this rank, bottom up) is frame row ((k / bandHeight) * world + rank) * bandHeight + k % bandHeight
            const uint32_t w = (uint32_t)a.width, bh = (uint32_t)a.bandHeight, height = (uint32_t)(a.count / w);
            const uint32_t period = bh * (uint32_t)a.world, first = bh * (uint32_t)a.rank;
            const uint32_t full = height / period, rest = height % period;
            const uint32_t ownRows = full * bh + (rest > first ? min(rest - first, bh) : 0u);
            const uint64_t items = (uint64_t)ownRows * w, stride = (uint64_t)gridDim.x * blockDim.x;
            const float4* __restrict__ src = a.snap[a.rank];
            for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < items; j += stride) {
                const uint32_t k = (uint32_t)(j / w), x = (uint32_t)(j - (uint64_t)k * w);
                const uint64_t i = (uint64_t)((k / bh) * period + first + k % bh) * w + x;
                a.result[i] = __ldcs(src + i);
            }
        } else {
            // slice `rank` of the frame, summed over the ranks in rank order
            const uint64_t per = (a.count + a.world - 1) / a.world;
            const uint64_t lo = per * a.rank, hi = lo + per < a.count ? lo + per : a.count;
            const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
            uint64_t i = lo + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
            for (; i < hi; i += stride) {
                float4 v[kMaxRanks];
#pragma unroll
                for (int p = 0; p < kMaxRanks; ++p) if (p < a.world) v[p] = loadPeer(a.snap[p] + i);      // all loads in flight before the first add
                float4 s = v[0];
#pragma unroll
                for (int p = 1; p < kMaxRanks; ++p) if (p < a.world) { s.x += v[p].x; s.y += v[p].y; s.z += v[p].z; s.w += v[p].w; }
                a.result[i] = s;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();                                   // this block's stores into rank 0's memory, before the count
        const uint32_t prev = atomicAdd(&mine->blocksDone, 1u);
        if (prev == gridDim.x - 1) {                              // last block of this rank
            mine->blocksDone = 0u;
            __threadfence_system();
            for (int p = 0; p < a.world; ++p) storeRelease(&a.flags[p]->done[a.rank], a.epoch);
            if (!waitAll(mine->done, a.world, a.epoch) && !mine->error) mine->error = 2u;
        }
    }
}

} // namespace

struct HrcExchange_t {
    int device = 0, rank = 0, world = 1;
    uint64_t count = 0;
    float4* snap = nullptr; float4* result = nullptr; Flags* flags = nullptr;      // local allocations (result: rank 0 only)
    void* peerSnap[kMaxRanks] = {}; void* peerFlags[kMaxRanks] = {}; void* peerResult = nullptr;      // opened IPC mappings (nullptr for the own rank)
    bool connected = false;
    uint32_t epoch = 0;
    int smCount = 148;
    std::string err;
};

namespace {
thread_local std::string g_exErr;
int exFail(HrcExchange ex, int code, const std::string& msg) { (ex ? ex->err : g_exErr) = msg; return code; }
int exCuda(HrcExchange ex, cudaError_t e, const char* what) { return exFail(ex, e == cudaErrorMemoryAllocation ? HRC_ERR_OUT_OF_MEMORY : HRC_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e)); }
#define EXCU(ex, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return exCuda(ex, e_, #call); } while (0)
struct Handles { cudaIpcMemHandle_t snap, flags, result; };
static_assert(sizeof(Handles) == HRC_EXCHANGE_HANDLE_BYTES, "three 64-byte IPC handles");
}

HRC_API const char* hrcExchangeGetLastError(HrcExchange ex) { return ex ? ex->err.c_str() : g_exErr.c_str(); }

HRC_API int hrcExchangeCreate(int32_t device, int32_t rank, int32_t world, uint64_t float4Count, HrcExchange* out, uint8_t* handle) {
    if (!out || !handle) return exFail(nullptr, HRC_ERR_INVALID_VALUE, "null pointer");
    if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return exFail(nullptr, HRC_ERR_INVALID_VALUE, "rank / world out of range (at most HRC_EXCHANGE_MAX_RANKS ranks)");
    if (float4Count == 0) return exFail(nullptr, HRC_ERR_INVALID_VALUE, "empty frame");
    EXCU(nullptr, cudaSetDevice(device));
    HrcExchange ex = new HrcExchange_t();
    ex->device = device; ex->rank = rank; ex->world = world; ex->count = float4Count;
    cudaDeviceGetAttribute(&ex->smCount, cudaDevAttrMultiProcessorCount, device);
    Handles h; std::memset(&h, 0, sizeof(h));
    cudaError_t e = cudaMalloc(&ex->snap, float4Count * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&ex->flags, sizeof(Flags));
    if (e == cudaSuccess) e = cudaMemset(ex->flags, 0, sizeof(Flags));
    if (e == cudaSuccess && rank == 0) e = cudaMalloc(&ex->result, float4Count * sizeof(float4));
    if (e == cudaSuccess && rank == 0) e = cudaMemset(ex->result, 0, float4Count * sizeof(float4));
    if (e == cudaSuccess && world > 1) e = cudaIpcGetMemHandle(&h.snap, ex->snap);
    if (e == cudaSuccess && world > 1) e = cudaIpcGetMemHandle(&h.flags, ex->flags);
    if (e == cudaSuccess && world > 1 && rank == 0) e = cudaIpcGetMemHandle(&h.result, ex->result);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { int r = exCuda(nullptr, e, "hrcExchangeCreate"); cudaFree(ex->snap); cudaFree(ex->flags); cudaFree(ex->result); delete ex; return r; }
    std::memcpy(handle, &h, sizeof(h));
    *out = ex;
    return HRC_OK;
}

HRC_API int hrcExchangeConnect(HrcExchange ex, const uint8_t* allHandles) {
    if (!ex || !allHandles) return exFail(ex, HRC_ERR_INVALID_VALUE, "null pointer");
    if (ex->connected) return exFail(ex, HRC_ERR_INVALID_OPERATION, "already connected");
    EXCU(ex, cudaSetDevice(ex->device));
    for (int p = 0; p < ex->world; ++p) {
        if (p == ex->rank) continue;
        Handles h; std::memcpy(&h, allHandles + (size_t)p * sizeof(Handles), sizeof(h));
        // ranks on different devices need peer access; two ranks on one device (tests) share it anyway
        cudaError_t e = cudaIpcOpenMemHandle(&ex->peerSnap[p], h.snap, cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&ex->peerFlags[p], h.flags, cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess && p == 0) e = cudaIpcOpenMemHandle(&ex->peerResult, h.result, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return exCuda(ex, e, "cudaIpcOpenMemHandle (the ranks must be on GPUs of one node with peer access)");
    }
    ex->connected = true;
    return HRC_OK;
}

static int runExchange(HrcExchange ex, const float* src, bool bands, int width, int bandHeight, void* stream) {
    if (!ex || !src) return exFail(ex, HRC_ERR_INVALID_VALUE, "null pointer");
    if (ex->world > 1 && !ex->connected) return exFail(ex, HRC_ERR_INVALID_OPERATION, "hrcExchangeConnect first");
    if (bands && (width <= 0 || bandHeight <= 0 || ex->count % (uint64_t)width)) return exFail(ex, HRC_ERR_INVALID_VALUE, "bad band geometry");
    EXCU(ex, cudaSetDevice(ex->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // SUM: the peers read this rank's frame, so it goes into the exported snapshot buffer first (the caller's buffer is not mapped by anyone).
    // BANDS: nobody reads it but this rank's own kernel: no snapshot, the kernel reads the caller's buffer.
    if (!bands) EXCU(ex, cudaMemcpyAsync(ex->snap, src, ex->count * sizeof(float4), cudaMemcpyDeviceToDevice, s));
    PeerArgs a; std::memset(&a, 0, sizeof(a));
    for (int p = 0; p < ex->world; ++p) {
        a.snap[p] = static_cast<const float4*>(p == ex->rank ? ex->snap : ex->peerSnap[p]);
        a.flags[p] = static_cast<Flags*>(p == ex->rank ? ex->flags : ex->peerFlags[p]);
    }
    if (bands) a.snap[ex->rank] = reinterpret_cast<const float4*>(src);
    a.result = ex->rank == 0 ? ex->result : static_cast<float4*>(ex->peerResult);
    a.count = ex->count; a.epoch = ++ex->epoch; a.rank = ex->rank; a.world = ex->world; a.width = width; a.bandHeight = bandHeight;
    const uint64_t items = (ex->count + ex->world - 1) / ex->world;      // both modes: a rank moves about 1 / world of the frame
    // A small grid that is resident as a whole (2 blocks of 256 per SM; measured best of 1 / 2 / 4 / 16 / 64: the per-block cost is the
    // system-scope fence that waits for the block's stores to reach the peer, 16 blocks per SM were 2.2x slower): block 0 signals the arrival.
    static const int gridMult = std::getenv("HRC_EXCHANGE_GRID_MULT") ? std::max(1, std::min(4, std::atoi(std::getenv("HRC_EXCHANGE_GRID_MULT")))) : 2;
    const int grid = (int)std::min<uint64_t>((items + 255) / 256, (uint64_t)ex->smCount * gridMult);
    if (bands) k_peer_exchange<true><<<grid, 256, 0, s>>>(a);
    else       k_peer_exchange<false><<<grid, 256, 0, s>>>(a);
    EXCU(ex, cudaGetLastError());
    return HRC_OK;
}

HRC_API int hrcExchangeSum(HrcExchange ex, const float* srcDevice, void* cudaStream) { return runExchange(ex, srcDevice, false, 0, 0, cudaStream); }
HRC_API int hrcExchangeBands(HrcExchange ex, const float* srcDevice, int32_t width, int32_t bandHeight, void* cudaStream) { return runExchange(ex, srcDevice, true, width, bandHeight, cudaStream); }

HRC_API int hrcExchangeResult(HrcExchange ex, const float** devicePtr) {
    if (!ex || !devicePtr) return exFail(ex, HRC_ERR_INVALID_VALUE, "null pointer");
    if (ex->rank != 0) return exFail(ex, HRC_ERR_INVALID_OPERATION, "the combined frame lives on rank 0");
    *devicePtr = reinterpret_cast<const float*>(ex->result);
    return HRC_OK;
}

HRC_API int hrcExchangeStatus(HrcExchange ex) {
    if (!ex) return exFail(nullptr, HRC_ERR_INVALID_VALUE, "null exchange");
    EXCU(ex, cudaSetDevice(ex->device));
    uint32_t err = 0;
    EXCU(ex, cudaDeviceSynchronize());
    EXCU(ex, cudaMemcpy(&err, &ex->flags->error, sizeof(err), cudaMemcpyDeviceToHost));
    if (err) return exFail(ex, HRC_ERR_INVALID_OPERATION, err == 1 ? "a peer never arrived at the exchange (every rank must call it the same number of times)" : "a peer never finished the exchange");
    return HRC_OK;
}

HRC_API int hrcExchangeDestroy(HrcExchange ex) {
    if (!ex) return HRC_OK;
    cudaSetDevice(ex->device);
    cudaDeviceSynchronize();
    for (int p = 0; p < kMaxRanks; ++p) { if (ex->peerSnap[p]) cudaIpcCloseMemHandle(ex->peerSnap[p]); if (ex->peerFlags[p]) cudaIpcCloseMemHandle(ex->peerFlags[p]); }
    if (ex->peerResult) cudaIpcCloseMemHandle(ex->peerResult);
    cudaFree(ex->snap); cudaFree(ex->flags); cudaFree(ex->result);
    delete ex;
    return HRC_OK;
}
