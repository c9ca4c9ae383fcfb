// The one cross-GPU exchange of the point-sharded layer: every rank needs every rank's packed slice state
// (B*H*G*(C+2) fp32, 278 KB at config 4) before the rank-ordered merge.  Instead of an NCCL all-gather this kernel works
// on PEER MEMORY over NVLink: each rank keeps a symmetric buffer (two slots + one flag, mapped by every peer),
//   1. copies its own partial into slot (epoch & 1) of its buffer, and -- once all its blocks have done so -- publishes
//      `epoch` in its flag with a system-scope release store,
//   2. waits (acquire loads over NVLink) until every peer's flag has reached `epoch`, and
//   3. pulls ALL peers' slots (one flat index space, several P2P loads in flight per thread) into the rank-major array the merge reads.
// The epoch lives in device memory and is advanced by the kernel itself, so the launch takes no host-side counter and
// can be replayed from a CUDA graph.  A rank is never more than one epoch ahead of a peer (it needs the peer's flag of
// its current epoch), which is why two slots suffice.  Waits are bounded (HAET_PEER_TIMEOUT_S, default 60 s of
// %globaltimer): a peer that never arrives makes the kernel raise the error word and trap -- the next CUDA call of the
// process fails loudly, as an NCCL watchdog abort would -- instead of hanging the GPU or merging a stale state.
#include "common.cuh"

namespace haet {

constexpr int PX_MAX_RANKS = 16;
constexpr int PX_THREADS = 256;
constexpr int PX_UNROLL = 4;      // P2P loads in flight per thread

struct PeerPtrs {
    float *buf[PX_MAX_RANKS];   // peer r's symmetric buffer: [2][n] floats, then the flag word (as mapped in THIS process)
};

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// state (local, 4 words): [0] epoch of the last completed exchange, [1] blocks that finished step 1, [2] blocks that
// finished the kernel, [3] error (1 = a peer did not arrive)
__global__ void __launch_bounds__(PX_THREADS) peer_exchange_kernel(PeerPtrs peers, int rank, int world, int64_t n,
                                                                    const float *__restrict__ mine,
                                                                    float *__restrict__ gathered, uint32_t *state,
                                                                    unsigned long long timeout_ns) {
    __shared__ uint32_t s_epoch;
    const int64_t n4 = n >> 2;   // n % 4 == 0 (checked by the host)
    if (threadIdx.x == 0) s_epoch = *reinterpret_cast<volatile uint32_t *>(state) + 1;
    __syncthreads();
    const uint32_t epoch = s_epoch;
    const int64_t slot = (int64_t)(epoch & 1) * n;
    uint32_t *my_flag = reinterpret_cast<uint32_t *>(peers.buf[rank] + 2 * n);

    // ---- 1. my partial -> my symmetric slot (and my row of the gathered array), then publish ----
    {
        const float4 *src = reinterpret_cast<const float4 *>(mine);
        float4 *d0 = reinterpret_cast<float4 *>(peers.buf[rank] + slot);
        float4 *d1 = reinterpret_cast<float4 *>(gathered + (int64_t)rank * n);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
            const float4 v = src[i];
            d0[i] = v;
            d1[i] = v;
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t t = atomicAdd(state + 1, 1u);
        if (t == gridDim.x - 1) {   // every block of this rank has written its share
            state[1] = 0;
            __threadfence_system();
            st_release_sys(my_flag, epoch);
        }
    }

    // ---- 2. wait for ALL peers (thread 0 of every block polls their flags in turn), then 3. pull their slots ----
    // Round 2: the first version pulled peer after peer, each thread copying ~4 float4 per peer through a loop whose loads
    // could not overtake its stores -- 4 dependent NVLink round trips x (world - 1) peers = +0.15 ms per step at 8 GPUs.  Now the
    // (world - 1) * n / 4 loads are one flat index space, PX_UNROLL loads in flight per thread before the first store.
    if (threadIdx.x == 0) {
        uint32_t ok = 1;
        int bad = -1;
        unsigned long long t0 = 0;
        for (int step = 1; step < world && ok; ++step) {
            const int r = (rank + step) % world;
            const uint32_t *flag = reinterpret_cast<const uint32_t *>(peers.buf[r] + 2 * n);
            for (uint32_t spin = 0;; ++spin) {
                // peers are at most one epoch ahead; the wrap-safe comparison keeps working after 2^32 exchanges
                if ((int32_t)(ld_acquire_sys(flag) - epoch) >= 0) break;
                __nanosleep(32);
                if ((spin & 1023u) == 1023u) {   // look at the wall clock every ~1000 polls
                    unsigned long long now;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                    if (t0 == 0) t0 = now;
                    else if (now - t0 > timeout_ns) { ok = 0; bad = r; break; }
                }
            }
        }
        if (!ok) {
            atomicExch(state + 3, 1u);
            printf("haet_b200: peer exchange: rank %d waited more than %llu s for rank %d (epoch %u)\n", rank,
                   timeout_ns / 1000000000ull, bad, epoch);
            __threadfence_system();
            __trap();
        }
    }
    __syncthreads();
    {
        const int64_t total = (int64_t)(world - 1) * n4, nthr = (int64_t)gridDim.x * blockDim.x;
        for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base < total; base += nthr * PX_UNROLL) {
            float4 v[PX_UNROLL];
            float4 *dst[PX_UNROLL];
#pragma unroll
            for (int u = 0; u < PX_UNROLL; ++u) {
                const int64_t idx = base + (int64_t)u * nthr;
                dst[u] = nullptr;
                if (idx < total) {
                    const int step = (int)(idx / n4) + 1;                 // start with the next rank: spreads the NVLink reads
                    const int64_t i = idx - (int64_t)(step - 1) * n4;
                    const int r = (rank + step) % world;
                    v[u] = reinterpret_cast<const float4 *>(peers.buf[r] + slot)[i];   // P2P load
                    dst[u] = reinterpret_cast<float4 *>(gathered + (int64_t)r * n) + i;
                }
            }
#pragma unroll
            for (int u = 0; u < PX_UNROLL; ++u)
                if (dst[u]) *dst[u] = v[u];
        }
    }

    // ---- the last block out advances the epoch (every block has read it by now) ----
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const uint32_t t = atomicAdd(state + 2, 1u);
        if (t == gridDim.x - 1) {
            state[2] = 0;
            __threadfence();
            *reinterpret_cast<volatile uint32_t *>(state) = epoch;
        }
    }
}

}  // namespace haet

using namespace haet;

extern "C" {

int64_t haet_peer_buffer_bytes(const haet_layer_shape *s) {
    return (2 * haet_partial_floats(s) + 4) * (int64_t)sizeof(float);
}

int haet_peer_exchange(const haet_layer_shape *s, void *const *peer_bufs, int32_t rank, int32_t world,
                       const float *partial_d, float *gathered_d, void *state_d, void *stream) {
    HAET_CHECK_ARG(world >= 1 && world <= PX_MAX_RANKS, "haet_peer_exchange: world size %d not in [1, %d]", world, PX_MAX_RANKS);
    HAET_CHECK_ARG(rank >= 0 && rank < world, "haet_peer_exchange: rank %d of %d", rank, world);
    const int64_t n = haet_partial_floats(s);
    HAET_CHECK_ARG(n % 4 == 0, "haet_peer_exchange: the packed slice state must be a multiple of 4 floats");
    if (n == 0) return HAET_OK;
    PeerPtrs pp;
    for (int r = 0; r < PX_MAX_RANKS; ++r) pp.buf[r] = r < world ? (float *)peer_bufs[r] : nullptr;
    // enough threads that the pulls of all peers are a handful of loads per thread; all blocks are co-resident (the flag of
    // step 1 is published by the last block to arrive): at most one block per SM here
    int dev = 0, sms = 148;
    HAET_CUDA(cudaGetDevice(&dev));
    HAET_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int64_t work = (n / 4) * (world > 1 ? world - 1 : 1);
    int blocks = (int)((work + (int64_t)PX_THREADS * PX_UNROLL - 1) / ((int64_t)PX_THREADS * PX_UNROLL));
    if (blocks > sms) blocks = sms;
    if (blocks < 1) blocks = 1;
    const unsigned long long timeout_ns = (unsigned long long)(options().peer_timeout_s * 1e9);
    peer_exchange_kernel<<<blocks, PX_THREADS, 0, (cudaStream_t)stream>>>(pp, rank, world, n, partial_d, gathered_d,
                                                                          (uint32_t *)state_d, timeout_ns);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // extern "C"
