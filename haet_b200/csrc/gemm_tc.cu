// tcgen05 GEMM with fp32-class accuracy: Y[M, Nn] = X[M, K] . W[Nn, K]^T + bias, X / Y fp32 in HBM.
// Used for in_project_x (IrregularMesh.py:157) and to_out (:235).
//
// Accuracy: the Rep-Slice softmax divides the logits by tau >= 0.1 and exponentiates, so rounding the
// projection's operands to bf16 (or tf32) costs 3e-2 (7e-3) relative error on the layer output
// (scripts/precision_study.py).  Both operands are therefore split x = hi + lo into two bf16 and the
// product is accumulated as hi.hi + lo.hi + hi.lo in the fp32 TMEM accumulator (error ~2^-17).
//
// Mapping (persistent, one CTA per SM): CTA = (worker, column half).  The CTA's half of W (hi and lo,
// <= 128 KB) is bulk-copied ONCE into shared memory as ready-made SWIZZLE_128B K-major panels and stays
// resident; the CTA then walks 128-row tiles of X.  Warp roles: 4 producer warps (coalesced LDG of fp32
// X, hi/lo split, swizzled STS), 1 MMA warp (one thread issues tcgen05.mma, accumulators double-buffered
// in TMEM), 4 epilogue warps (tcgen05.ld, + bias, store).  mbarrier pipelines: a_full/a_empty per stage,
// t_full/t_empty per accumulator buffer.
#include "common.cuh"
#include "tc.cuh"

namespace haet {
using namespace tc;

constexpr int GT_M = 128;          // rows per tile (UMMA M)
constexpr int GT_N = 128;          // output columns per CTA (UMMA N)
constexpr int GT_KP = 64;          // k per panel: 64 bf16 = one 128-byte swizzle row
constexpr int GT_STAGES = 2;
constexpr int GT_PANEL_BYTES = GT_M * 128;   // 16 KB: 128 rows x 128 B
constexpr int GT_THREADS = 288;

// W [Nn, K] fp32 -> image [Nn/128 halves][K/64 panels][hi, lo][128 rows x 128 B swizzled]
__global__ void gemm_tc_pack_kernel(const float *__restrict__ w, int Nn, int K, uint8_t *__restrict__ img) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 8 consecutive k (one 16-B chunk)
    const int chunks_per_row = K / 8;
    if (idx >= Nn * chunks_per_row) return;
    const int n = idx / chunks_per_row, ck = idx - n * chunks_per_row;
    const int half = n / GT_N, row = n % GT_N;
    const int k0 = ck * 8, panel = k0 / GT_KP, chunk = (k0 % GT_KP) / 8;
    const int panels = K / GT_KP;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) split_bf16(w[(size_t)n * K + k0 + i], hi[i], lo[i]);
    uint8_t *base = img + ((size_t)half * panels + panel) * (2 * GT_PANEL_BYTES) + sw128_off(row, chunk);
    *reinterpret_cast<uint4 *>(base) = *reinterpret_cast<const uint4 *>(hi);
    *reinterpret_cast<uint4 *>(base + GT_PANEL_BYTES) = *reinterpret_cast<const uint4 *>(lo);
}

struct GemmSmem {   // control block placed after the tiles
    uint64_t w_full;
    uint64_t a_full[GT_STAGES], a_empty[GT_STAGES];
    uint64_t t_full[2], t_empty[2];
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(GT_THREADS, 1)
gemm_tc_kernel(const float *__restrict__ X, const uint8_t *__restrict__ Wimg, const float *__restrict__ bias,
               float *__restrict__ Y, int64_t M, int K, int Nn, int n_workers) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int panels = K / GT_KP;
    const int n_halves = Nn / GT_N;
    uint8_t *sW = smem;                                         // [panels][hi|lo][16 KB]
    uint8_t *sA = sW + (size_t)panels * 2 * GT_PANEL_BYTES;     // [stages][hi|lo][16 KB]
    GemmSmem *cb = reinterpret_cast<GemmSmem *>(sA + (size_t)GT_STAGES * 2 * GT_PANEL_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int half = blockIdx.x % n_halves, worker = blockIdx.x / n_halves;
    const int64_t n_tiles = (M + GT_M - 1) / GT_M;

    if (warp == 4) {
        if (lane == 0) {
            mbar_init(&cb->w_full, 1);
            for (int s = 0; s < GT_STAGES; ++s) { mbar_init(&cb->a_full[s], 128); mbar_init(&cb->a_empty[s], 1); }
            for (int b = 0; b < 2; ++b) { mbar_init(&cb->t_full[b], 1); mbar_init(&cb->t_empty[b], 128); }
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc<256>(&cb->tmem_base);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = cb->tmem_base;

    if (warp < 4) {
        // ---------------- producers: X tile panel -> bf16 hi/lo, SWIZZLE_128B K-major ----------------
        const int tid = threadIdx.x;
        int stage = 0, phase = 0;
        for (int64_t tile = worker; tile < n_tiles; tile += n_workers) {
            for (int p = 0; p < panels; ++p) {
                mbar_wait(&cb->a_empty[stage], phase ^ 1);
                uint8_t *hi_base = sA + (size_t)stage * 2 * GT_PANEL_BYTES;
                float4 v[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int idx = i * 128 + tid, row = idx >> 4, c4 = idx & 15;
                    const int64_t m = tile * GT_M + row;
                    v[i] = (m < M) ? __ldg(reinterpret_cast<const float4 *>(X + m * K + p * GT_KP + c4 * 4))
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int idx = i * 128 + tid, row = idx >> 4, c4 = idx & 15;
                    __nv_bfloat16 h0, h1, h2, h3, l0, l1, l2, l3;
                    split_bf16(v[i].x, h0, l0); split_bf16(v[i].y, h1, l1);
                    split_bf16(v[i].z, h2, l2); split_bf16(v[i].w, h3, l3);
                    __nv_bfloat162 ha = __halves2bfloat162(h0, h1), hb = __halves2bfloat162(h2, h3);
                    __nv_bfloat162 la = __halves2bfloat162(l0, l1), lb = __halves2bfloat162(l2, l3);
                    const uint32_t off = sw128_off(row, c4 >> 1) + (c4 & 1) * 8;
                    uint2 hv = make_uint2(*reinterpret_cast<uint32_t *>(&ha), *reinterpret_cast<uint32_t *>(&hb));
                    uint2 lv = make_uint2(*reinterpret_cast<uint32_t *>(&la), *reinterpret_cast<uint32_t *>(&lb));
                    *reinterpret_cast<uint2 *>(hi_base + off) = hv;
                    *reinterpret_cast<uint2 *>(hi_base + GT_PANEL_BYTES + off) = lv;
                }
                fence_proxy_async();
                mbar_arrive(&cb->a_full[stage]);
                if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 4) {
        // ---------------- MMA issuer ----------------
        if (lane == 0) {
            const uint32_t wbytes = (uint32_t)panels * 2 * GT_PANEL_BYTES;
            mbar_arrive_expect_tx(&cb->w_full, wbytes);
            const uint8_t *src = Wimg + (size_t)half * wbytes;
            for (int c = 0; c < panels * 2; ++c)
                bulk_g2s(sW + (size_t)c * GT_PANEL_BYTES, src + (size_t)c * GT_PANEL_BYTES, GT_PANEL_BYTES, &cb->w_full);
            mbar_wait(&cb->w_full, 0);
            const uint32_t idesc = instr_desc(FMT_BF16, GT_M, GT_N, false, false);
            int stage = 0, phase = 0;
            int64_t it = 0;
            for (int64_t tile = worker; tile < n_tiles; tile += n_workers, ++it) {
                const int buf = (int)(it & 1);
                mbar_wait(&cb->t_empty[buf], (uint32_t)(((it >> 1) & 1) ^ 1));
                tc_fence_after();
                const uint32_t d = tmem_base + (uint32_t)buf * GT_N;
                for (int p = 0; p < panels; ++p) {
                    mbar_wait(&cb->a_full[stage], phase);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(sA + (size_t)stage * 2 * GT_PANEL_BYTES), a_lo = a_hi + GT_PANEL_BYTES;
                    const uint32_t b_hi = smem_u32(sW + (size_t)p * 2 * GT_PANEL_BYTES), b_lo = b_hi + GT_PANEL_BYTES;
#pragma unroll
                    for (int ks = 0; ks < GT_KP / 16; ++ks) {
                        const uint64_t dah = smem_desc_sw128(a_hi + ks * 32, 16, 1024);
                        const uint64_t dal = smem_desc_sw128(a_lo + ks * 32, 16, 1024);
                        const uint64_t dbh = smem_desc_sw128(b_hi + ks * 32, 16, 1024);
                        const uint64_t dbl = smem_desc_sw128(b_lo + ks * 32, 16, 1024);
                        mma_bf16_ss(d, dah, dbh, idesc, (p | ks) != 0);
                        mma_bf16_ss(d, dal, dbh, idesc, true);
                        mma_bf16_ss(d, dah, dbl, idesc, true);
                    }
                    tc_commit(&cb->a_empty[stage]);
                    if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
                }
                tc_commit(&cb->t_full[buf]);
            }
        }
        __syncwarp();
    } else {
        // ---------------- epilogue: TMEM -> registers -> + bias -> global ----------------
        const int q = warp & 3;   // TMEM lane quarter this warp may access
        int64_t it = 0;
        for (int64_t tile = worker; tile < n_tiles; tile += n_workers, ++it) {
            const int buf = (int)(it & 1);
            mbar_wait(&cb->t_full[buf], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            const int64_t m = tile * GT_M + q * 32 + lane;
            float *yrow = Y + m * Nn + half * GT_N;
#pragma unroll 1
            for (int c0 = 0; c0 < GT_N; c0 += 32) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * GT_N + c0), r);
                tmem_wait_ld();
                if (m < M) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        float4 o;
                        o.x = __uint_as_float(r[j]) + __ldg(bias + half * GT_N + c0 + j);
                        o.y = __uint_as_float(r[j + 1]) + __ldg(bias + half * GT_N + c0 + j + 1);
                        o.z = __uint_as_float(r[j + 2]) + __ldg(bias + half * GT_N + c0 + j + 2);
                        o.w = __uint_as_float(r[j + 3]) + __ldg(bias + half * GT_N + c0 + j + 3);
                        *reinterpret_cast<float4 *>(yrow + c0 + j) = o;
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(&cb->t_empty[buf]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

bool gemm_tc_supported(int K, int Nn) { return K % GT_KP == 0 && K >= GT_KP && K <= 256 && Nn % GT_N == 0 && Nn >= GT_N; }

int64_t gemm_tc_image_bytes(int K, int Nn) { return (int64_t)Nn * K * 2 * 2; }

static int sm_count() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

// img: scratch of gemm_tc_image_bytes(K, Nn) bytes (the packed weights; rebuilt each call, ~2 us)
int launch_gemm_tc(const float *x, const float *w, const float *bias, float *y, int64_t M, int K, int Nn, void *img,
                   cudaStream_t st) {
    if (M == 0) return HAET_OK;
    HAET_CHECK_ARG(gemm_tc_supported(K, Nn), "gemm_tc: unsupported K=%d Nn=%d", K, Nn);
    const int chunks = Nn * (K / 8);
    gemm_tc_pack_kernel<<<(chunks + 255) / 256, 256, 0, st>>>(w, Nn, K, (uint8_t *)img);
    HAET_LAUNCHED();
    const int panels = K / GT_KP, n_halves = Nn / GT_N;
    const size_t smem = 1024 + (size_t)(panels + GT_STAGES) * 2 * GT_PANEL_BYTES + sizeof(GemmSmem);
    static size_t smem_set = 0;
    if (smem > smem_set) {
        HAET_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    const int64_t n_tiles = (M + GT_M - 1) / GT_M;
    int n_workers = sm_count() / n_halves;
    if (n_workers < 1) n_workers = 1;
    if (n_workers > n_tiles) n_workers = (int)n_tiles;
    gemm_tc_kernel<<<n_workers * n_halves, GT_THREADS, smem, st>>>(x, (const uint8_t *)img, bias, y, M, K, Nn, n_workers);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet
