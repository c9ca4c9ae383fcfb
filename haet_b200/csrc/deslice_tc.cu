// Pass 2 on tcgen05: o[b, n, h*C + c] = sum_g W[n,g] S'[b,h,g,c],  W[n,g] = 2^(z2[n,g] - lse2[g])
// (IrregularMesh.py:221-232; the N x G weight matrix is recomputed tile by tile, never materialised).
//
// Points on the TMEM lanes (thread = point, so s_n = log2e / tau_n is a register):
//   MMA1 (kind::f16, 3-term fp16 split): D[n, g] = x_proj[n,:] . Ws[g,:]      M = 128 points, N = 128 slices (64 when G <= 64), K' = 96 (+ 16: bias)
//        (N = 128: with N = 64 / 32 every tcgen05.mma costs ~50 cycles of tensor-pipe time whatever its size -- the pipe was
//        the bound of the kernel at one item per ~750 cycles -- so each instruction is made to carry twice the logits)
//        (A = a 128-point tile of the spill's operand image, common.cuh, brought in by one bulk copy)
//   softmax warps: P = 2^(D * s_n - lse2_g + 14) -> TMEM as packed fp16 (the bias bc_g rides in MMA1 as an extra K step; W <= 1, so
//        2^14 W keeps 11 bits down to W = 2^-28 -- same mantissa as tf32 -- and the epilogue multiplies by 2^-14)
//   MMA2 (kind::f16, A = P from TMEM, B = S'_h as MN-major fp16 rows): O[n, c] += sum_g P[n, g] S'[g, c]
//        (S'_h is scaled by a per-(b, h) power of two so that its largest magnitude sits in [2^13, 2^14): the single fp16 term
//        then keeps 2^-12 relative precision whatever the range of the Erwin output -- no overflow above 65504, no flush
//        of small rows -- and the epilogue multiplies the scale back out)
// CTA = (b, h, chunk of points): the Ws and S'_h operand images are built once and stay resident; the x_proj
// head slice streams through a 4-stage ring of bulk copies; O is double-buffered in TMEM and stored by 4 epilogue warps.
#include <cuda_fp16.h>

#include <cuda_bf16.h>

#include "common.cuh"
#include "tc.cuh"

namespace haet {
using namespace tc;

#ifdef HAET_TRACE   // development build only (scripts/trace_pass2.py): cycle counters per warp role, never in the shipped library
__device__ unsigned long long g_dt_trace[64];
#define DT_T(var) const long long var = clock64()
#define DT_DECL unsigned long long dt_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define DT_ACC(slot, v) dt_acc[(slot) & 7] += (unsigned long long)(v)
#define DT_FLUSH(base) do { if (lane == 0) for (int i_ = 0; i_ < 8; ++i_) atomicAdd(&g_dt_trace[(base) + i_], dt_acc[i_]); } while (0)
#else
#define DT_T(var)
#define DT_DECL
#define DT_ACC(slot, v)
#define DT_FLUSH(base)
#endif

constexpr int DT_C = 32;
constexpr int DT_TILE = 128;        // points per tile (UMMA M)
constexpr int DT_GCMAX = 128;       // slices per item (UMMA N of MMA1, K of MMA2): 128, or 64 when G <= 64
constexpr int DT_STAGES = 4;
constexpr int DT_THREADS = 640;     // 12 softmax warps (3 warpgroups) + 4 epilogue warps + MMA1 warp + loader warp + 2 MMA2 warps
constexpr int DT_MAX_G = 512;
constexpr float DT_PSHIFT = 14.f;   // P is handed to MMA2 as fp16 2^14 W (W <= 1): full precision down to W = 2^-28, flush below 2^-38
constexpr int DT_NBUF = 3;          // D / P buffers (128 TMEM columns each)
constexpr int DT_TILE_BYTES = DT_TILE * 128;

struct DesliceCtl {
    uint64_t a_full[DT_STAGES], a_empty[DT_STAGES];
    uint64_t d_full[DT_NBUF], p_full[DT_NBUF], p_free[DT_NBUF];
    uint64_t o_full[2], o_empty[2];
    uint32_t tmem_base;
    float red[20];       // block reduction of max |S'_h| (one slot per warp)
    float o_scale;       // 2^-DT_PSHIFT / (the power-of-two scale applied to S'_h)
};

// TMEM columns: O accumulators [issuer][buffer] x 32 at 0..127, D / P buffers [stream][3] x 64 at 128..511
// TMEM columns: O accumulators [MMA2 warp][double buffer] x 32 at 0..127, D / P buffers at 128 + 128 buf
__device__ __forceinline__ uint32_t dt_dbuf_col(int buf) { return 128u + (uint32_t)buf * DT_GCMAX; }
__host__ __device__ __forceinline__ int dt_gpad(int G) { return G <= 64 ? 64 : (G + 127) / 128 * 128; }

// WG: softmax warpgroups (blockDim.x = (4 WG + 8) * 32).  3: warpgroup i owns D / P buffer i.  2: the warpgroups alternate over the
// three buffers (item q: buffer q % 3, warpgroup q & 1), so a warpgroup never waits for the MMA phases of the buffer it just filled.
// (Measured and dropped, DESIGN 10: WG = 2 removes the wait for d_full -- 1060 -> 126 cycles per item -- but two warps per scheduler
// feed the MUFU pipe no better, 0.667 against 0.651 ms; software-pipelined TMEM reads in 16-column units were slower with either.)
// POLY: eighths (0..4) of every thread's logits whose 2^t comes from poly_ex2_pair (FMA / ALU pipes) instead of MUFU.EX2, interleaved
// pair by pair in the same instruction stream (0 = all MUFU).
template <bool F2, int WG, int POLY>
__global__ void __launch_bounds__(DT_THREADS, 1)
deslice_tc_kernel(const uint8_t *__restrict__ img, const float *__restrict__ scale, const float *__restrict__ lse2,
                  const float *__restrict__ states, const float *__restrict__ ws, const float *__restrict__ bs, float shift,
                  int64_t N, int64_t npad, int H, int G, int C, int64_t chunk_pts, float *__restrict__ o, int o_image) {
    // C = dim_head: 32, or 16 -- then the image rows, Ws and S' are zero-padded to 32 channels and only 16 are written out --
    // or 64 ("pair mode", slice_tc.cu): H counts pseudo-heads of 32 channels, the logits run over both sub-row tiles of the
    // head, and this CTA produces the 32 output channels of its own sub-row from S'[.., 32 srow : 32 srow + 32]
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS / STS)
    const int g_pad = dt_gpad(G), cw = g_pad < DT_GCMAX ? g_pad : DT_GCMAX;   // chunk width: slices per item
    const int n_gc = g_pad / cw;
    const bool pair = C == 64;
    const int Cw = pair ? 32 : C;                           // channels per image row / output channels of this CTA
    const int n_stages = pair ? DT_STAGES - 1 : DT_STAGES;  // pair mode: a stage holds the tiles of both sub-rows
    const uint32_t stage_bytes = pair ? 2 * DT_TILE_BYTES : DT_TILE_BYTES;
    uint8_t *sW = smem;                                     // [sub-row][g_pad rows][hi(32 fp16) | lo(32 fp16)]  K-major SW128
    uint8_t *sV = sW + (size_t)g_pad * 128 * (pair ? 2 : 1);   // [g_pad rows][32 fp32]  S'_h, MN-major SW128_BASE32B
    uint8_t *sA = sV + (size_t)g_pad * 128;                 // [stages][128 rows][hi | lo] (x 2 sub-rows)
    // bias operands of MMA1's extra K step (no swizzle: 8-row x 16-B core matrices, K halves 128 B apart, 8-row groups 256 B
    // apart): sBb row g = {bc_g hi, bc_g lo, 0 ..} with bc = bs_g - log(-log eps), sOne row n = {1, 1, 0 ..}
    uint8_t *sBb = sA + (size_t)n_stages * stage_bytes;     // [g_pad / 8][256 B]
    uint8_t *sOne = sBb + (size_t)g_pad * 32;               // [128 / 8][256 B]
    float *sL = reinterpret_cast<float *>(sOne + DT_TILE * 32);                        // [g_pad] -lse2_g
    float *sS = sL + g_pad;                                                            // [stages][128] s_n
    DesliceCtl *ctl = reinterpret_cast<DesliceCtl *>(sS + DT_STAGES * DT_TILE);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform role index
    constexpr int W_EPI = 4 * WG, W_MMA1 = W_EPI + 4, W_LOAD = W_EPI + 5, W_MMA2 = W_EPI + 6, NTHR = (4 * WG + 8) * 32;
    const int bh = blockIdx.y, b = bh / H, h = bh - b * H;   // pair mode: h = pseudo-head 2 h_real + srow
    const int bhr = pair ? (bh >> 1) : bh, srow = pair ? (bh & 1) : 0;
    const int inner = H * Cw;
    const int64_t n_begin = (int64_t)blockIdx.x * chunk_pts;
    const int64_t n_end = min(N, n_begin + chunk_pts);
    const int n_tiles = (int)((n_end - n_begin + DT_TILE - 1) / DT_TILE);
    DT_T(k_start);

    if (warp == W_MMA1) {
        if (lane == 0) {
            for (int s = 0; s < DT_STAGES; ++s) { mbar_init(&ctl->a_full[s], 1); mbar_init(&ctl->a_empty[s], 2); }
            for (int k = 0; k < DT_NBUF; ++k) { mbar_init(&ctl->d_full[k], 1); mbar_init(&ctl->p_full[k], 128); mbar_init(&ctl->p_free[k], 2); }
            for (int k = 0; k < 2; ++k) { mbar_init(&ctl->o_full[k], 2); mbar_init(&ctl->o_empty[k], 128); }
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc<512>(&ctl->tmem_base);
    }
    // power-of-two scale of S'_h: 2^(14 - e) with max |S'_h| in [2^(e-1), 2^e)
    {
        float amax = 0.f;
        const float *sp = states + (size_t)bhr * G * C + srow * 32;
        for (int idx = threadIdx.x; idx < G * Cw; idx += NTHR) {
            const int g = idx / Cw, c = idx - g * Cw;
            amax = fmaxf(amax, fabsf(sp[(size_t)g * C + c]));
        }
#pragma unroll
        for (int o2 = 16; o2; o2 >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o2));
        if (lane == 0) ctl->red[threadIdx.x >> 5] = amax;
    }
    __syncthreads();
    float v_scale = 1.f;
    {
        float amax = 0.f;
#pragma unroll
        for (int w2 = 0; w2 < NTHR / 32; ++w2) amax = fmaxf(amax, ctl->red[w2]);
        int e = 14;
        if (amax > 0.f && amax <= 3.0e38f) frexpf(amax, &e);   // amax = f 2^e, f in [0.5, 1); zero / inf / NaN states: no scaling
        e = e < -90 ? -90 : e;
        v_scale = ldexpf(1.f, 14 - e);
        if (threadIdx.x == 0) ctl->o_scale = ldexpf(1.f, e - 14 - (int)DT_PSHIFT);
    }
    // resident operand images: Ws rows [hi | lo] and S'_h rows (fp16), per-column constants
    for (int idx = threadIdx.x; idx < g_pad * 4; idx += NTHR) {
        const int g = idx >> 2, ck = idx & 3;
        __align__(16) __half hi[8], lo[8];
        __align__(16) __half sv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const bool live = g < G && ck * 8 + i < Cw;
            const float v = live ? ws[(size_t)g * C + ck * 8 + i] : 0.f;
            hi[i] = __float2half_rn(v);
            lo[i] = __float2half_rn(v - __half2float(hi[i]));
            const float s = live ? states[((size_t)bhr * G + g) * C + srow * 32 + ck * 8 + i] : 0.f;
            sv[i] = __float2half_rn(s * v_scale);
        }
        *reinterpret_cast<uint4 *>(sW + sw128_off(g, ck)) = *reinterpret_cast<const uint4 *>(hi);
        *reinterpret_cast<uint4 *>(sW + sw128_off(g, ck + 4)) = *reinterpret_cast<const uint4 *>(lo);
        if (pair) {   // Ws channels 32 .. 63: the second sub-row's image
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float v = g < G ? ws[(size_t)g * C + 32 + ck * 8 + i] : 0.f;
                hi[i] = __float2half_rn(v);
                lo[i] = __float2half_rn(v - __half2float(hi[i]));
            }
            uint8_t *sW1 = sW + (size_t)g_pad * 128;
            *reinterpret_cast<uint4 *>(sW1 + sw128_off(g, ck)) = *reinterpret_cast<const uint4 *>(hi);
            *reinterpret_cast<uint4 *>(sW1 + sw128_off(g, ck + 4)) = *reinterpret_cast<const uint4 *>(lo);
        }
        *reinterpret_cast<uint4 *>(sV + sw128_off(g, ck)) = *reinterpret_cast<const uint4 *>(sv);   // MN-major operand: row g = 32 fp16 channels
        *reinterpret_cast<uint4 *>(sV + sw128_off(g, ck + 4)) = make_uint4(0, 0, 0, 0);
    }
    for (int g = threadIdx.x; g < g_pad; g += NTHR) {
        sL[g] = g < G ? DT_PSHIFT - lse2[(size_t)bhr * G + g] : -INFINITY;
        uint32_t b2 = 0;
        if (g < G) {
            const float bc = bs[g] - shift;
            const __half bhi = __float2half_rn(bc), blo = __float2half_rn(bc - __half2float(bhi));
            b2 = (uint32_t)__half_as_ushort(bhi) | ((uint32_t)__half_as_ushort(blo) << 16);
        }
        uint8_t *dstb = sBb + (g >> 3) * 256 + (g & 7) * 16;
        *reinterpret_cast<uint4 *>(dstb) = make_uint4(b2, 0, 0, 0);
        *reinterpret_cast<uint4 *>(dstb + 128) = make_uint4(0, 0, 0, 0);
    }
    for (int n = threadIdx.x; n < DT_TILE; n += NTHR) {
        uint8_t *dst1 = sOne + (n >> 3) * 256 + (n & 7) * 16;
        *reinterpret_cast<uint4 *>(dst1) = make_uint4(0x3c003c00u, 0, 0, 0);   // fp16 1.0, 1.0
        *reinterpret_cast<uint4 *>(dst1 + 128) = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp >= W_EPI && warp < W_EPI + 4) {
        // =========================== O epilogue: TMEM -> global ===========================
        const int r = threadIdx.x - W_EPI * 32;   // row of the tile == TMEM lane
        const float o_scale = ctl->o_scale;
        const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
        float *dst = o + ((int64_t)b * N) * inner + h * Cw;
        for (int te = 0; te < n_tiles; ++te) {
            const int ob = te & 1;
            mbar_wait(&ctl->o_full[ob], (uint32_t)((te >> 1) & 1));
            tc_fence_after();
            uint32_t v[32];
            tmem_ld32(tmem_base + lane_base + (uint32_t)ob * DT_C, v);
            {   // the second MMA2 warp's half of every chunk's K steps
                uint32_t v1[32];
                tmem_ld32(tmem_base + lane_base + (uint32_t)(2 + ob) * DT_C, v1);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __float_as_uint((__uint_as_float(v[i]) + __uint_as_float(v1[i])) * o_scale);   // 2^-DT_PSHIFT / scale of S'
            }
            tc_fence_before();
            mbar_arrive(&ctl->o_empty[ob]);
            if (o_image) {
                // the to_out GEMM's operand image (common.cuh: OPI): lane l4 of a 4-lane group holds channels 8 l4 .. 8 l4 + 7
                // of 4 rows and writes them as one 16-byte bf16 hi chunk and one lo chunk of the (tile, panel) block
                xpose_stage<0, 3>(v, lane);
                xpose_stage<1, 4>(v, lane);
                const int64_t n0 = n_begin + (int64_t)te * DT_TILE + (r & ~3);
                // global 16-byte chunk (8 channels) of this lane inside the row of `inner` channels
                const int l4 = lane & 3, gch = (h * Cw >> 3) + l4, cc = gch & 7, pan = gch >> 3, panels = inner >> 6;
                uint8_t *oimg = reinterpret_cast<uint8_t *>(o);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const float f0 = __uint_as_float(v[8 * i + 2 * c]), f1 = __uint_as_float(v[8 * i + 2 * c + 1]);
                        const __nv_bfloat162 hb = __floats2bfloat162_rn(f0, f1);
                        const uint32_t ub = *reinterpret_cast<const uint32_t *>(&hb);
                        const __nv_bfloat162 lb = __floats2bfloat162_rn(f0 - __uint_as_float(ub << 16), f1 - __uint_as_float(ub & 0xffff0000u));
                        hi[c] = ub;
                        lo[c] = *reinterpret_cast<const uint32_t *>(&lb);
                    }
                    if (n0 + i < n_end && 8 * l4 < Cw) {
                        const int64_t m = (int64_t)b * N + n0 + i;
                        const int rr = (int)(m & 127);
                        uint8_t *blk = oimg + (((m >> 7) * panels + pan) * 2) * (int64_t)(128 * 128) + sw128_off(rr, cc);
                        *reinterpret_cast<uint4 *>(blk) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        *reinterpret_cast<uint4 *>(blk + 128 * 128) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    }
                }
                continue;
            }
            // in-warp transpose (tc.cuh): 4 adjacent lanes hold 4 rows and write 64 contiguous bytes of one row per
            // store (8 rows per instruction) instead of 32 lanes writing 16 bytes of 32 different rows
            xpose_stage<0, 2>(v, lane);
            xpose_stage<1, 3>(v, lane);
            const int64_t n0 = n_begin + (int64_t)te * DT_TILE + (r & ~3);
            const int l4 = lane & 3;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (n0 + i < n_end) {
                    float *p = dst + (n0 + i) * inner + 4 * l4;
                    *reinterpret_cast<float4 *>(p) = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                                 __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                    if (Cw > 16)
                        *reinterpret_cast<float4 *>(p + 16) = make_float4(__uint_as_float(v[4 * i + 16]), __uint_as_float(v[4 * i + 17]),
                                                                          __uint_as_float(v[4 * i + 18]), __uint_as_float(v[4 * i + 19]));
                }
            }
        }
    } else if (warp == W_LOAD) {
        // =========================== loader: one bulk copy per 128-point tile (+ its s_n) ===========================
        if (lane == 0) {
            const int64_t bh0 = pair ? (bh & ~1) : bh;   // pair mode: sub-row 0 of this head; sub-row 1 follows npad rows later
            const uint8_t *src = img + ((int64_t)bh0 * npad + n_begin) * XPI_ROW_BYTES;
            const float *ssrc = scale + (int64_t)bh * npad + n_begin;
            int stage = 0, phase = 0;
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait(&ctl->a_empty[stage], phase ^ 1);
                mbar_arrive_expect_tx(&ctl->a_full[stage], stage_bytes + DT_TILE * 4);
                bulk_g2s(sA + (size_t)stage * stage_bytes, src + (size_t)t * DT_TILE_BYTES, DT_TILE_BYTES, &ctl->a_full[stage]);
                if (pair)
                    bulk_g2s(sA + (size_t)stage * stage_bytes + DT_TILE_BYTES, src + ((size_t)npad + (size_t)t * DT_TILE) * XPI_ROW_BYTES,
                             DT_TILE_BYTES, &ctl->a_full[stage]);
                bulk_g2s(sS + stage * DT_TILE, ssrc + (size_t)t * DT_TILE, DT_TILE * 4, &ctl->a_full[stage]);
                if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == W_MMA1) {
        // =========================== MMA1 issuer ===========================
        // Items q = (tile t, slice chunk jc) in order; item q uses D / P buffer q % 3 = softmax warpgroup q % 3.  One warp can
        // issue a tcgen05.mma only every ~75 cycles whatever its size (clock64 trace: 7 + 8 small MMAs = 1300 cycles per item was
        // the bound of the kernel while the tensor pipe sat at 40 %), so the instruction streams are spread over THREE
        // warps: this one issues the logits MMAs, warps 18 / 19 each half of the K steps of P . S'.  The whole warp runs the
        // warp-uniform loop and one elected lane issues (slice_tc.cu).
        const int Q = n_tiles * n_gc;
        const uint32_t id1 = instr_desc(FMT_F16, DT_TILE, cw, false, false);
        const uint32_t wbase = smem_u32(sW) >> 4, abase = smem_u32(sA) >> 4;
        const uint32_t bb_a = smem_u32(sBb) >> 4, one_a = smem_u32(sOne) >> 4;
        const uint32_t w_step = (uint32_t)cw * 128 >> 4, b_step = (uint32_t)cw * 32 >> 4;   // one chunk of Ws rows, of bias rows
        int jc1 = 0, st1 = 0, ph1 = 0, buf1 = 0, fph = 0;   // chunk, stage (+phase), D buffer, phase of p_free
        DT_DECL;
        for (int q = 0; q < Q; ++q) {    // D = x_proj . Ws[chunk]^T  (hi.hi + lo.hi + hi.lo) + bc
            DT_T(m0);
            if (q >= DT_NBUF) mbar_wait(&ctl->p_free[buf1], (uint32_t)fph);   // both halves of P . S' of item q - 3 have completed
            DT_T(m1);
            if (jc1 == 0) mbar_wait(&ctl->a_full[st1], (uint32_t)ph1);
            tc_fence_after();
            DT_T(m2);
            DT_ACC(16, m1 - m0); DT_ACC(17, m2 - m1); DT_ACC(19, 1);
            if (elect_one()) {
                const uint32_t a = abase + (uint32_t)st1 * (stage_bytes >> 4);
                const uint32_t wb = wbase + (uint32_t)jc1 * w_step;
                const uint32_t d = tmem_base + dt_dbuf_col(buf1);
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {
                    mma_bf16_ss(d, desc_sw128_u(a + ks * 2, 16, 1024), desc_sw128_u(wb + ks * 2, 16, 1024), id1, ks != 0);
                    mma_bf16_ss(d, desc_sw128_u(a + 4 + ks * 2, 16, 1024), desc_sw128_u(wb + ks * 2, 16, 1024), id1, true);
                    mma_bf16_ss(d, desc_sw128_u(a + ks * 2, 16, 1024), desc_sw128_u(wb + 4 + ks * 2, 16, 1024), id1, true);
                }
                // + bc_g: ones[n, 0:16] . bias[g, 0:16]  (the softmax warps then need no per-element bias add)
                mma_bf16_ss(d, desc_u(one_a, 128, 256, 0), desc_u(bb_a + (uint32_t)jc1 * b_step, 128, 256, 0), id1, true);
                if (pair) {   // channels 32 .. 63: second tile of the stage against the second Ws image
                    const uint32_t a1 = a + (DT_TILE_BYTES >> 4), wb1 = wb + ((uint32_t)g_pad * 128 >> 4);
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        mma_bf16_ss(d, desc_sw128_u(a1 + ks * 2, 16, 1024), desc_sw128_u(wb1 + ks * 2, 16, 1024), id1, true);
                        mma_bf16_ss(d, desc_sw128_u(a1 + 4 + ks * 2, 16, 1024), desc_sw128_u(wb1 + ks * 2, 16, 1024), id1, true);
                        mma_bf16_ss(d, desc_sw128_u(a1 + ks * 2, 16, 1024), desc_sw128_u(wb1 + 4 + ks * 2, 16, 1024), id1, true);
                    }
                }
                tc_commit(&ctl->d_full[buf1]);
            }
            __syncwarp();
            DT_T(m3);
            DT_ACC(18, m3 - m2);
            if (++buf1 == DT_NBUF) { buf1 = 0; if (q >= DT_NBUF) fph ^= 1; }
            if (++jc1 == n_gc) { jc1 = 0; if (++st1 == n_stages) { st1 = 0; ph1 ^= 1; } }
        }
        DT_FLUSH(16);
    } else if (warp >= W_MMA2) {
        // =========================== MMA2 issuers: O_u += P . S'_h[chunk], K steps [u, u + 1) * cw / 16 ===========================
        const int u = warp - W_MMA2;
        const int Q = n_tiles * n_gc;
        const uint32_t id2 = instr_desc(FMT_F16, DT_TILE, DT_C, false, true);
        const uint32_t vbase = smem_u32(sV) >> 4;
        const uint32_t w_step = (uint32_t)cw * 128 >> 4;
        const int kh = cw / 32;   // K steps (of 16 slices) per warp
        int t2 = 0, jc2 = 0, st2 = 0, buf2 = 0, ph2 = 0;
        DT_DECL;
        for (int k = 0; k < Q; ++k) {
            const int ob = t2 & 1;
            DT_T(n0);
            if (jc2 == 0) mbar_wait(&ctl->o_empty[ob], (uint32_t)(((t2 >> 1) & 1) ^ 1));
            DT_T(n1);
            mbar_wait(&ctl->p_full[buf2], (uint32_t)ph2);
            tc_fence_after();
            DT_T(n2);
            DT_ACC(24, n1 - n0); DT_ACC(25, n2 - n1); DT_ACC(27, 1);
            if (elect_one()) {
                const uint32_t pa = tmem_base + dt_dbuf_col(buf2) + (uint32_t)(u * kh) * 8, od = tmem_base + (uint32_t)(u * 2 + ob) * DT_C;
                const uint32_t vb = vbase + (uint32_t)jc2 * w_step + (uint32_t)(u * kh) * 128;
#pragma unroll 4
                for (int ks = 0; ks < kh; ++ks)   // K = 16 slices = 8 packed TMEM columns of P, two 8-row swizzle atoms of S'
                    mma_f16_ts(od, pa + ks * 8, desc_sw128_u(vb + ks * 128, 1024, 1024), id2, (jc2 | ks) != 0);
                tc_commit(&ctl->p_free[buf2]);
                if (jc2 == n_gc - 1) { tc_commit(&ctl->o_full[ob]); tc_commit(&ctl->a_empty[st2]); }
            }
            __syncwarp();
            DT_T(n3);
            DT_ACC(26, n3 - n2);
            if (++buf2 == DT_NBUF) { buf2 = 0; ph2 ^= 1; }
            if (++jc2 == n_gc) { jc2 = 0; ++t2; if (++st2 == n_stages) st2 = 0; }
        }
        DT_FLUSH(24);
    } else if (warp < W_EPI) {
        // =========================== softmax warpgroups ===========================
        // 3 warpgroups, warpgroup i owns D / P buffer i and takes the items q = i, i + 3, ..: it sees EVERY phase of its two
        // barriers (a warpgroup that skipped phases could alias the mbarrier parity when it runs two uses ahead).
        const int wg = warp >> 2, q4 = warp & 3, row = q4 * 32 + lane;
        const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
        const int Q = n_tiles * n_gc, halves = cw / 32;
        DT_DECL;
        DT_T(l_start);
        // running item coordinates (tile t, slice chunk jc, stage + parity): the divisions they replace cost ~10 % of an item
        int t = wg / n_gc, jc = wg - t * n_gc, stage = t % n_stages, sph = (t / n_stages) & 1, dph = 0, buf = wg;
        for (int q = wg; q < Q; q += WG) {
            DT_T(t0);
            mbar_wait(&ctl->a_full[stage], (uint32_t)sph);
            const float sn = sS[stage * DT_TILE + row];
            DT_T(t1);
            mbar_wait(&ctl->d_full[buf], (uint32_t)dph);
            tc_fence_after();
            DT_T(t2);
            DT_ACC(0, t1 - t0); DT_ACC(1, t2 - t1); DT_ACC(7, 1);
            const uint32_t dcol = tmem_base + lane_base + dt_dbuf_col(buf);
            const float4 *ll = reinterpret_cast<const float4 *>(sL + jc * cw);
#pragma unroll 1
            for (int hh = 0; hh < halves; ++hh) {
                uint32_t r[32], pk[16];
                DT_T(ta);
                tmem_ld32(dcol + hh * 32, r);
                tmem_wait_ld();
                DT_T(tb);
                DT_ACC(2, tb - ta);
                if (F2) {   // packed fp32 pairs: 2 FFMA2 instead of 4 FFMA per 4 logits, bitwise the same values
                    const uint64_t sn2 = f2_pack(sn, sn);
#pragma unroll
                    for (int c = 0; c < 32; c += 4) {
                        const float4 k4 = ll[hh * 8 + (c >> 2)];   // 14 - lse2 of four slices
                        float v0, v1, v2, v3;
                        f2_unpack(f2_fma(f2_pack(__uint_as_float(r[c]), __uint_as_float(r[c + 1])), sn2, f2_pack(k4.x, k4.y)), v0, v1);
                        f2_unpack(f2_fma(f2_pack(__uint_as_float(r[c + 2]), __uint_as_float(r[c + 3])), sn2, f2_pack(k4.z, k4.w)), v2, v3);
                        // pair index within the half: 2 * (c / 4) and + 1; pair j goes to the polynomial when poly_pair<POLY>(j) (tc.cuh)
                        float p0, p1, p2, p3;
                        if (poly_pair<POLY>(c >> 1)) poly_ex2_pair(v0, v1, p0, p1); else { p0 = fast_ex2(v0); p1 = fast_ex2(v1); }
                        if (poly_pair<POLY>((c >> 1) + 1)) poly_ex2_pair(v2, v3, p2, p3); else { p2 = fast_ex2(v2); p3 = fast_ex2(v3); }
                        const __half2 ha = __floats2half2_rn(p0, p1), hb = __floats2half2_rn(p2, p3);
                        pk[c >> 1] = *reinterpret_cast<const uint32_t *>(&ha);
                        pk[(c >> 1) + 1] = *reinterpret_cast<const uint32_t *>(&hb);
                    }
                } else {
#pragma unroll
                for (int c = 0; c < 32; c += 4) {
                    const float4 k4 = ll[hh * 8 + (c >> 2)];   // 14 - lse2 of four slices
                    const float p0 = fast_ex2(fmaf(__uint_as_float(r[c]), sn, k4.x)), p1 = fast_ex2(fmaf(__uint_as_float(r[c + 1]), sn, k4.y));
                    const float p2 = fast_ex2(fmaf(__uint_as_float(r[c + 2]), sn, k4.z)), p3 = fast_ex2(fmaf(__uint_as_float(r[c + 3]), sn, k4.w));
                    const __half2 ha = __floats2half2_rn(p0, p1), hb = __floats2half2_rn(p2, p3);
                    pk[c >> 1] = *reinterpret_cast<const uint32_t *>(&ha);
                    pk[(c >> 1) + 1] = *reinterpret_cast<const uint32_t *>(&hb);
                }
                }
                DT_T(tc_);
                tmem_st16(dcol + hh * 16, pk);   // packed P trails the logits it replaces (columns already read)
                DT_T(td);
                DT_ACC(3, tc_ - tb); DT_ACC(4, td - tc_);
            }
            DT_T(t3);
            tmem_wait_st();
            tc_fence_before();
            mbar_arrive(&ctl->p_full[buf]);
            DT_T(t4);
            DT_ACC(5, t4 - t3); DT_ACC(6, t4 - t0);
            buf += WG;
            if (buf >= DT_NBUF) { buf -= DT_NBUF; dph ^= 1; }   // buffer q % 3 and the parity (q / 3) & 1 of its barriers
            jc += WG;
            while (jc >= n_gc) {
                jc -= n_gc;
                ++t;
                if (++stage == n_stages) { stage = 0; sph ^= 1; }
            }
        }
        DT_FLUSH(0);
#ifdef HAET_TRACE
        { DT_T(l_end); if (lane == 0) { atomicAdd(&g_dt_trace[8], (unsigned long long)(l_start - k_start)); atomicAdd(&g_dt_trace[9], (unsigned long long)(l_end - l_start)); atomicAdd(&g_dt_trace[10], 1ull); } }
#endif
    }
    tc_fence_before();
    __syncthreads();
#ifdef HAET_TRACE
    if (threadIdx.x == 0) { atomicAdd(&g_dt_trace[11], (unsigned long long)(clock64() - k_start)); atomicAdd(&g_dt_trace[12], 1ull); }
#endif
    if (warp == W_MMA1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

bool deslice_tc_supported(const haet_layer_shape *s) {
    if (s->dim_head == 64) return s->slice_num <= 256;   // pair mode (two Ws images, double stages): shared memory holds G <= 256
    return (s->dim_head == DT_C || s->dim_head == 16) && s->slice_num <= DT_MAX_G;
}

int launch_deslice_tc(const haet_layer_shape *s, const uint8_t *img, const float *scale, const float *lse2,
                      const float *states, const float *ws, const float *bs, float *o, int o_image, cudaStream_t st) {
    if (s->N == 0 || s->B == 0) return HAET_OK;
    const int g_pad = dt_gpad(s->slice_num);
    const bool pair = s->dim_head == 64;
    const size_t smem = 1024 + (size_t)g_pad * (pair ? 384 : 256) + (size_t)(pair ? 2 * (DT_STAGES - 1) : DT_STAGES) * DT_TILE_BYTES + (size_t)g_pad * 32 +
                        DT_TILE * 32 + sizeof(float) * g_pad + sizeof(float) * DT_STAGES * DT_TILE + sizeof(DesliceCtl);
    HAET_DYN_SMEM((deslice_tc_kernel<true, 3, 0>), smem);
    HAET_DYN_SMEM((deslice_tc_kernel<true, 3, 1>), smem);
    HAET_DYN_SMEM((deslice_tc_kernel<true, 3, 2>), smem);
    HAET_DYN_SMEM((deslice_tc_kernel<true, 3, 3>), smem);
    HAET_DYN_SMEM((deslice_tc_kernel<true, 3, 4>), smem);
    HAET_DYN_SMEM((deslice_tc_kernel<false, 3, 0>), smem);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t BH = s->B * s->heads * (pair ? 2 : 1);   // pair mode: one CTA per (head, sub-row)
    int64_t n_chunks = sms / BH;
    if (n_chunks < 1) n_chunks = 1;
    const int64_t max_chunks = cdiv(s->N, DT_TILE);
    if (n_chunks > max_chunks) n_chunks = max_chunks;
    const int64_t chunk_pts = align_up(cdiv(s->N, n_chunks), DT_TILE);
    dim3 grid((unsigned)cdiv(s->N, chunk_pts), (unsigned)BH);
    constexpr int WG = 3;
    const int poly = options().no_f32x2 ? 0 : options().poly_p2;
    auto kern = options().no_f32x2 ? deslice_tc_kernel<false, WG, 0>
                : poly == 1 ? deslice_tc_kernel<true, WG, 1> : poly == 2 ? deslice_tc_kernel<true, WG, 2>
                : poly == 3 ? deslice_tc_kernel<true, WG, 3> : poly == 4 ? deslice_tc_kernel<true, WG, 4> : deslice_tc_kernel<true, WG, 0>;
    kern<<<grid, (4 * WG + 8) * 32, smem, st>>>(img, scale, lse2, states, ws, bs, rep_slice_shift(s->epsilon), s->N, xpi_npad(s->N),
                                               s->heads * (pair ? 2 : 1), s->slice_num, s->dim_head, chunk_pts, o, o_image);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet

#ifdef HAET_TRACE
extern "C" int haet_debug_dt_trace(unsigned long long *out64, int reset) {
    if (out64 && cudaMemcpyFromSymbol(out64, haet::g_dt_trace, sizeof(unsigned long long) * 64) != cudaSuccess) return -2;
    if (reset) { unsigned long long z[64] = {0}; if (cudaMemcpyToSymbol(haet::g_dt_trace, z, sizeof(z)) != cudaSuccess) return -2; }
    return 0;
}
#endif
