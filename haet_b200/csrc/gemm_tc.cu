// tcgen05 GEMM with fp32-class accuracy: Y[M, Nn] = X[M, K] . W[Nn, K]^T + bias, X fp32 in HBM.
// Used for in_project_x (IrregularMesh.py:157) and to_out (:235).
//
// Accuracy: the Rep-Slice softmax divides the logits by tau >= 0.1 and exponentiates, so rounding the
// projection's operands to bf16 (or tf32) costs 3e-2 (7e-3) relative error on the layer output
// (scripts/precision_study.py).  Both operands are therefore split x = hi + lo into two bf16 and the
// product is accumulated as hi.hi + lo.hi + hi.lo in the fp32 TMEM accumulator (error ~2^-17).
//
// Mapping (persistent, one CTA per SM): CTA = (worker, column half).  The CTA's half of W (hi and lo,
// <= 128 KB) is bulk-copied ONCE into shared memory as ready-made SWIZZLE_128B K-major panels and stays
// resident; the CTA then walks 128-row tiles of X.  Warp roles: 8 producer warps (coalesced LDG of fp32 X two
// panels ahead in registers, hi/lo split, swizzled STS), 1 MMA warp (one thread issues tcgen05.mma, accumulators
// double-buffered in TMEM), 4 epilogue warps (tcgen05.ld, + bias, store).  mbarrier pipelines: a_full/a_empty
// per stage, t_full/t_empty per accumulator buffer.
// Epilogues: EPI_F32 stores fp32 rows (to_out); EPI_IMG is the in-projection of the tensor-core path: per head it
// adds the bias, evaluates Ada-Temp (s_n = log2 e / tau_n) and writes the spill's fp16 hi/lo operand image
// (common.cuh) instead of fp32 rows, so that passes 1 and 2 bulk-copy their MMA operands.
#include <stdlib.h>

#include "gemm_tc.cuh"

namespace haet {
using namespace tc;

// W [Nn, K] fp32 -> image [Nn/128 halves][K/64 panels][hi, lo][128 rows x 128 B swizzled]
__global__ void gemm_tc_pack_kernel(const float *__restrict__ w, int Nn, int K, uint8_t *__restrict__ img, int ldw) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 8 consecutive k (one 16-B chunk)
    const int chunks_per_row = K / 8;
    if (idx >= Nn * chunks_per_row) return;
    const int n = idx / chunks_per_row, ck = idx - n * chunks_per_row;
    const int half = n / GT_N, row = n % GT_N;
    const int k0 = ck * 8, panel = k0 / GT_KP, chunk = (k0 % GT_KP) / 8;
    const int panels = K / GT_KP;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) split_bf16(w[(size_t)n * ldw + k0 + i], hi[i], lo[i]);
    uint8_t *base = img + ((size_t)half * panels + panel) * (2 * GT_PANEL_BYTES) + sw128_off(row, chunk);
    *reinterpret_cast<uint4 *>(base) = *reinterpret_cast<const uint4 *>(hi);
    *reinterpret_cast<uint4 *>(base + GT_PANEL_BYTES) = *reinterpret_cast<const uint4 *>(lo);
}

int launch_gemm_tc_pack(const float *w, int Nn, int K, uint8_t *img, cudaStream_t st, int ldw) {
    const int chunks = Nn * (K / 8);
    gemm_tc_pack_kernel<<<(chunks + 255) / 256, 256, 0, st>>>(w, Nn, K, img, ldw ? ldw : K);
    HAET_LAUNCHED();
    return HAET_OK;
}

// zero the pad rows [N, npad) of every (b, h) of the spill image and their scales
__global__ void xpi_pad_kernel(uint8_t *__restrict__ img, float *__restrict__ scale, int64_t N, int64_t npad, int64_t bh_count) {
    const int64_t pad = npad - N;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 16-B chunk of a pad row
    if (t >= bh_count * pad * 8) return;
    const int64_t rowi = t >> 3, bh = rowi / pad, n = N + (rowi - bh * pad);
    const int64_t row = bh * npad + n;
    *reinterpret_cast<uint4 *>(img + row * XPI_ROW_BYTES + ((t & 7) << 4)) = make_uint4(0, 0, 0, 0);
    if ((t & 7) == 0) scale[row] = 0.f;
}

template <int EPI>
__global__ void __launch_bounds__(GT_THREADS, 1)
gemm_tc_kernel(const float *__restrict__ X, const uint8_t *__restrict__ Wimg, const float *__restrict__ bias,
               float *__restrict__ Y, int64_t M, int K, int Nn, int n_workers, GemmEpi ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS / STS)
    const int panels = K / GT_KP;
    const int n_halves = Nn / GT_N;
    uint8_t *sW = smem;                                         // [panels][hi|lo][16 KB]
    uint8_t *sA = sW + (size_t)panels * 2 * GT_PANEL_BYTES;     // [stages][hi|lo][16 KB]
    GemmSmem *cb = reinterpret_cast<GemmSmem *>(sA + (size_t)GT_STAGES * 2 * GT_PANEL_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int half = blockIdx.x % n_halves, worker = blockIdx.x / n_halves;
    const int64_t n_tiles = (M + GT_M - 1) / GT_M;
    const int64_t my_tiles = worker < n_tiles ? (n_tiles - worker + n_workers - 1) / n_workers : 0;

    if (warp == 8) {
        if (lane == 0) {
            mbar_init(&cb->w_full, 1);
            for (int s = 0; s < GT_STAGES; ++s) { mbar_init(&cb->a_full[s], GT_PROD); mbar_init(&cb->a_empty[s], 1); }
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

    if (warp < 8) {
        // ---------------- producers: X tile panel -> bf16 hi/lo, SWIZZLE_128B K-major ----------------
        gemm_produce(X, M, K, panels, my_tiles, sA,
                     [&](int64_t ti) { return (worker + ti * n_workers) * (int64_t)GT_M; },
                     [&](int st, int parity) { mbar_wait(&cb->a_empty[st], (uint32_t)parity); },
                     [&](int st) { mbar_arrive(&cb->a_full[st]); }, 0, NoPanelHook(), ep.pf);
    } else if (warp == 8) {
        // ---------------- MMA issuer ----------------
        if (lane == 0 && my_tiles > 0) {
            const uint32_t wbytes = (uint32_t)panels * 2 * GT_PANEL_BYTES;
            mbar_arrive_expect_tx(&cb->w_full, wbytes);
            const uint8_t *src = Wimg + (size_t)half * wbytes;
            for (int c = 0; c < panels * 2; ++c)
                bulk_g2s(sW + (size_t)c * GT_PANEL_BYTES, src + (size_t)c * GT_PANEL_BYTES, GT_PANEL_BYTES, &cb->w_full);
            mbar_wait(&cb->w_full, 0);
            const uint32_t idesc = instr_desc(FMT_BF16, GT_M, GT_N, false, false);
            int stage = 0, phase = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
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
        const int q = warp & 3;   // TMEM lane quarter this warp may access (warps 9..12 -> 1, 2, 3, 0)
        for (int64_t it = 0; it < my_tiles; ++it) {
            const int64_t tile = worker + it * n_workers;
            const int buf = (int)(it & 1);
            mbar_wait(&cb->t_full[buf], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            const int64_t m0 = tile * GT_M + q * 32 + (lane & ~3);   // first of this lane's 4 rows after the transpose
            int64_t bi0 = 0, n0 = 0;
            if (EPI == EPI_IMG) { bi0 = m0 / ep.N; n0 = m0 - bi0 * ep.N; }
#pragma unroll 1
            for (int c0 = 0; c0 < GT_N; c0 += 32) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * GT_N + c0), r);
                tmem_wait_ld();
                gemm_epilogue_chunk<EPI>(r, lane, m0, bi0, n0, M, bias + half * GT_N + c0, ep.wt, Y, Nn, half * GT_N + c0, ep);
            }
            tc_fence_before();
            mbar_arrive(&cb->t_empty[buf]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

bool gemm_tc_supported(int K, int Nn) { return K % GT_KP == 0 && K >= GT_KP && K <= 256 && Nn % GT_N == 0 && Nn >= GT_N; }

int64_t gemm_tc_image_bytes(int K, int Nn) { return (int64_t)Nn * K * 2 * 2; }

static int sm_count() {   // of the CURRENT device (no process-wide cache: devices may differ)
    int dev = 0, n = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n > 0 ? n : 148;
}

// img: scratch of gemm_tc_image_bytes(K, Nn) bytes (the packed weights; rebuilt each call, ~2 us)
template <int EPI>
static int launch_gemm_impl(const float *x, const float *w, const float *bias, float *y, int64_t M, int K, int Nn, void *img,
                            const GemmEpi &ep, cudaStream_t st) {
    if (M == 0) return HAET_OK;
    HAET_CHECK_ARG(gemm_tc_supported(K, Nn), "gemm_tc: unsupported K=%d Nn=%d", K, Nn);
    const int chunks = Nn * (K / 8);
    gemm_tc_pack_kernel<<<(chunks + 255) / 256, 256, 0, st>>>(w, Nn, K, (uint8_t *)img, K);
    HAET_LAUNCHED();
    if (gemm2_tc_supported(K, Nn) && !options().gemm_1cta)
        return launch_gemm2_tc(EPI, x, (const uint8_t *)img, bias, y, M, K, Nn, ep, st);
    const int panels = K / GT_KP, n_halves = Nn / GT_N;
    const size_t smem = 1024 + (size_t)(panels + GT_STAGES) * 2 * GT_PANEL_BYTES + sizeof(GemmSmem);
    HAET_DYN_SMEM(gemm_tc_kernel<EPI>, smem);
    const int64_t n_tiles = (M + GT_M - 1) / GT_M;
    int n_workers = sm_count() / n_halves;
    if (n_workers < 1) n_workers = 1;
    if (n_workers > n_tiles) n_workers = (int)n_tiles;
    GemmEpi epd = ep;
    epd.pf = options().gemm_pf;
    gemm_tc_kernel<EPI><<<n_workers * n_halves, GT_THREADS, smem, st>>>(x, (const uint8_t *)img, bias, y, M, K, Nn, n_workers, epd);
    HAET_LAUNCHED();
    if (EPI == EPI_F32 && ep.resid) return launch_add_rows(y, ep.resid, M * Nn, st);   // only the pair kernel fuses the residual
    return HAET_OK;
}

int launch_gemm_tc(const float *x, const float *w, const float *bias, float *y, int64_t M, int K, int Nn, void *img,
                   cudaStream_t st, const float *resid) {
    GemmEpi ep{};
    ep.resid = resid;
    return launch_gemm_impl<EPI_F32>(x, w, bias, y, M, K, Nn, img, ep, st);
}

// to_out with A given as the operand image pass 2 wrote (common.cuh: OPI); only the CTA-pair kernel reads it
bool gemm_opi_supported(int K, int Nn) { return gemm2_tc_supported(K, Nn) && !options().gemm_1cta && !options().no_opi; }
int launch_gemm_tc_opi(const uint8_t *a_img, const float *w, const float *bias, float *y, int64_t M, int K, int Nn, void *img,
                       cudaStream_t st, const float *resid) {
    HAET_CHECK_ARG(gemm_opi_supported(K, Nn), "gemm_tc_opi: unsupported K=%d Nn=%d", K, Nn);
    GemmEpi ep{};
    ep.a_img = 1;
    ep.resid = resid;
    return launch_gemm_impl<EPI_F32>(reinterpret_cast<const float *>(a_img), w, bias, y, M, K, Nn, img, ep, st);
}

// ---- block fusion helpers ------------------------------------------------------------------------------
// ln_1 folded into the in-projection (one block per output row o, K <= 256 threads each own one k):
//   Wf[o, k] = W[o, k] gamma[k],  bf[o] = b[o] + sum_k W[o, k] beta[k],
//   wsum[o] = sum_k (hi + lo)(Wf[o, k])  -- of the bf16 pair the MMA actually multiplies, so that the mean term cancels exactly
__global__ void ln_fold_kernel(const float *__restrict__ w, const float *__restrict__ b, const float *__restrict__ gamma,
                               const float *__restrict__ beta, int K, float *__restrict__ wf, float *__restrict__ bf,
                               float *__restrict__ wsum) {
    __shared__ float sa[256], sb[256];
    const int o = blockIdx.x, k = threadIdx.x;
    float a = 0.f, c = 0.f;
    if (k < K) {
        const float wv = w[(size_t)o * K + k], f = wv * gamma[k];
        wf[(size_t)o * K + k] = f;
        __nv_bfloat16 hi, lo;
        split_bf16(f, hi, lo);
        a = __bfloat162float(hi) + __bfloat162float(lo);
        c = wv * beta[k];
    }
    sa[k] = a; sb[k] = c;
    __syncthreads();
    for (int o2 = 128; o2; o2 >>= 1) {   // fixed-order tree: bit-reproducible
        if (k < o2) { sa[k] += sa[k + o2]; sb[k] += sb[k + o2]; }
        __syncthreads();
    }
    if (k == 0) { wsum[o] = sa[0]; bf[o] = b[o] + sb[0]; }
}

bool ln_fused_supported(const haet_layer_shape *s) {
    const int inner = s->heads * s->dim_head;
    return hooks().path == 0 && s->conv_nd == 0 && s->dim_head == 32 && gemm2_tc_supported(s->dim, inner) && !options().gemm_1cta &&
           !options().no_ln_fuse;
}
static int64_t ln_fold_floats(const haet_layer_shape *s) {
    const int64_t inner = s->heads * s->dim_head;
    return align_up(inner * s->dim, 64) + 2 * align_up(inner, 64);
}
int64_t ln_fused_scratch_bytes(const haet_layer_shape *s) {
    return align_up(ln_fold_floats(s) * 4 + s->B * s->N * (int64_t)sizeof(float2), 256);
}

// nn.LayerNorm over the last dimension, one warp per row (the unfused route: conv front-ends, SIMT path, other shapes)
__global__ void layernorm_rows_kernel(const float *__restrict__ x, const float *__restrict__ gamma, const float *__restrict__ beta,
                                      float eps, float *__restrict__ y, int64_t M, int K) {
    const int64_t m = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (m >= M) return;
    const float *xr = x + m * K;
    float s1 = 0.f;
    for (int k = lane; k < K; k += 32) s1 += xr[k];
#pragma unroll
    for (int o = 16; o; o >>= 1) s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    const float mean = s1 / (float)K;
    float s2 = 0.f;
    for (int k = lane; k < K; k += 32) { const float d = xr[k] - mean; s2 = fmaf(d, d, s2); }
#pragma unroll
    for (int o = 16; o; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    const float rstd = rsqrtf(s2 / (float)K + eps);
    for (int k = lane; k < K; k += 32) y[m * K + k] = (xr[k] - mean) * rstd * gamma[k] + beta[k];
}
int launch_layernorm_rows(const float *x, const float *gamma, const float *beta, float eps, float *y, int64_t M, int K, cudaStream_t st) {
    if (M == 0) return HAET_OK;
    layernorm_rows_kernel<<<(unsigned)cdiv(M, 8), 256, 0, st>>>(x, gamma, beta, eps, y, M, K);
    HAET_LAUNCHED();
    return HAET_OK;
}
__global__ void add_rows_kernel(float *__restrict__ y, const float *__restrict__ r, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] += r[i];
}
int launch_add_rows(float *y, const float *resid, int64_t n, cudaStream_t st) {
    if (n == 0) return HAET_OK;
    add_rows_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(y, resid, n);
    HAET_LAUNCHED();
    return HAET_OK;
}

int launch_xpi_pad(const haet_layer_shape *s, uint8_t *xpi, float *scale, cudaStream_t st) {
    const int64_t npad = xpi_npad(s->N), pad = npad - s->N, bh = s->B * s->heads;
    if (pad > 0 && bh > 0) {
        xpi_pad_kernel<<<(unsigned)cdiv(bh * pad * 8, 256), 256, 0, st>>>(xpi, scale, s->N, npad, bh);
        HAET_LAUNCHED();
    }
    return HAET_OK;
}

// in_project_x of the tensor-core path: X [B*N, dim] -> the spill's operand image + scales (no fp32 x_proj)
int launch_inproj_tc(const haet_layer_shape *s, const float *x, const float *wx, const float *bx, const float *wt, float bt,
                     uint8_t *xpi, float *scale, void *img, cudaStream_t st, const LnArgs *ln) {
    HAET_CHECK_ARG(s->dim_head == 32 || s->dim_head == 16, "the operand image needs dim_head 16 or 32");
    GemmEpi ep{};
    ep.c16 = s->dim_head == 16;
    if (ln) {
        HAET_CHECK_ARG(ln_fused_supported(s), "fused LayerNorm: unsupported shape");
        const int inner = s->heads * s->dim_head;
        float *wf = (float *)ln->scratch, *bf = wf + align_up((int64_t)inner * s->dim, 64), *wsum = bf + align_up(inner, 64);
        ln_fold_kernel<<<inner, 256, 0, st>>>(wx, bx, ln->gamma, ln->beta, s->dim, wf, bf, wsum);
        HAET_LAUNCHED();
        ep.wsum = wsum;
        ep.ln_stats = reinterpret_cast<float2 *>(wsum + align_up(inner, 64));
        ep.ln_eps = ln->eps;
        wx = wf;
        bx = bf;
    }
    ep.img = xpi; ep.scale = scale; ep.wt = wt; ep.bt = bt; ep.base_temp = s->base_temp;
    ep.N = s->N; ep.npad = xpi_npad(s->N); ep.H = s->heads;
    if (int rc = launch_xpi_pad(s, xpi, scale, st)) return rc;
    return launch_gemm_impl<EPI_IMG>(x, wx, bx, nullptr, s->B * s->N, s->dim, s->heads * s->dim_head, img, ep, st);
}

// ---- block MLP: out = W2 . gelu(W1 . LayerNorm(x) + b1) + b2 + x  (HAETransolver_Irregular_Mesh.py:128,
// models/components/mlp.py:67-96 with n_layers = 0), on the CTA-pair kernel, 256 hidden columns per launch:
//   fc1 block j: ln_2 fused as in the in-projection, epilogue = GELU -> panels [4 j, 4 j + 4) of the operand image h
//   fc2 block j: A = those panels (bulk-copied), out (+)= h_j . W2[:, 256 j : 256 j + 256]^T (+ b2 + x for j = 0, + out after)
bool mlp_tc_supported(int dim, int hidden) {
    return hooks().path == 0 && dim == 256 && hidden >= 256 && hidden % 256 == 0 && hidden <= 2048 && !options().gemm_1cta;
}
int64_t mlp_tc_workspace_bytes(int64_t M, int dim, int hidden) {
    const int64_t img_b = gemm_tc_image_bytes(dim, 256) > 4 * (int64_t)dim * hidden ? gemm_tc_image_bytes(dim, 256) : 4 * (int64_t)dim * hidden;
    return align_up(opi_bytes(M, hidden), 256) + align_up(img_b, 256) +
           align_up((int64_t)(256 * dim + 2 * 256) * 4, 256) + align_up(M * (int64_t)sizeof(float2), 256);
}
int launch_mlp_tc(const float *x, int64_t M, int dim, int hidden, const float *ln_g, const float *ln_b, float ln_eps,
                  const float *w1, const float *b1, const float *w2, const float *b2, float *out, void *workspace, cudaStream_t st) {
    HAET_CHECK_ARG(mlp_tc_supported(dim, hidden), "block MLP: unsupported dim=%d hidden=%d (dim 256, hidden a multiple of 256)", dim, hidden);
    if (M == 0) return HAET_OK;
    char *p = (char *)workspace;
    uint8_t *h = (uint8_t *)p;      p += align_up(opi_bytes(M, hidden), 256);
    const int64_t img_b = gemm_tc_image_bytes(dim, 256) > 4 * (int64_t)dim * hidden ? gemm_tc_image_bytes(dim, 256) : 4 * (int64_t)dim * hidden;
    uint8_t *img = (uint8_t *)p;    p += align_up(img_b, 256);
    float *wf = (float *)p;         p += align_up((int64_t)(256 * dim + 2 * 256) * 4, 256);
    float *bf = wf + 256 * dim, *wsum = bf + 256;
    float2 *stats = (float2 *)p;
    const int blocks = hidden / 256;
    int rc;
    for (int j = 0; j < blocks; ++j) {
        ln_fold_kernel<<<256, 256, 0, st>>>(w1 + (size_t)j * 256 * dim, b1 + j * 256, ln_g, ln_b, dim, wf, bf, wsum);
        HAET_LAUNCHED();
        if ((rc = launch_gemm_tc_pack(wf, 256, dim, img, st))) return rc;
        GemmEpi ep{};
        ep.wsum = wsum; ep.ln_stats = stats; ep.ln_eps = ln_eps; ep.ln_reuse = j > 0;
        ep.opi = h; ep.opi_panels = hidden / 64; ep.opi_panel0 = 4 * j;
        if ((rc = launch_gemm2_tc(EPI_OPI, x, img, bf, nullptr, M, dim, 256, ep, st))) return rc;
    }
    // fc2.  hidden >= 512: ONE pass with W2 streamed from L2 (conv_tc.cu: launch_gemm_stream_opi) -- h is read once, out written
    // once: 4 GB per 1M rows instead of 6 GB for one CTA-pair launch per 256 hidden columns with `out` read back in between
    if (blocks >= 2 && !options().mlp_fc2_split && gemm_stream_opi_supported(hidden, dim))
        return launch_gemm_stream_opi(h, w2, b2, x, out, M, hidden, dim, img, st);
    for (int j = 0; j < blocks; ++j) {
        if ((rc = launch_gemm_tc_pack(w2 + j * 256, dim, 256, img, st, hidden))) return rc;
        GemmEpi ep{};
        ep.a_img = 1; ep.a_panels = hidden / 64; ep.a_panel0 = 4 * j;
        ep.resid = j == 0 ? x : out;   // in place after the first block: every element is read, then written, by the same thread
        if ((rc = launch_gemm2_tc(EPI_F32, reinterpret_cast<const float *>(h), img, j == 0 ? b2 : nullptr, out, M, 256, dim, ep, st))) return rc;
    }
    return HAET_OK;
}

}  // namespace haet
