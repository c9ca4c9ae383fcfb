// Model prologue / epilogue around the blocks.
//
// Reference: models/HAETransolver_Irregular_Mesh.py:338-376 (Model.forward), :297-336 (get_grid), :131-132 (last block:
// mlp2(ln_3(fx))), models/components/mlp.py:67-96 (MLP, n_layers = 0), models/components/embedding.py:72-99
// (timestep_embedding); the structured models differ only in where the unified-position features come from
// (HAETransolver_Structured_Mesh_2D.py:384-410).
//
//   prologue:  fx0 = W2 . gelu(W1 . [feat_a | feat_f] + b1) + b2 + add            preprocess MLP + placeholder (+ time embedding)
//      feat_a = the coordinates x [M, Ka]   or, with ref > 0 (unified_pos of the irregular model), the ref^2 distances of
//      the 2-D coordinates to the ref x ref reference grid on [0, 1]^2, computed on the fly
//      kernel 1 (SIMT, the contraction is only Ka + Kf <= 128 deep): hidden activations written straight as the bf16 hi/lo
//      operand image (common.cuh: OPI) of kernel 2 = the CTA-pair tcgen05 GEMM (gemm2_tc.cu), 256 hidden columns per launch
//   time embedding: add[b, :] = time_fc(timestep_embedding(T[b]))  -- B rows, one CTA each
//   epilogue:  y = W . LayerNorm(fx) + b   with out_dim <= 8 outputs: one warp per row, one pass over fx
#include <cuda_bf16.h>

#include "common.cuh"
#include "gemm_tc.cuh"

namespace haet {

constexpr int PRO_THREADS = 256;
constexpr int PRO_ROWS = 8;      // rows per warp step (register block 8 rows x 8 hidden columns per lane)
constexpr int PRO_MAXK = 128;

// One CTA = one 128-row tile of the operand image; warp w owns rows [16 w, 16 w + 16).  Lane l owns the 8 hidden columns
// [256 cb + 8 l, + 8) of column block cb: its 16-byte hi chunk and lo chunk of the row; the 8 lanes of a panel write 128
// contiguous bytes.
__global__ void __launch_bounds__(PRO_THREADS)
prologue_fc1_kernel(const float *__restrict__ fa, int Ka, const float *__restrict__ ff, int Kf, int ref, int64_t M,
                    const float *__restrict__ w1, const float *__restrict__ b1, int hidden, uint8_t *__restrict__ opi) {
    extern __shared__ __align__(16) float psm[];
    const int K = (ref > 0 ? ref * ref : Ka) + Kf;
    float *sWt = psm;                         // [K][hidden]   W1 transposed: a lane's 8 columns are two float4
    float *sF = sWt + (size_t)K * hidden;     // [K][128]      the tile's features, k-major: 8 rows are two float4
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t m0 = (int64_t)blockIdx.x * 128;
    for (int i = tid; i < K * hidden; i += PRO_THREADS) {
        const int o = i / K, k = i - o * K;   // coalesced read of W1 [hidden, K]
        sWt[(size_t)k * hidden + o] = w1[i];
    }
    const int Kc = ref > 0 ? ref * ref : Ka;
    for (int i = tid; i < 128 * K; i += PRO_THREADS) {
        const int r = i / K, k = i - r * K;
        const int64_t m = m0 + r;
        float v = 0.f;
        if (m < M) {
            if (k >= Kc) v = ff[m * Kf + (k - Kc)];
            else if (ref > 0) {
                // get_grid (HAETransolver_Irregular_Mesh.py:297-336): reference point k = (i, j) = (k / ref, k % ref) at
                // (linspace(0, 1, ref)[i], linspace(0, 1, ref)[j]); feature = Euclidean distance of the point to it
                const int gi = k / ref, gj = k - gi * ref;
                const float step = ref > 1 ? 1.f / (float)(ref - 1) : 0.f;
                const float dx = fa[m * 2] - (float)gi * step, dy = fa[m * 2 + 1] - (float)gj * step;
                v = sqrtf(dx * dx + dy * dy);
            } else v = fa[m * Ka + k];
        }
        sF[k * 128 + r] = v;
    }
    __syncthreads();
    const int n_cb = hidden >> 8, panels = hidden >> 6;
    for (int cb = 0; cb < n_cb; ++cb) {
        const int col = cb * 256 + lane * 8;
        const float4 ba = *reinterpret_cast<const float4 *>(b1 + col), bb = *reinterpret_cast<const float4 *>(b1 + col + 4);
#pragma unroll 1
        for (int rb = 0; rb < 16; rb += PRO_ROWS) {
            const int r0 = warp * 16 + rb;
            float acc[PRO_ROWS][8];
#pragma unroll
            for (int r = 0; r < PRO_ROWS; ++r) {
                acc[r][0] = ba.x; acc[r][1] = ba.y; acc[r][2] = ba.z; acc[r][3] = ba.w;
                acc[r][4] = bb.x; acc[r][5] = bb.y; acc[r][6] = bb.z; acc[r][7] = bb.w;
            }
            for (int k = 0; k < K; ++k) {
                const float4 wa = *reinterpret_cast<const float4 *>(sWt + (size_t)k * hidden + col);
                const float4 wb = *reinterpret_cast<const float4 *>(sWt + (size_t)k * hidden + col + 4);
                const float4 fa4 = *reinterpret_cast<const float4 *>(sF + k * 128 + r0);        // warp-uniform: broadcast
                const float4 fb4 = *reinterpret_cast<const float4 *>(sF + k * 128 + r0 + 4);
                const float f[8] = {fa4.x, fa4.y, fa4.z, fa4.w, fb4.x, fb4.y, fb4.z, fb4.w};
                const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
                for (int r = 0; r < PRO_ROWS; ++r)
#pragma unroll
                    for (int c = 0; c < 8; ++c) acc[r][c] = fmaf(f[r], w[c], acc[r][c]);
            }
            const int panel = col >> 6, cc = (col & 63) >> 3;
#pragma unroll
            for (int r = 0; r < PRO_ROWS; ++r) {
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float v[2];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float y = acc[r][2 * c + e];
                        v[e] = 0.5f * y * (1.f + erff(y * 0.70710678118654752f));   // nn.GELU() (exact; mlp.py:11)
                    }
                    const __nv_bfloat162 hb = __floats2bfloat162_rn(v[0], v[1]);
                    const uint32_t ub = *reinterpret_cast<const uint32_t *>(&hb);
                    const __nv_bfloat162 lb = __floats2bfloat162_rn(v[0] - __uint_as_float(ub << 16), v[1] - __uint_as_float(ub & 0xffff0000u));
                    hi[c] = ub;
                    lo[c] = *reinterpret_cast<const uint32_t *>(&lb);
                }
                const int rr = r0 + r;   // every row of the tile is written (rows past M carry gelu(b1): finite, never read back)
                uint8_t *blk = opi + (((int64_t)blockIdx.x * panels + panel) * 2) * (int64_t)GT_PANEL_BYTES + tc::sw128_off(rr, cc);
                *reinterpret_cast<uint4 *>(blk) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4 *>(blk + GT_PANEL_BYTES) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
        }
    }
}

// add[b, o] = base[o] + time_fc(timestep_embedding(T[b], dim))[o]   (embedding.py:72-99; time_fc = Linear -> SiLU -> Linear,
// HAETransolver_Irregular_Mesh.py:220-223).  One CTA per batch element, thread o = output channel.
__global__ void __launch_bounds__(1024)
time_embed_kernel(const float *__restrict__ T, int dim, const float *__restrict__ w1, const float *__restrict__ b1,
                  const float *__restrict__ w2, const float *__restrict__ b2, const float *__restrict__ base,
                  float *__restrict__ add) {
    extern __shared__ float tsm[];
    float *emb = tsm, *hid = tsm + dim;
    const int b = blockIdx.x, half = dim / 2;
    const float t = T[b];
    for (int i = threadIdx.x; i < dim; i += blockDim.x) {
        float v = 0.f;
        if (i < 2 * half) {
            const int j = i < half ? i : i - half;
            const float freq = expf(-logf(10000.f) * (float)j / (float)half);   // fp32, as the reference computes it
            const float a = t * freq;
            v = i < half ? cosf(a) : sinf(a);
        }
        emb[i] = v;
    }
    __syncthreads();
    for (int o = threadIdx.x; o < dim; o += blockDim.x) {
        float acc = b1[o];
        for (int k = 0; k < dim; ++k) acc = fmaf(emb[k], w1[(size_t)o * dim + k], acc);
        hid[o] = acc / (1.f + expf(-acc));   // SiLU
    }
    __syncthreads();
    for (int o = threadIdx.x; o < dim; o += blockDim.x) {
        float acc = b2[o];
        for (int k = 0; k < dim; ++k) acc = fmaf(hid[k], w2[(size_t)o * dim + k], acc);
        add[(size_t)b * dim + o] = acc + (base ? base[o] : 0.f);
    }
}

// y[b, n, :] += add[b, :]
__global__ void add_batch_vec_kernel(float *__restrict__ y, const float *__restrict__ add, int64_t N, int dim4, int64_t total4) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total4) return;
    const int64_t row = i / dim4;
    const int c = (int)(i - row * dim4);
    const int64_t b = row / N;
    float4 v = reinterpret_cast<float4 *>(y)[i];
    const float4 a = reinterpret_cast<const float4 *>(add)[b * dim4 + c];
    v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    reinterpret_cast<float4 *>(y)[i] = v;
}

// y[m, o] = b[o] + sum_k W[o, k] LN(x[m, :])[k]   (last block: mlp2(ln_3(fx)), HAETransolver_Irregular_Mesh.py:131-132).
// One warp per row (dim % 128 == 0, dim <= 1024), out_dim <= 8; persistent grid, the next row's loads are in flight while
// the current row is reduced (one short-lived block per 8 rows measured 2.4 TB/s: block turnover, not bandwidth).
// N4 = float4 per lane held in registers: 2 for dim == 256, 8 for any dim <= 1024
template <int OUT, int N4>
__global__ void __launch_bounds__(256)
head_ln_linear_kernel(const float *__restrict__ x, const float *__restrict__ gamma, const float *__restrict__ beta, float eps,
                      const float *__restrict__ w, const float *__restrict__ bias, float *__restrict__ y, int64_t M, int dim) {
    const int lane = threadIdx.x & 31;
    const int n4 = dim >> 7;   // float4 per lane
    const int64_t stride = (int64_t)gridDim.x * 8;
    int64_t m = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    float4 nxt[N4];
    auto load = [&](int64_t row) {
#pragma unroll
        for (int i = 0; i < N4; ++i)
            if (i < n4 && row < M) nxt[i] = __ldg(reinterpret_cast<const float4 *>(x + row * dim) + i * 32 + lane);
    };
    load(m);
    for (; m < M; m += stride) {
        float4 v[N4];
        float s1 = 0.f;
#pragma unroll
        for (int i = 0; i < N4; ++i)
            if (i < n4) {
                v[i] = nxt[i];
                s1 += (v[i].x + v[i].y) + (v[i].z + v[i].w);
            }
        load(m + stride);
        // two more rows ahead into L2 (the register pipeline alone keeps ~64 KB per SM in flight: latency, not bandwidth, bound)
        if (m + 3 * stride < M && lane * 32 < dim) asm volatile("prefetch.global.L2 [%0];" ::"l"(x + (m + 3 * stride) * dim + lane * 32));
#pragma unroll
        for (int o = 16; o; o >>= 1) s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        const float mean = s1 / (float)dim;
        float s2 = 0.f;
#pragma unroll
        for (int i = 0; i < N4; ++i)
            if (i < n4) {
                v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
                s2 = fmaf(v[i].x, v[i].x, s2); s2 = fmaf(v[i].y, v[i].y, s2); s2 = fmaf(v[i].z, v[i].z, s2); s2 = fmaf(v[i].w, v[i].w, s2);
            }
#pragma unroll
        for (int o = 16; o; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        const float rstd = rsqrtf(s2 / (float)dim + eps);
        float acc[OUT];
#pragma unroll
        for (int o = 0; o < OUT; ++o) acc[o] = 0.f;
#pragma unroll
        for (int i = 0; i < N4; ++i)
            if (i < n4) {
                const int k = (i * 32 + lane) * 4;
                const float4 g = __ldg(reinterpret_cast<const float4 *>(gamma + k)), bt = __ldg(reinterpret_cast<const float4 *>(beta + k));
                const float n0 = fmaf(v[i].x * rstd, g.x, bt.x), n1 = fmaf(v[i].y * rstd, g.y, bt.y);
                const float n2 = fmaf(v[i].z * rstd, g.z, bt.z), n3 = fmaf(v[i].w * rstd, g.w, bt.w);
#pragma unroll
                for (int o = 0; o < OUT; ++o) {
                    const float4 wv = __ldg(reinterpret_cast<const float4 *>(w + (size_t)o * dim + k));
                    acc[o] = fmaf(n0, wv.x, acc[o]); acc[o] = fmaf(n1, wv.y, acc[o]);
                    acc[o] = fmaf(n2, wv.z, acc[o]); acc[o] = fmaf(n3, wv.w, acc[o]);
                }
            }
#pragma unroll
        for (int o = 0; o < OUT; ++o) {
#pragma unroll
            for (int s = 16; s; s >>= 1) acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], s);
        }
        if (lane == 0) {
#pragma unroll
            for (int o = 0; o < OUT; ++o) y[m * OUT + o] = acc[o] + bias[o];
        }
    }
}

static bool prologue_supported(int K, int hidden, int dim) {
    return hooks().path == 0 && !options().gemm_1cta && dim == 256 && hidden >= 256 && hidden % 256 == 0 && hidden <= 2048 &&
           K >= 1 && K <= PRO_MAXK && (size_t)K * (hidden + 128) * 4 <= 200 * 1024;
}

}  // namespace haet

using namespace haet;

extern "C" {

int64_t haet_model_prologue_workspace_bytes(int64_t M, int32_t Ka, int32_t Kf, int32_t ref, int32_t hidden, int32_t dim) {
    const int K = (ref > 0 ? ref * ref : Ka) + Kf;
    if (M < 0 || !prologue_supported(K, hidden, dim)) {
        set_error("haet_model_prologue: unsupported shape (features %d, hidden %d, dim %d; needs dim 256, hidden a multiple of 256, "
                  "<= %d input features, tensor-core path)", K, hidden, dim, PRO_MAXK);
        return HAET_ENOSUP;
    }
    const int64_t img_b = gemm_tc_image_bytes(256, dim) > 4 * (int64_t)dim * hidden ? gemm_tc_image_bytes(256, dim) : 4 * (int64_t)dim * hidden;
    return align_up(opi_bytes(M, hidden), 256) + align_up(img_b, 256) + 256 * 4;
}

int haet_model_prologue(const float *feat_a_d, int32_t Ka, const float *feat_f_d, int32_t Kf, int32_t ref, int64_t B, int64_t N,
                        int32_t hidden, int32_t dim, const float *w1_d, const float *b1_d, const float *w2_d, const float *b2_d,
                        const float *add_d, int32_t add_per_batch, float *out_d, void *workspace_d, void *stream) {
    const int K = (ref > 0 ? ref * ref : Ka) + Kf;
    HAET_CHECK_ARG(B >= 0 && N >= 0 && Ka >= 0 && Kf >= 0 && ref >= 0, "haet_model_prologue: negative size");
    HAET_CHECK_ARG(ref == 0 || Ka == 2, "haet_model_prologue: the reference-grid distances (unified_pos) need 2-D coordinates");
    HAET_CHECK_ARG((Ka == 0 || feat_a_d) && (Kf == 0 || feat_f_d) && w1_d && b1_d && w2_d && b2_d && out_d && workspace_d,
                   "haet_model_prologue: null argument");
    if (!prologue_supported(K, hidden, dim)) {
        set_error("haet_model_prologue: unsupported shape (features %d, hidden %d, dim %d)", K, hidden, dim);
        return HAET_ENOSUP;
    }
    const int64_t M = B * N;
    if (M == 0) return HAET_OK;
    cudaStream_t st = (cudaStream_t)stream;
    char *p = (char *)workspace_d;
    uint8_t *h = (uint8_t *)p;      p += align_up(opi_bytes(M, hidden), 256);
    const int64_t img_b = gemm_tc_image_bytes(256, dim) > 4 * (int64_t)dim * hidden ? gemm_tc_image_bytes(256, dim) : 4 * (int64_t)dim * hidden;
    uint8_t *img = (uint8_t *)p;    p += align_up(img_b, 256);
    float *bias0 = (float *)p;
    const size_t smem = (size_t)K * (hidden + 128) * sizeof(float);
    if (smem > 48 * 1024) HAET_DYN_SMEM(prologue_fc1_kernel, smem);
    prologue_fc1_kernel<<<(unsigned)cdiv(M, 128), PRO_THREADS, smem, st>>>(feat_a_d, Ka, feat_f_d, Kf, ref, M, w1_d, b1_d, hidden, h);
    HAET_LAUNCHED();
    // fc2 on the CTA-pair GEMM, 256 hidden columns per launch; the batch-independent additive vector rides in the bias
    const float *bias = b2_d;
    if (add_d && !add_per_batch) {
        HAET_CUDA(cudaMemcpyAsync(bias0, b2_d, sizeof(float) * dim, cudaMemcpyDeviceToDevice, st));
        if (int rc = launch_add_rows(bias0, add_d, dim, st)) return rc;
        bias = bias0;
    }
    int rc;
    if (hidden >= 512 && !options().mlp_fc2_split && gemm_stream_opi_supported(hidden, dim)) {
        // one pass over the hidden activations with W2 streamed from L2 (conv_tc.cu: launch_gemm_stream_opi), as the block MLP's fc2
        if ((rc = launch_gemm_stream_opi(h, w2_d, bias, nullptr, out_d, M, hidden, dim, img, st))) return rc;
    } else
    for (int j = 0; j < hidden / 256; ++j) {
        if ((rc = launch_gemm_tc_pack(w2_d + j * 256, dim, 256, img, st, hidden))) return rc;
        GemmEpi ep{};
        ep.a_img = 1; ep.a_panels = hidden / 64; ep.a_panel0 = 4 * j;
        ep.resid = j == 0 ? nullptr : out_d;   // in place after the first block: every element is read, then written, by the same thread
        if ((rc = launch_gemm2_tc(EPI_F32, reinterpret_cast<const float *>(h), img, j == 0 ? bias : nullptr, out_d, M, 256, dim, ep, st))) return rc;
    }
    if (add_d && add_per_batch) {
        const int64_t total4 = M * (dim / 4);
        add_batch_vec_kernel<<<(unsigned)cdiv(total4, 256), 256, 0, st>>>(out_d, add_d, N, dim / 4, total4);
        HAET_LAUNCHED();
    }
    return HAET_OK;
}

int haet_time_embed(const float *t_d, int64_t B, int32_t dim, const float *w1_d, const float *b1_d, const float *w2_d,
                    const float *b2_d, const float *base_d, float *add_d, void *stream) {
    HAET_CHECK_ARG(B >= 0 && dim > 0 && dim <= 4096, "haet_time_embed: bad size");
    HAET_CHECK_ARG(t_d && w1_d && b1_d && w2_d && b2_d && add_d, "haet_time_embed: null argument");
    if (B == 0) return HAET_OK;
    time_embed_kernel<<<(unsigned)B, dim < 1024 ? (dim + 31) / 32 * 32 : 1024, 2 * dim * sizeof(float), (cudaStream_t)stream>>>(
        t_d, dim, w1_d, b1_d, w2_d, b2_d, base_d, add_d);
    HAET_LAUNCHED();
    return HAET_OK;
}

int haet_head_ln_linear(const float *x_d, int64_t M, int32_t dim, const float *ln_g_d, const float *ln_b_d, float ln_eps,
                        const float *w_d, const float *b_d, int32_t out_dim, float *y_d, void *stream) {
    HAET_CHECK_ARG(M >= 0 && x_d && ln_g_d && ln_b_d && w_d && b_d && y_d, "haet_head_ln_linear: null argument");
    HAET_CHECK_ARG(ln_eps > 0.f, "haet_head_ln_linear: eps must be positive");
    if (dim % 128 != 0 || dim > 1024 || out_dim < 1 || out_dim > 8) {
        set_error("haet_head_ln_linear: unsupported dim=%d out_dim=%d (dim a multiple of 128 up to 1024, out_dim 1..8)", dim, out_dim);
        return HAET_ENOSUP;
    }
    if (M == 0) return HAET_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t want = cdiv(M, 8);
    const unsigned grid = (unsigned)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
#define HAET_HEAD(O)                                                                                                  \
    case O:                                                                                                           \
        if (dim == 256) head_ln_linear_kernel<O, 2><<<grid, 256, 0, st>>>(x_d, ln_g_d, ln_b_d, ln_eps, w_d, b_d, y_d, M, dim); \
        else head_ln_linear_kernel<O, 8><<<grid, 256, 0, st>>>(x_d, ln_g_d, ln_b_d, ln_eps, w_d, b_d, y_d, M, dim);            \
        break
    switch (out_dim) {
        HAET_HEAD(1); HAET_HEAD(2); HAET_HEAD(3); HAET_HEAD(4); HAET_HEAD(5); HAET_HEAD(6); HAET_HEAD(7); HAET_HEAD(8);
    }
#undef HAET_HEAD
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // extern "C"
