// Conv2d / Conv3d front-end (kernel k, padding k/2; StructuredMesh2D.py:109,173-189, StructuredMesh3D.py:113,177-193) as
// an implicit GEMM on tcgen05:
//   xp[m, o] = b[o] + sum_tap sum_ci x[nbr(m, tap), ci] W[o, tap, ci]        M = B * N grid points, K = taps * dim
// with the 3-term bf16 split of gemm_tc.cu (fp32-class accuracy, fp32 accumulation in TMEM).  Same roles as the dense
// GEMM -- 8 producer warps gather and split the A rows (the neighbour's channels, zeros outside the grid), one MMA warp,
// 4 epilogue warps -- except that W (taps * dim * 4 bytes per output row: 7 MB at 64^3 x 256) cannot stay in shared
// memory: a loader warp streams it, one pre-swizzled 32 KB [hi | lo] panel per (tap, 64 channels), through the same
// 3-stage ring as A.  One CTA = NH 128-column halves of 128-row tiles (NH = 2 when the output has a multiple of 256
// columns: the A rows are then gathered and converted once for 256 columns, UMMA N = 256, W through a 2-stage ring);
// persistent grid.
#include "gemm_tc.cuh"

namespace haet {
using namespace tc;

constexpr int CT_THREADS = GT_PROD + 32 + 128 + 32;   // producers (0-7), MMA warp (8), epilogue warps (9-12), W loader (13)
constexpr int CT_MAX_PANELS = 1024;                   // entries of the per-panel tap table in shared memory (16 KB)

struct ConvSmem {
    uint64_t a_full[GT_STAGES], a_empty[GT_STAGES], w_full[GT_STAGES], w_empty[GT_STAGES];
    uint64_t t_full[2], t_empty[2];
    uint32_t tmem_base;
};

struct ConvGeom {
    int64_t N;            // grid points per batch element
    int g0, g1, g2;       // grid extents (g2 = 1 for 2D), point n = (i * g1 + j) * g2 + k
    int ks, nd, din;      // kernel size, 2 or 3, input channels
};

// IMG: the epilogue writes the spill's fp16 hi / lo operand image + Ada-Temp scales directly (gemm_tc.cuh: EPI_IMG, as the Linear
// front-end does) instead of fp32 x_proj that xpi_pack_kernel would read back: no x_proj round trip, one launch less.
// AIMG: X is not fp32 rows but a bf16 hi / lo operand image (common.cuh: OPI, [tile][k panel][hi | lo][128 x 128 B]) whose
// (tile, panel) blocks are bulk-copied straight into the A stages -- a plain GEMM with K too deep for resident weights
// (block MLP fc2, K = hidden >= 512: W streamed from L2, one pass over A); the epilogue adds bias and the residual rows.
template <int NH, bool IMG, bool AIMG = false, bool RES = false>
__global__ void __launch_bounds__(CT_THREADS, 1)
conv_tc_kernel(const float *__restrict__ X, const uint8_t *__restrict__ Wimg, const float *__restrict__ bias,
               float *__restrict__ Y, int64_t M, int Nn, int n_workers, ConvGeom geo, GemmEpi epi) {
    __shared__ __align__(16) float sWtc[32];   // Ada-Temp weights of one head (twice for dim_head 16)
    if (IMG && threadIdx.x < 32) sWtc[threadIdx.x] = epi.wt[epi.c16 ? (threadIdx.x & 15) : threadIdx.x];
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int kp = geo.din / GT_KP;                       // 64-channel panels per tap
    const int taps = geo.nd == 3 ? geo.ks * geo.ks * geo.ks : geo.ks * geo.ks;
    const int panels = taps * kp;
    constexpr int CN = NH * GT_N;                         // columns per CTA
    constexpr int WST = NH == 2 ? 2 : GT_STAGES;          // W ring depth (NH = 2: 64 KB panels)
    constexpr size_t WSTAGE = (size_t)NH * 2 * GT_PANEL_BYTES;
    const int n_groups = Nn / CN;
    uint8_t *sA = smem;                                          // [stages][hi|lo][16 KB]: gathered x rows
    uint8_t *sWs = sA + (size_t)GT_STAGES * 2 * GT_PANEL_BYTES;  // [stages][hi|lo][16 KB]: W panel of this column half
    ConvSmem *cb = reinterpret_cast<ConvSmem *>(sWs + (size_t)WST * WSTAGE);
    // Per-panel tap table: (point offset of the neighbour, di, dj, dk, first channel).  ncu on the first version: 41 % of all
    // samples and 45 % of the instructions were the integer divisions that decode (tap, channel block) from the panel index --
    // by run-time divisors, per item, in every producer thread -- while the tensor pipe sat at 37 %.
    int4 *tapinfo = reinterpret_cast<int4 *>(reinterpret_cast<uint8_t *>(cb) + ((sizeof(ConvSmem) + 15) & ~size_t(15)));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int group = blockIdx.x % n_groups, worker = blockIdx.x / n_groups;
    for (int p = threadIdx.x; p < panels; p += CT_THREADS) {
        const int tap = p / kp, c0 = (p - tap * kp) * GT_KP, pad = geo.ks / 2;
        int di, dj, dk = 0;
        if (geo.nd == 3) { di = tap / (geo.ks * geo.ks) - pad; dj = (tap / geo.ks) % geo.ks - pad; dk = tap % geo.ks - pad; }
        else { di = tap / geo.ks - pad; dj = tap % geo.ks - pad; }
        tapinfo[p] = make_int4((di * geo.g1 + dj) * geo.g2 + dk, (di & 0xff) | ((dj & 0xff) << 8) | ((dk & 0xff) << 16), c0, 0);
    }
    const int64_t n_tiles = (M + GT_M - 1) / GT_M;
    const int64_t my_tiles = worker < n_tiles ? (n_tiles - worker + n_workers - 1) / n_workers : 0;

    if (warp == 8) {
        if (lane == 0) {
            for (int s = 0; s < GT_STAGES; ++s) {
                mbar_init(&cb->a_full[s], AIMG ? 1 : GT_PROD); mbar_init(&cb->a_empty[s], 1);
                mbar_init(&cb->w_full[s], 1); mbar_init(&cb->w_empty[s], 1);
            }
            for (int b = 0; b < 2; ++b) { mbar_init(&cb->t_full[b], 1); mbar_init(&cb->t_empty[b], 128); }
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc<2 * CN>(&cb->tmem_base);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = cb->tmem_base;

    if (AIMG && warp < 8) {
        // ---------------- A loader: one 32 KB [hi | lo] block of the operand image per stage ----------------
        if (warp == 0 && lane == 0) {
            const uint8_t *opi = reinterpret_cast<const uint8_t *>(X);
            int stage = 0, phase = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
                const int64_t tile = worker + it * n_workers;
                for (int p = 0; p < panels; ++p) {
                    mbar_wait(&cb->a_empty[stage], (uint32_t)(phase ^ 1));
                    mbar_arrive_expect_tx(&cb->a_full[stage], 2 * GT_PANEL_BYTES);
                    bulk_g2s(sA + (size_t)stage * 2 * GT_PANEL_BYTES, opi + ((size_t)tile * panels + p) * (2 * GT_PANEL_BYTES), 2 * GT_PANEL_BYTES,
                             &cb->a_full[stage]);
                    if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp < 8) {
        // ---------------- producers: implicit im2col.  Each thread owns 4 (row, chunk) items of the tile; the grid coordinates
        // of its rows are recomputed when the tile changes, the tap offset per panel. ----------------
        int64_t cur_ti = -1;
        int ci_[GT_LD / 2], cj_[GT_LD / 2], ck_[GT_LD / 2];
        int64_t base_[GT_LD / 2];   // (b * N + n) of the row, -1 past M
        auto src = [&](int64_t ti, int p, int i, int row, int c8) -> const float * {
            if (ti != cur_ti) {   // first item of a new tile: (this lambda is called with i = 0..3 in order)
                if (i == GT_LD / 2 - 1) cur_ti = ti;
                const int64_t m = (worker + ti * n_workers) * (int64_t)GT_M + row;
                if (m < M) {
                    const int64_t b = m / geo.N, n = m - b * geo.N;
                    const int64_t ij = n / geo.g2;
                    ck_[i] = (int)(n - ij * geo.g2);
                    ci_[i] = (int)(ij / geo.g1);
                    cj_[i] = (int)(ij - (int64_t)ci_[i] * geo.g1);
                    base_[i] = m;
                } else {
                    base_[i] = -1;
                }
            }
            if (base_[i] < 0) return nullptr;
            const int4 t = tapinfo[p];
            const int ii = ci_[i] + (int)(int8_t)(t.y & 0xff), jj = cj_[i] + (int)(int8_t)((t.y >> 8) & 0xff), kk = ck_[i] + (int)(int8_t)((t.y >> 16) & 0xff);
            if ((unsigned)ii >= (unsigned)geo.g0 || (unsigned)jj >= (unsigned)geo.g1 || (unsigned)kk >= (unsigned)geo.g2) return nullptr;   // zero padding
            return X + (base_[i] + t.x) * geo.din + t.z + c8 * 8;
        };
        gemm_produce_g<int>(panels, my_tiles, sA, src,
                       [&](int st, int parity) { mbar_wait(&cb->a_empty[st], (uint32_t)parity); },
                       [&](int st) { mbar_arrive(&cb->a_full[st]); });
    } else if (warp == 13) {
        // ---------------- W loader: one 32 KB panel [hi | lo] of this column half per stage ----------------
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (int64_t it = 0; it < my_tiles; ++it)
                for (int p = 0; p < panels; ++p) {
                    mbar_wait(&cb->w_empty[stage], (uint32_t)(phase ^ 1));
                    mbar_arrive_expect_tx(&cb->w_full[stage], (uint32_t)WSTAGE);
                    uint8_t *dst = sWs + (size_t)stage * WSTAGE;   // [hi: NH x 16 KB][lo: NH x 16 KB] = 128 NH rows, contiguous per half
#pragma unroll
                    for (int h = 0; h < NH; ++h) {
                        const uint8_t *src = Wimg + ((size_t)(group * NH + h) * panels + p) * (2 * GT_PANEL_BYTES);
                        bulk_g2s(dst + (size_t)h * GT_PANEL_BYTES, src, GT_PANEL_BYTES, &cb->w_full[stage]);
                        bulk_g2s(dst + (size_t)(NH + h) * GT_PANEL_BYTES, src + GT_PANEL_BYTES, GT_PANEL_BYTES, &cb->w_full[stage]);
                    }
                    if (++stage == WST) { stage = 0; phase ^= 1; }
                }
        }
        __syncwarp();
    } else if (warp == 8) {
        // ---------------- MMA issuer ----------------
        if (lane == 0 && my_tiles > 0) {
            const uint32_t idesc = instr_desc(FMT_BF16, GT_M, CN, false, false);
            int stage = 0, phase = 0, wst = 0, wph = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
                const int buf = (int)(it & 1);
                mbar_wait(&cb->t_empty[buf], (uint32_t)(((it >> 1) & 1) ^ 1));
                tc_fence_after();
                const uint32_t d = tmem_base + (uint32_t)buf * CN;
                for (int p = 0; p < panels; ++p) {
                    mbar_wait(&cb->a_full[stage], phase);
                    mbar_wait(&cb->w_full[wst], wph);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(sA + (size_t)stage * 2 * GT_PANEL_BYTES), a_lo = a_hi + GT_PANEL_BYTES;
                    const uint32_t b_hi = smem_u32(sWs + (size_t)wst * WSTAGE), b_lo = b_hi + NH * GT_PANEL_BYTES;
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
                    tc_commit(&cb->w_empty[wst]);
                    if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
                    if (++wst == WST) { wst = 0; wph ^= 1; }
                }
                tc_commit(&cb->t_full[buf]);
            }
        }
        __syncwarp();
    } else {
        // ---------------- epilogue: TMEM -> registers -> + bias -> xp [M, Nn] ----------------
        const int q = warp & 3;
        for (int64_t it = 0; it < my_tiles; ++it) {
            const int64_t tile = worker + it * n_workers;
            const int buf = (int)(it & 1);
            mbar_wait(&cb->t_full[buf], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            const int64_t m0 = tile * GT_M + q * 32 + (lane & ~3);
#pragma unroll 1
            for (int c0 = 0; c0 < CN; c0 += 32) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * CN + c0), r);
                tmem_wait_ld();
                if (AIMG) {
                    gemm_epilogue_chunk<EPI_F32, false, RES>(r, lane, m0, 0, 0, M, bias + group * CN + c0, nullptr, Y, Nn, group * CN + c0, epi);
                } else if (IMG) {
                    const int64_t bi0 = m0 / epi.N, n0 = m0 - bi0 * epi.N;
                    gemm_epilogue_chunk<EPI_IMG>(r, lane, m0, bi0, n0, M, bias + group * CN + c0, sWtc, nullptr, Nn, group * CN + c0, epi);
                } else {
                    gemm_epilogue_chunk<EPI_F32>(r, lane, m0, 0, 0, M, bias + group * CN + c0, nullptr, Y, Nn, group * CN + c0, epi);
                }
            }
            tc_fence_before();
            mbar_arrive(&cb->t_empty[buf]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc<2 * CN>(tmem_base);
    }
}

bool conv_tc_supported(const haet_layer_shape *s) {
    const int inner = s->heads * s->dim_head;
    const int taps = s->conv_nd == 3 ? s->kernel * s->kernel * s->kernel : s->kernel * s->kernel;
    return (s->conv_nd == 2 || s->conv_nd == 3) && s->dim % GT_KP == 0 && inner % GT_N == 0 && (s->kernel & 1) == 1 && s->kernel <= 7 &&
           taps * (s->dim / GT_KP) <= CT_MAX_PANELS && (int64_t)s->grid[0] * s->grid[1] * (s->conv_nd == 3 ? s->grid[2] : 1) < (1ll << 30);
}

int64_t conv_tc_image_bytes(const haet_layer_shape *s) {
    const int taps = s->conv_nd == 3 ? s->kernel * s->kernel * s->kernel : s->kernel * s->kernel;
    return (int64_t)s->heads * s->dim_head * taps * s->dim * 4;
}

// wx: [inner, taps, dim] (tap-major, as the SIMT path reads it); img: scratch of conv_tc_image_bytes(s)
// epi != nullptr: write the operand image / scales it names instead of xp (dim_head 16 or 32)
int launch_conv_tc(const haet_layer_shape *s, const float *x, const float *wx, const float *bx, float *xp, void *img,
                   cudaStream_t st, const GemmEpi *epi) {
    HAET_CHECK_ARG(conv_tc_supported(s), "conv_tc: unsupported shape");
    const int inner = s->heads * s->dim_head;
    ConvGeom geo;
    geo.N = s->N; geo.g0 = s->grid[0]; geo.g1 = s->grid[1]; geo.g2 = s->conv_nd == 3 ? s->grid[2] : 1;
    geo.ks = s->kernel; geo.nd = s->conv_nd; geo.din = s->dim;
    HAET_CHECK_ARG((int64_t)geo.g0 * geo.g1 * geo.g2 == s->N, "grid %dx%dx%d does not match N=%lld", geo.g0, geo.g1, geo.g2,
                   (long long)s->N);
    const int taps = geo.nd == 3 ? geo.ks * geo.ks * geo.ks : geo.ks * geo.ks;
    const int K = taps * s->dim;
    const int64_t M = s->B * s->N;
    if (M == 0) return HAET_OK;
    if (int rc = launch_gemm_tc_pack(wx, inner, K, (uint8_t *)img, st)) return rc;
    const int nh = (inner % (2 * GT_N) == 0 && !options().conv_nh1) ? 2 : 1;
    const size_t smem = 1024 + (size_t)GT_STAGES * 2 * GT_PANEL_BYTES + (size_t)(nh == 2 ? 2 : GT_STAGES) * nh * 2 * GT_PANEL_BYTES + sizeof(ConvSmem) + 16 +
                        (size_t)(K / GT_KP) * sizeof(int4);
    GemmEpi ep{};
    if (epi) ep = *epi;
    void (*kern)(const float *, const uint8_t *, const float *, float *, int64_t, int, int, ConvGeom, GemmEpi) =
        nh == 2 ? (epi ? conv_tc_kernel<2, true> : conv_tc_kernel<2, false>) : (epi ? conv_tc_kernel<1, true> : conv_tc_kernel<1, false>);
    if (nh == 2) { if (epi) HAET_DYN_SMEM((conv_tc_kernel<2, true>), smem); else HAET_DYN_SMEM((conv_tc_kernel<2, false>), smem); }
    else { if (epi) HAET_DYN_SMEM((conv_tc_kernel<1, true>), smem); else HAET_DYN_SMEM((conv_tc_kernel<1, false>), smem); }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int n_groups = inner / (nh * GT_N);
    const int64_t n_tiles = (M + GT_M - 1) / GT_M;
    int n_workers = sms / n_groups;
    if (n_workers < 1) n_workers = 1;
    if (n_workers > n_tiles) n_workers = (int)n_tiles;
    kern<<<n_workers * n_groups, CT_THREADS, smem, st>>>(x, (const uint8_t *)img, bx, xp, M, inner, n_workers, geo, ep);
    HAET_LAUNCHED();
    return HAET_OK;
}

int launch_conv_tc_img(const haet_layer_shape *s, const float *x, const float *wx, const float *bx, const float *wt, float bt,
                       uint8_t *xpi, float *scale, void *img, cudaStream_t st) {
    HAET_CHECK_ARG(s->dim_head == 32 || s->dim_head == 16, "conv_tc: the image epilogue needs dim_head 16 or 32");
    GemmEpi ep{};
    ep.c16 = s->dim_head == 16;
    ep.img = xpi; ep.scale = scale; ep.wt = wt; ep.bt = bt; ep.base_temp = s->base_temp;
    ep.N = s->N; ep.npad = xpi_npad(s->N); ep.H = s->heads;
    if (int rc = launch_xpi_pad(s, xpi, scale, st)) return rc;
    return launch_conv_tc(s, x, wx, bx, nullptr, img, st, &ep);
}

// out[M, 256] = A . W^T + bias (+ resid, may be NULL) with A the operand image `opi` ([tile][K / 64 panels][hi | lo]) and W [256, K] fp32:
// the block MLP's fc2 for hidden >= 512 in ONE pass (W streamed as 64 KB panels instead of one resident 256-column slice per launch)
bool gemm_stream_opi_supported(int K, int Nn) { return Nn == 2 * GT_N && K % GT_KP == 0 && K >= GT_KP && K / GT_KP <= CT_MAX_PANELS; }
int launch_gemm_stream_opi(const uint8_t *opi, const float *w, const float *bias, const float *resid, float *out, int64_t M, int K, int Nn,
                           void *img, cudaStream_t st) {
    HAET_CHECK_ARG(gemm_stream_opi_supported(K, Nn), "gemm_stream_opi: unsupported K=%d Nn=%d", K, Nn);
    HAET_CHECK_ARG(bias != nullptr, "gemm_stream_opi: a bias is required");
    if (M == 0) return HAET_OK;
    if (int rc = launch_gemm_tc_pack(w, Nn, K, (uint8_t *)img, st)) return rc;
    ConvGeom geo{};
    geo.N = M; geo.g0 = 1; geo.g1 = 1; geo.g2 = 1; geo.ks = 1; geo.nd = 2; geo.din = K;
    GemmEpi ep{};
    ep.resid = resid;
    const size_t smem = 1024 + (size_t)GT_STAGES * 2 * GT_PANEL_BYTES + (size_t)2 * 2 * 2 * GT_PANEL_BYTES + sizeof(ConvSmem) + 16 +
                        (size_t)(K / GT_KP) * sizeof(int4);
    HAET_DYN_SMEM((conv_tc_kernel<2, false, true, true>), smem);
    HAET_DYN_SMEM((conv_tc_kernel<2, false, true, false>), smem);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t n_tiles = (M + GT_M - 1) / GT_M;
    int n_workers = sms;
    if (n_workers > n_tiles) n_workers = (int)n_tiles;
    auto kern = resid ? conv_tc_kernel<2, false, true, true> : conv_tc_kernel<2, false, true, false>;
    kern<<<n_workers, CT_THREADS, smem, st>>>(reinterpret_cast<const float *>(opi), (const uint8_t *)img, bias, out, M, Nn, n_workers, geo, ep);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet
