// CTA-pair (cta_group::2) variant of the split-precision tcgen05 GEMM of gemm_tc.cu for Nn == 256:
//   Y[M, 256] = X[M, K] . W[256, K]^T + bias      (in_project_x, IrregularMesh.py:157; to_out, :235)
//
// A cluster of two CTAs (one TPC) works on 256-row tiles: CTA r loads and splits ITS 128 rows of x (hi/lo bf16,
// SWIZZLE_128B K-major, as in gemm_tc.cu) and keeps W rows [128 r, 128 r + 128) resident (hi and lo, 128 KB).  One
// thread of the leader CTA issues tcgen05.mma.cta_group::2 (M = 256, N = 256): the hardware reads each CTA's A rows
// and B half from its own shared memory and leaves rows [128 r, ..) x all 256 columns in CTA r's TMEM.  Compared with
// the single-CTA kernel (each CTA = one column half of a 128-row tile) every x element is fetched and converted ONCE
// instead of twice, which is what that kernel was bound by (producer load latency; profiles/r1_ncu_summary.md).
//
// Cross-CTA protocol (all mbarriers live at the same offset in both CTAs):
//   a_full[stage]  (leader's, count 512): every producer thread of both CTAs arrives (remote mbarrier.arrive on the shared::cluster address)
//   a_empty[stage] (both, count 1)      : tcgen05.commit multicast after the stage's MMAs
//   t_full[buf]    (both, count 1)      : tcgen05.commit multicast after the tile's last MMA
//   t_empty[buf]   (leader's, count 512): every epilogue thread of both CTAs arrives
#include "gemm_tc.cuh"

namespace haet {
using namespace tc;

constexpr int G2_N = 256;
constexpr int G2_EPI = 256;                          // epilogue threads: 2 warpgroups, alternating 32-column chunks
constexpr int G2_THREADS = GT_PROD + 32 + G2_EPI;    // producers (warps 0-7), MMA warp (8), epilogue warps (9-16)
constexpr int G2_CB_BYTES = 128;                     // GemmSmem rounded up to 16 bytes; bias[256] and wt[32] follow

template <int EPI, bool LN, bool RES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G2_THREADS, 1)
gemm2_tc_kernel(const float *__restrict__ X, const uint8_t *__restrict__ Wimg, const float *__restrict__ bias,
                float *__restrict__ Y, int64_t M, int K, int n_pairs, GemmEpi ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int panels = K / GT_KP;
    uint8_t *sW = smem;                                         // [panels][hi|lo][16 KB]: W rows of this CTA's half
    uint8_t *sA = sW + (size_t)panels * 2 * GT_PANEL_BYTES;     // [stages][hi|lo][16 KB]: this CTA's 128 x rows
    GemmSmem *cb = reinterpret_cast<GemmSmem *>(sA + (size_t)GT_STAGES * 2 * GT_PANEL_BYTES);
    static_assert(sizeof(GemmSmem) <= G2_CB_BYTES, "control block");
    float *sBias = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(cb) + G2_CB_BYTES);   // [256]
    float *sWt = sBias + G2_N;                                                                 // [32] (EPI_IMG)
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1;
    const int64_t n_ptiles = (M + 2 * GT_M - 1) / (2 * GT_M);   // 256-row pair tiles
    const int64_t my_tiles = pair < n_ptiles ? (n_ptiles - pair + n_pairs - 1) / n_pairs : 0;

    if (warp == 8) {
        if (lane == 0) {
            mbar_init(&cb->w_full, 1);
            for (int s = 0; s < GT_STAGES; ++s) {
                mbar_init(&cb->a_full[s], ep.a_img ? 2 : 2 * GT_PROD);
                mbar_init(&cb->a_empty[s], 1);
                mbar_init(&cb->a_land[s], 1);
            }
            for (int b = 0; b < 2; ++b) { mbar_init(&cb->t_full[b], 1); mbar_init(&cb->t_empty[b], 2 * G2_EPI); }
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc2<512>(&cb->tmem_base);
        // this CTA's half of W (rows [128 rank, +128), hi and lo panels): resident for the whole kernel
        const uint32_t wbytes = (uint32_t)panels * 2 * GT_PANEL_BYTES;
        if (lane == 0) {
            mbar_arrive_expect_tx(&cb->w_full, wbytes);
            const uint8_t *src = Wimg + (size_t)rank * wbytes;
            for (int c = 0; c < panels * 2; ++c)
                bulk_g2s(sW + (size_t)c * GT_PANEL_BYTES, src + (size_t)c * GT_PANEL_BYTES, GT_PANEL_BYTES, &cb->w_full);
        }
        mbar_wait(&cb->w_full, 0);
    }
    for (int i = threadIdx.x; i < G2_N; i += G2_THREADS) sBias[i] = bias ? bias[i] : 0.f;
    if (EPI == EPI_IMG && threadIdx.x < 32) sWt[threadIdx.x] = ep.wt[ep.c16 ? (threadIdx.x & 15) : threadIdx.x];
    // LN: the epilogue re-reads wsum[256] for every tile; from global memory those loads sat behind the streaming x rows in L1 / L2
    // and their latency was 15 % of the kernel's stall samples.  The layout is 256 bytes short of holding a copy -- unless the
    // 1 KB reserved for aligning the dynamic segment is not needed (its base is 1 KB aligned on this architecture): checked here,
    // per launch, and used when it fits.
    GemmEpi epl = ep;
    if (LN) {
        float *sWsum = sWt + 32;
        const uint32_t used = (uint32_t)(reinterpret_cast<uint8_t *>(sWsum + G2_N) - smem_raw);
        if (used <= (uint32_t)ep.smem_total) {
            for (int i = threadIdx.x; i < G2_N; i += G2_THREADS) sWsum[i] = ep.wsum[i];
            epl.wsum_s = smem_u32(sWsum);
        }
    }
    // barriers initialised, TMEM allocated and W resident in BOTH CTAs before any remote arrive / pair MMA
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = cb->tmem_base;

    if (warp < 8) {
        // ---------------- producers: this CTA's 128 rows of the pair tile -> bf16 hi/lo, SWIZZLE_128B K-major ----------------
        const uint32_t full0 = mapa_u32(&cb->a_full[0], 0);   // the LEADER's a_full barriers (8 bytes apart)
        if (ep.a_img) {
            // A arrives as the ready-made operand image (pass 2 wrote [tile128][panel][hi | lo][128 rows x 128 B, swizzled]):
            // warp 0 bulk-copies one 32 KB stage per (tile, panel), warp 1 relays "landed here" to the leader's a_full.
            const uint8_t *Aimg = reinterpret_cast<const uint8_t *>(X);
            const int64_t total = my_tiles * panels;
            const int a_panels = ep.a_panels ? ep.a_panels : panels;   // panels per tile of the image / first one this GEMM reads
            int stage = 0, phase = 0;
            if (warp == 0) {
                if (lane == 0) {
                    for (int64_t it = 0; it < total; ++it) {
                        const int64_t ti = it / panels;
                        const int p = (int)(it - ti * panels);
                        const int64_t t128 = (pair + ti * n_pairs) * 2 + rank;
                        mbar_wait(&cb->a_empty[stage], (uint32_t)(phase ^ 1));
                        mbar_arrive_expect_tx(&cb->a_land[stage], 2 * GT_PANEL_BYTES);
                        bulk_g2s(sA + (size_t)stage * 2 * GT_PANEL_BYTES, Aimg + (size_t)(t128 * a_panels + ep.a_panel0 + p) * 2 * GT_PANEL_BYTES,
                                 2 * GT_PANEL_BYTES, &cb->a_land[stage]);
                        if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
                    }
                }
                __syncwarp();
            } else if (warp == 1) {
                for (int64_t it = 0; it < total; ++it) {
                    mbar_wait(&cb->a_land[stage], (uint32_t)phase);
                    fence_proxy_async();
                    if (lane == 0) mbar_arrive_remote(full0 + 8u * stage);
                    __syncwarp();
                    if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
                }
            }
            // (RES: pulling the next tile's residual rows into L2 from the idle warps -- bulk or per-line prefetch -- made them
            // cross HBM 1.75 times, ncu dram__bytes_read 2.78 GB against 2.05 GB, and was slower than no prefetch: 0.58 against
            // 0.50 ms.  What fixed the naive 0.81 ms was issuing the epilogue's residual loads before its transposes.)
        } else if (LN) {
            // fused ln_1: the producers see every element of their rows anyway; they leave (mean, rstd) per row for the epilogue
            auto row0 = [&](int64_t ti) { return (pair + ti * n_pairs) * (int64_t)(2 * GT_M) + rank * GT_M; };
            LnStatsHook ln;
            gemm_produce<int>(X, M, K, panels, my_tiles, sA, row0,
                         [&](int st, int parity) { mbar_wait(&cb->a_empty[st], (uint32_t)parity); },
                         [&](int st) { mbar_arrive_remote(full0 + 8u * st); }, ep.dbg,
                         [&](const float4 (&v)[GT_LD], int item) { ln.run(v, item, panels, row0, ep, M, K); }, ep.pf);
        } else
        gemm_produce(X, M, K, panels, my_tiles, sA,
                     [&](int64_t ti) { return (pair + ti * n_pairs) * (int64_t)(2 * GT_M) + rank * GT_M; },
                     [&](int st, int parity) { mbar_wait(&cb->a_empty[st], (uint32_t)parity); },
                     [&](int st) { mbar_arrive_remote(full0 + 8u * st); }, ep.dbg, NoPanelHook(), ep.pf);
    } else if (warp == 8) {
        // ---------------- MMA issuer: the leader CTA only, whole warp in the loop, one elected lane issues ----------------
        if (rank == 0) {
            const uint32_t idesc = instr_desc(FMT_BF16, 2 * GT_M, G2_N, false, false);
            const uint32_t a_base = smem_u32(sA) >> 4, b_base = smem_u32(sW) >> 4;
            int stage = 0, phase = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
                const int buf = (int)(it & 1);
                mbar_wait(&cb->t_empty[buf], (uint32_t)(((it >> 1) & 1) ^ 1));
                tc_fence_after();
                const uint32_t d = tmem_base + (uint32_t)buf * G2_N;
                for (int p = 0; p < panels; ++p) {
                    mbar_wait(&cb->a_full[stage], (uint32_t)phase);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t a_hi = a_base + (uint32_t)stage * (2 * GT_PANEL_BYTES >> 4), a_lo = a_hi + (GT_PANEL_BYTES >> 4);
                        const uint32_t b_hi = b_base + (uint32_t)p * (2 * GT_PANEL_BYTES >> 4), b_lo = b_hi + (GT_PANEL_BYTES >> 4);
#pragma unroll
                        for (int ks = 0; ks < GT_KP / 16; ++ks) {
                            const uint64_t dah = desc_sw128_u(a_hi + ks * 2, 16, 1024), dal = desc_sw128_u(a_lo + ks * 2, 16, 1024);
                            const uint64_t dbh = desc_sw128_u(b_hi + ks * 2, 16, 1024), dbl = desc_sw128_u(b_lo + ks * 2, 16, 1024);
                            mma2_f16_ss(d, dah, dbh, idesc, (p | ks) != 0);
                            mma2_f16_ss(d, dal, dbh, idesc, true);
                            mma2_f16_ss(d, dah, dbl, idesc, true);
                        }
                        tc_commit2(&cb->a_empty[stage], 3);
                        if (p == panels - 1) tc_commit2(&cb->t_full[buf], 3);
                    }
                    __syncwarp();
                    if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else {
        // ---------------- epilogue (both CTAs): own 128 rows x 256 columns, TMEM -> registers -> + bias -> global ----------------
        const int q = warp & 3;          // TMEM lane quarter this warp may access
        const int eg = (warp - 9) >> 2;  // epilogue warpgroup: takes the 32-column chunks c0 = 32 (2 i + eg)
        const uint32_t empty0 = mapa_u32(&cb->t_empty[0], 0);   // the LEADER's t_empty barriers (8 bytes apart)
        for (int64_t it = 0; it < my_tiles; ++it) {
            const int64_t ptile = pair + it * n_pairs;
            const int buf = (int)(it & 1);
            mbar_wait(&cb->t_full[buf], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            const int64_t m0 = ptile * (2 * GT_M) + rank * GT_M + q * 32 + (lane & ~3);   // first of this lane's 4 rows
            int64_t bi0 = 0, n0 = 0;
            if (EPI == EPI_IMG) { bi0 = m0 / ep.N; n0 = m0 - bi0 * ep.N; }
            float2 rs[4];   // LN: (rstd, -mean rstd) of this lane's 4 rows, written by this CTA's producers (L2 read: not L1-cached)
            if (LN) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 mr = (m0 + i < M && !(ep.dbg & 32)) ? __ldcg(ep.ln_stats + m0 + i) : make_float2(0.f, 1.f);
                    rs[i] = make_float2(mr.y, -mr.x * mr.y);
                }
            }
#pragma unroll 1
            for (int c0 = eg * 32; c0 < G2_N; c0 += 64) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * G2_N + c0), r);
                tmem_wait_ld();
                gemm_epilogue_chunk<EPI, LN, RES>(r, lane, m0, bi0, n0, M, sBias + c0, sWt, Y, G2_N, c0, epl, rs);
            }
            tc_fence_before();
            mbar_arrive_remote(empty0 + 8u * buf);
        }
    }
    // neither CTA may leave (or free its TMEM) while the pair's MMAs / remote arrives can still touch it
    tc_fence_before();
    cluster_sync_all();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc2<512>(tmem_base);
    }
}

bool gemm2_tc_supported(int K, int Nn) { return Nn == G2_N && K % GT_KP == 0 && K >= GT_KP && K <= 256; }

int launch_gemm2_tc(int epi, const float *x, const uint8_t *wimg, const float *bias, float *y, int64_t M, int K, int Nn,
                    const GemmEpi &ep, cudaStream_t st) {
    HAET_CHECK_ARG(gemm2_tc_supported(K, Nn), "gemm2_tc: unsupported K=%d Nn=%d", K, Nn);
    const int panels = K / GT_KP;
    const size_t smem_base = 1024 + (size_t)(panels + GT_STAGES) * 2 * GT_PANEL_BYTES + G2_CB_BYTES + sizeof(float) * (G2_N + 32);
    // LN variants ask for up to 1 KB more (their shared-memory copy of wsum), capped at the 227 KB opt-in limit
    const size_t smem_ln = smem_base + sizeof(float) * G2_N <= 232448 ? smem_base + sizeof(float) * G2_N : 232448;
    const size_t smem = (epi == EPI_OPI || ep.wsum != nullptr) ? smem_ln : smem_base;
    if (epi == EPI_OPI) {   // block MLP fc1: ln_2 + GELU + operand image
        HAET_CHECK_ARG(ep.ln_stats != nullptr && ep.wsum != nullptr && ep.opi != nullptr, "EPI_OPI needs the LayerNorm scratch and an image");
        HAET_DYN_SMEM((gemm2_tc_kernel<EPI_OPI, true, false>), smem);
        int dev0 = 0, sms0 = 148;
        cudaGetDevice(&dev0);
        cudaDeviceGetAttribute(&sms0, cudaDevAttrMultiProcessorCount, dev0);
        const int64_t tiles = (M + 2 * GT_M - 1) / (2 * GT_M);
        int np = sms0 / 2;
        if (np > tiles) np = (int)tiles;
        if (np < 1) np = 1;
        GemmEpi epo = ep;
        epo.pf = options().gemm_pf;
        epo.smem_total = (int)smem;
        gemm2_tc_kernel<EPI_OPI, true, false><<<2 * np, G2_THREADS, smem, st>>>(x, wimg, bias, y, M, K, np, epo);
        HAET_LAUNCHED();
        return HAET_OK;
    }
    const bool ln = epi == EPI_IMG && ep.ln_stats != nullptr;
    const bool res = epi == EPI_F32 && ep.resid != nullptr;
    if (ln)
        HAET_DYN_SMEM((gemm2_tc_kernel<EPI_IMG, true, false>), smem);
    else if (res)
        HAET_DYN_SMEM((gemm2_tc_kernel<EPI_F32, false, true>), smem);
    else if (epi == EPI_IMG)
        HAET_DYN_SMEM((gemm2_tc_kernel<EPI_IMG, false, false>), smem);
    else
        HAET_DYN_SMEM((gemm2_tc_kernel<EPI_F32, false, false>), smem);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t n_ptiles = (M + 2 * GT_M - 1) / (2 * GT_M);
    int n_pairs = sms / 2;
    if (n_pairs > n_ptiles) n_pairs = (int)n_ptiles;
    if (n_pairs < 1) n_pairs = 1;
    GemmEpi epd = ep;
    epd.pf = options().gemm_pf;
    epd.smem_total = (int)smem;
#ifdef HAET_DEV_HOOKS   // timing ablations (results are garbage): never compiled into the shipped library
    if (const char *e = getenv("HAET_GEMM_DBG")) epd.dbg |= atoi(e);
#endif
    if (ln) {
        HAET_CHECK_ARG(ep.wsum != nullptr, "fused LayerNorm needs wsum");
        gemm2_tc_kernel<EPI_IMG, true, false><<<2 * n_pairs, G2_THREADS, smem, st>>>(x, wimg, bias, y, M, K, n_pairs, epd);
    } else if (res)
        gemm2_tc_kernel<EPI_F32, false, true><<<2 * n_pairs, G2_THREADS, smem, st>>>(x, wimg, bias, y, M, K, n_pairs, epd);
    else if (epi == EPI_IMG)
        gemm2_tc_kernel<EPI_IMG, false, false><<<2 * n_pairs, G2_THREADS, smem, st>>>(x, wimg, bias, y, M, K, n_pairs, epd);
    else
        gemm2_tc_kernel<EPI_F32, false, false><<<2 * n_pairs, G2_THREADS, smem, st>>>(x, wimg, bias, y, M, K, n_pairs, epd);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet
