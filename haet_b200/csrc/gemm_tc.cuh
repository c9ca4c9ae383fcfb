// Shared declarations of the tcgen05 GEMM kernels (gemm_tc.cu: one CTA per tile half; gemm2_tc.cu: CTA pairs).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc.cuh"

namespace haet {

constexpr int GT_M = 128;          // rows per CTA tile (UMMA M per CTA)
constexpr int GT_N = 128;          // W rows (output columns) held by one CTA
constexpr int GT_KP = 64;          // k per panel: 64 bf16 = one 128-byte swizzle row
constexpr int GT_STAGES = 3;
constexpr int GT_PANEL_BYTES = GT_M * 128;   // 16 KB: 128 rows x 128 B
constexpr int GT_PROD = 256;       // producer threads (warps 0-7)
constexpr int GT_THREADS = GT_PROD + 32 + 128;   // + MMA warp (8) + epilogue warps (9-12)
constexpr int GT_LD = GT_M * GT_KP / 4 / GT_PROD;   // float4 per producer thread per panel (8)
enum { EPI_F32 = 0, EPI_IMG = 1, EPI_OPI = 2 };   // EPI_OPI: GELU, then the NEXT GEMM's bf16 hi/lo operand image (block MLP)

struct GemmEpi {   // EPI_IMG arguments
    uint8_t *img;
    float *scale;
    const float *wt;
    float bt, base_temp;
    int64_t N, npad;   // rows m = b * N + n
    int H;
    int pf;    // producers prefetch the fp32 rows of the item `pf` (tile, panel) items ahead of their register loads into L2 (0: off)
    int dbg;   // development switches (HAET_GEMM_DBG): 1 = no epilogue stores, 2 = no global loads
    int c16;   // EPI_IMG: dim_head == 16 -- a 32-column chunk holds TWO heads; each head's image row keeps its 16 channels in
               // chunks 0, 1 (hi) and 4, 5 (lo), the other chunks are zero (the passes run their dim_head-32 kernels on it)
    int a_img; // X is the bf16 hi/lo operand image written by pass 2 (common.cuh: OPI), bulk-copied instead of converted
    // Block fusion (HAETransolver_Irregular_Mesh.py:125  fx = Attn(ln_1(fx)) + fx):
    //   EPI_F32: resid [M, ldy] is added to the output rows (nullptr: none)
    //   pair kernel, LN = true: X is the RAW residual stream; W / bias arrive folded (W.gamma, b + W.beta), the producers
    //   leave (mean_m, rstd_m) of every row in ln_stats [M] and the epilogue applies
    //   xp[m, o] = rstd_m (acc[m, o] - mean_m wsum[o]) + bias'[o],  wsum[o] = sum_k W'[o, k]
    const float *resid;
    const float *wsum;
    uint32_t wsum_s;  // shared-memory address of the CTA's copy of wsum (0: read ep.wsum from global memory); set by the kernel
    int smem_total;   // pair kernel, LN: dynamic shared memory of the launch; the kernel keeps wsum in its tail when the alignment pad leaves room
    float2 *ln_stats;
    float ln_eps;
    int ln_reuse;   // ln_stats was filled by an earlier launch over the same X (later column blocks of the block MLP's fc1)
    // block MLP (mlp.py:67-96): EPI_OPI writes gelu(y) as panels [opi_panel0, opi_panel0 + Nn / 64) of an operand image with
    // opi_panels panels per 128-row tile; an a_img GEMM reads panels [a_panel0, a_panel0 + K / 64) of an image with a_panels
    // (0: K / 64) panels per tile -- a hidden layer wider than 256 is produced / consumed 256 columns per launch
    uint8_t *opi;
    int opi_panels, opi_panel0;
    int a_panels, a_panel0;
};

struct GemmSmem {   // control block placed after the tiles
    uint64_t w_full;
    uint64_t a_full[GT_STAGES], a_empty[GT_STAGES];
    uint64_t t_full[2], t_empty[2];
    uint64_t a_land[GT_STAGES];   // pair kernel, operand-image A: this CTA's bulk copy of a stage has landed
    uint32_t tmem_base;
};


// Producer side shared by both GEMM kernels: this CTA's 128 rows of each of its tiles, k panel by k panel, fp32 from
// global -> bf16 hi/lo -> SWIZZLE_128B K-major operand stage, two panels ahead in registers.
// (Tried and dropped: a quarter-panel software pipeline with 8 loads in flight -- 2x slower, a warp has 6 scoreboard
// slots so independent loads cannot be retired independently; cp.async.bulk.prefetch.L2 of the tiles ahead -- ncu
// showed the x tiles then crossing HBM twice, 0.50 ms instead of 0.43 ms.)
//   row0(i)        first global row of this CTA's i-th tile
//   wait_empty(s, parity) / arrive_full(s): the stage hand-shake with the MMA issuer
//   srcf(ti, p, i, row, c8) address of the 8 consecutive fp32 (32 B) of tile ti, k panel p, tile row `row`, 16-byte bf16
//                  chunk c8 (i = which of the thread's 4 items), or nullptr for zeros (rows past M, conv padding)
struct NoPanelHook {
    __device__ __forceinline__ void operator()(const float4 (&)[8], int64_t) const {}
};
//   hook(v, it): sees the fp32 registers of item it = tile * panels + panel right before they are split and stored (LayerNorm statistics)
//   Idx: type of the item counter.  int64_t is what the tuned plain kernels were measured with; the LN variants use int
//   (a 64-bit division is a subroutine call and the extra live registers pushed them into 100-230 B of spills: 0.72 -> 0.60 ms)
template <class Idx = int64_t, class SrcFn, class WaitEmpty, class ArriveFull, class PanelHook = NoPanelHook>
__device__ __forceinline__ void gemm_produce_g(int panels, int64_t my_tiles, uint8_t *sA, SrcFn srcf, WaitEmpty wait_empty,
                                               ArriveFull arrive_full, PanelHook hook = PanelHook(), int pf = 0) {
    using namespace tc;
    const int tid = threadIdx.x;
    const Idx total = (Idx)(my_tiles * panels);   // (tile, panel) items of this CTA, in order
    auto load = [&](Idx it, float4 (&v)[GT_LD]) {
        if (it >= total) return;
        const Idx ti = it / (Idx)panels;
        const int p = (int)(it - ti * panels);
#pragma unroll
        for (int i = 0; i < GT_LD / 2; ++i) {   // 8 consecutive k (32 B) per item: one 16-B bf16 chunk of the row
            const int idx = i * GT_PROD + tid, row = idx >> 3, c8 = idx & 7;
            const float4 *src = reinterpret_cast<const float4 *>(srcf(ti, p, i, row, c8));
            if (src) {
                v[2 * i] = __ldg(src);
                v[2 * i + 1] = __ldg(src + 1);
            } else {
                v[2 * i] = v[2 * i + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        // The register pipeline holds two panels (64 KB per SM) in flight: at a loaded HBM latency of ~2.7 us that is 3.5 TB/s
        // of reads.  A short-distance L2 prefetch (one 128-byte line per 4 threads) turns the latency of the later register
        // loads into L2 latency; the distance must stay within a few microseconds -- at 6.5 TB/s the 126 MB L2 turns over in
        // ~20 us (round 1's bulk prefetch "tiles ahead" was evicted before use and read x from HBM twice).
        if (pf > 0 && it + (Idx)pf < total) {
            const Idx it2 = it + (Idx)pf, ti2 = it2 / (Idx)panels;
            const int p2 = (int)(it2 - ti2 * panels);
#pragma unroll
            for (int i = 0; i < GT_LD / 2; ++i) {
                const int idx = i * GT_PROD + tid, row = idx >> 3, c8 = idx & 7;
                if ((c8 & 3) == 0) {
                    const float *src = srcf(ti2, p2, i, row, c8);
                    if (src) asm volatile("prefetch.global.L2 [%0];" ::"l"(src));
                }
            }
        }
    };
    int stage = 0, phase = 0;
    auto store = [&](const float4 (&v)[GT_LD], Idx item) {
        hook(v, item);
        wait_empty(stage, phase ^ 1);
        uint8_t *hi_base = sA + (size_t)stage * 2 * GT_PANEL_BYTES;
#pragma unroll
        for (int i = 0; i < GT_LD / 2; ++i) {
            const int idx = i * GT_PROD + tid, row = idx >> 3, c8 = idx & 7;
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float4 f = v[2 * i + j];
                const __nv_bfloat162 ha = __floats2bfloat162_rn(f.x, f.y), hb = __floats2bfloat162_rn(f.z, f.w);
                const uint32_t ua = *reinterpret_cast<const uint32_t *>(&ha), ub = *reinterpret_cast<const uint32_t *>(&hb);
                const __nv_bfloat162 la = __floats2bfloat162_rn(f.x - __uint_as_float(ua << 16), f.y - __uint_as_float(ua & 0xffff0000u));
                const __nv_bfloat162 lb = __floats2bfloat162_rn(f.z - __uint_as_float(ub << 16), f.w - __uint_as_float(ub & 0xffff0000u));
                hi[2 * j] = ua; hi[2 * j + 1] = ub;
                lo[2 * j] = *reinterpret_cast<const uint32_t *>(&la); lo[2 * j + 1] = *reinterpret_cast<const uint32_t *>(&lb);
            }
            const uint32_t off = sw128_off(row, c8);
            *reinterpret_cast<uint4 *>(hi_base + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4 *>(hi_base + GT_PANEL_BYTES + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        fence_proxy_async();
        arrive_full(stage);
        if (++stage == GT_STAGES) { stage = 0; phase ^= 1; }
    };
    float4 va[GT_LD], vb[GT_LD];
    load(0, va);
    load(1, vb);
    for (Idx it = 0; it < total; it += 2) {
        store(va, it);
        load(it + 2, va);
        if (it + 1 < total) {
            store(vb, it + 1);
            load(it + 3, vb);
        }
    }
}

// dense A: row m of X [M, K]
template <class Idx = int64_t, class Row0, class WaitEmpty, class ArriveFull, class PanelHook = NoPanelHook>
__device__ __forceinline__ void gemm_produce(const float *__restrict__ X, int64_t M, int K, int panels, int64_t my_tiles,
                                             uint8_t *sA, Row0 row0, WaitEmpty wait_empty, ArriveFull arrive_full, int dbg = 0,
                                             PanelHook hook = PanelHook(), int pf = 0) {
    gemm_produce_g<Idx>(panels, my_tiles, sA,
                   [&](int64_t ti, int p, int, int row, int c8) -> const float * {
                       const int64_t m = row0(ti) + row;
                       return (m < M && !(dbg & 2)) ? X + m * K + p * GT_KP + c8 * 8 : nullptr;
                   },
                   wait_empty, arrive_full, hook, pf);
}

// LayerNorm statistics of the rows a producer thread touches (pair kernel, LN = true).  Thread tid holds, per panel, the 8
// floats (row i * 32 + tid / 8, chunk tid & 7) for i < 4: the 8 threads of a row reduce with 3 shuffles after the last
// panel and lane (tid & 7) == 0 writes (mean, rstd) -- nn.LayerNorm: biased variance, eps inside the root.
struct LnStatsHook {
    float s1[4], s2[4];
    // everything else comes from the kernel parameters (constant bank): no registers held across the load / store pipeline
    template <class Row0>
    __device__ __forceinline__ void run(const float4 (&v)[8], int64_t item, int panels, Row0 row0, const GemmEpi &ep, int64_t M, int K) {
        if ((ep.dbg & 16) || ep.ln_reuse) return;   // timing ablation / ln_stats already holds these rows' statistics
        const unsigned it32 = (unsigned)item, ti32 = it32 / (unsigned)panels;   // 32-bit: a 64-bit division is a subroutine call
        const int p = (int)(it32 - ti32 * (unsigned)panels);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (p == 0) { s1[i] = 0.f; s2[i] = 0.f; }
            const float4 a = v[2 * i], b = v[2 * i + 1];
            s1[i] += (a.x + a.y) + (a.z + a.w) + (b.x + b.y) + (b.z + b.w);
            s2[i] = fmaf(a.x, a.x, fmaf(a.y, a.y, fmaf(a.z, a.z, fmaf(a.w, a.w, s2[i]))));
            s2[i] = fmaf(b.x, b.x, fmaf(b.y, b.y, fmaf(b.z, b.z, fmaf(b.w, b.w, s2[i]))));
        }
        if (p != panels - 1) return;
        const int tid = threadIdx.x;
        const int64_t mrow = row0((int64_t)ti32) + (tid >> 3);
        const float invK = 1.f / (float)K;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float a = s1[i], b = s2[i];
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
            const int64_t m = mrow + i * 32;
            if ((tid & 7) == 0 && m < M) {
                const float mean = a * invK;
                const float var = fmaxf(fmaf(b, invK, -mean * mean), 0.f);
                ep.ln_stats[m] = make_float2(mean, rsqrtf(var + ep.ln_eps));
            }
        }
        __threadfence_block();   // ordered before this thread's arrive on the stage barrier (the epilogue reads after t_full)
    }
};

// Epilogue of one 32-row x 32-column accumulator chunk as tcgen05.ld.32x32b delivers it (r[j] = row `lane`, column j).
// Written row-per-lane, every store instruction would touch 32 different rows with 16 bytes each: measured, those
// stores bounded the GEMM at ~2 TB/s.  Two butterfly stages regroup the chunk so that 4 adjacent lanes hold 4 rows and
// write 64 contiguous bytes of one row per instruction (whole 32-byte sectors, 8 rows per instruction).
//   m0 / bi0 / n0: global row (and its batch / point index, EPI_IMG) of this lane's FIRST row = warp row + (lane & ~3)
//   bias32 / wt32: the chunk's 32 bias values / the Ada-Temp weights (EPI_IMG); col0: first global column of the chunk
//   LN (EPI_IMG only): rs[i] = (rstd, -mean * rstd) of row m0 + i; ep.wsum as documented in GemmEpi
//   RES (EPI_F32 only): ep.resid rows are added; their loads are issued BEFORE the transposes so that the memory latency
//   overlaps the shuffles (issued after them, the loads serialised the epilogue: to_out 0.34 -> 0.81 ms; now 0.50 ms)
template <int EPI, bool LN = false, bool RES = false>
__device__ __forceinline__ void gemm_epilogue_chunk(uint32_t (&r)[32], int lane, int64_t m0, int64_t bi0, int64_t n0, int64_t M,
                                                    const float *bias32, const float *wt32, float *Y, int ldy, int col0,
                                                    const GemmEpi &ep, const float2 *rs = nullptr) {
    const int l4 = lane & 3;
    // LN: the chunk's wsum values come from global memory (L1 / L2): issued before the transposes so that their latency
    // overlaps the shuffles (the same fix as the residual loads below)
    float su[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (LN && EPI != EPI_F32 && !(ep.dbg & 64)) {
        float4 sa, sb;
        if (ep.wsum_s) {   // the CTA's shared-memory copy (gemm2_tc.cu): LDS, not a generic load
            const uint32_t a = ep.wsum_s + (uint32_t)(col0 + 8 * l4) * 4u;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(sa.x), "=f"(sa.y), "=f"(sa.z), "=f"(sa.w) : "r"(a));
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(sb.x), "=f"(sb.y), "=f"(sb.z), "=f"(sb.w) : "r"(a + 16u));
        } else {
            sa = __ldg(reinterpret_cast<const float4 *>(ep.wsum + col0 + 8 * l4));
            sb = __ldg(reinterpret_cast<const float4 *>(ep.wsum + col0 + 8 * l4 + 4));
        }
        su[0] = sa.x; su[1] = sa.y; su[2] = sa.z; su[3] = sa.w; su[4] = sb.x; su[5] = sb.y; su[6] = sb.z; su[7] = sb.w;
    }
    if (EPI == EPI_F32) {
        float4 q0[4], q1[4];
        if (RES) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (m0 + i < M) {
                    const float *rp = ep.resid + (m0 + i) * ldy + col0 + 4 * l4;
                    q0[i] = __ldg(reinterpret_cast<const float4 *>(rp));
                    q1[i] = __ldg(reinterpret_cast<const float4 *>(rp + 16));
                } else {
                    q0[i] = q1[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
        }
        xpose_stage<0, 2>(r, lane);
        xpose_stage<1, 3>(r, lane);
        // register 4 i + 16 h + c = row m0 + i, column 16 h + 4 l4 + c
        const float4 b0 = *reinterpret_cast<const float4 *>(bias32 + 4 * l4), b1 = *reinterpret_cast<const float4 *>(bias32 + 16 + 4 * l4);
        if (ep.dbg & 1) return;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (m0 + i >= M) break;
            float *y = Y + (m0 + i) * ldy + col0 + 4 * l4;
            float4 o0 = make_float4(__uint_as_float(r[4 * i]) + b0.x, __uint_as_float(r[4 * i + 1]) + b0.y,
                                    __uint_as_float(r[4 * i + 2]) + b0.z, __uint_as_float(r[4 * i + 3]) + b0.w);
            float4 o1 = make_float4(__uint_as_float(r[4 * i + 16]) + b1.x, __uint_as_float(r[4 * i + 17]) + b1.y,
                                    __uint_as_float(r[4 * i + 18]) + b1.z, __uint_as_float(r[4 * i + 19]) + b1.w);
            if (RES) {   // + fx (the block's residual connection), read with the same full-sector pattern as the store
                o0.x += q0[i].x; o0.y += q0[i].y; o0.z += q0[i].z; o0.w += q0[i].w;
                o1.x += q1[i].x; o1.y += q1[i].y; o1.z += q1[i].z; o1.w += q1[i].w;
            }
            *reinterpret_cast<float4 *>(y) = o0;
            *reinterpret_cast<float4 *>(y + 16) = o1;
        }
    } else if (EPI == EPI_OPI) {
        xpose_stage<0, 3>(r, lane);
        xpose_stage<1, 4>(r, lane);
        // register 8 i + c = row m0 + i, column col0 + 8 l4 + c: one 16-byte bf16 chunk (hi) + one (lo) of the next GEMM's A row
        const float4 ba = *reinterpret_cast<const float4 *>(bias32 + 8 * l4), bb = *reinterpret_cast<const float4 *>(bias32 + 8 * l4 + 4);
        const float bs[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
        const int panel = ep.opi_panel0 + (col0 >> 6), cc = ((col0 & 63) >> 3) + l4;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float v[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const float a = __uint_as_float(r[8 * i + 2 * c + e]);
                    const float y = LN ? fmaf(a, rs[i].x, fmaf(rs[i].y, su[2 * c + e], bs[2 * c + e])) : a + bs[2 * c + e];
                    v[e] = 0.5f * y * (1.f + erff(y * 0.70710678118654752f));   // nn.GELU() (exact, mlp.py:8)
                }
                const __nv_bfloat162 hb = __floats2bfloat162_rn(v[0], v[1]);
                const uint32_t ub = *reinterpret_cast<const uint32_t *>(&hb);
                const __nv_bfloat162 lb = __floats2bfloat162_rn(v[0] - __uint_as_float(ub << 16), v[1] - __uint_as_float(ub & 0xffff0000u));
                hi[c] = ub;
                lo[c] = *reinterpret_cast<const uint32_t *>(&lb);
            }
            const int64_t m = m0 + i;
            if (m < M) {
                const int rr = (int)(m & 127);
                uint8_t *blk = ep.opi + (((m >> 7) * ep.opi_panels + panel) * 2) * (int64_t)GT_PANEL_BYTES + tc::sw128_off(rr, cc);
                *reinterpret_cast<uint4 *>(blk) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4 *>(blk + GT_PANEL_BYTES) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
        }
    } else {
        xpose_stage<0, 3>(r, lane);
        xpose_stage<1, 4>(r, lane);
        // register 8 i + c = row m0 + i, column 8 l4 + c: lane l4 owns 16-byte chunk l4 (hi) and 4 + l4 (lo) of the image row
        const float4 ba = *reinterpret_cast<const float4 *>(bias32 + 8 * l4), bb = *reinterpret_cast<const float4 *>(bias32 + 8 * l4 + 4);
        const int c16 = ep.c16, lc = c16 ? (l4 & 1) : l4;   // this lane's 16-byte chunk inside its head's row
        const float4 wa = *reinterpret_cast<const float4 *>(wt32 + 8 * lc), wb = *reinterpret_cast<const float4 *>(wt32 + 8 * lc + 4);
        const float bs[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
        const float ws[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
        const int h = c16 ? (col0 >> 4) + (l4 >> 1) : col0 >> 5;
        int64_t bi = bi0, n = n0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float d = 0.f;
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float v0, v1;
                if (LN) {
                    v0 = fmaf(__uint_as_float(r[8 * i + 2 * c]), rs[i].x, fmaf(rs[i].y, su[2 * c], bs[2 * c]));
                    v1 = fmaf(__uint_as_float(r[8 * i + 2 * c + 1]), rs[i].x, fmaf(rs[i].y, su[2 * c + 1], bs[2 * c + 1]));
                } else {
                    v0 = __uint_as_float(r[8 * i + 2 * c]) + bs[2 * c];
                    v1 = __uint_as_float(r[8 * i + 2 * c + 1]) + bs[2 * c + 1];
                }
                d = fmaf(v0, ws[2 * c], d);
                d = fmaf(v1, ws[2 * c + 1], d);
                // the image is fp16 hi + lo: the conversions saturate at the largest finite half instead of storing (inf, -inf) -> NaN
                hi[c] = tc::pack_h2_sat(v0, v1);
                const float2 hf = __half22float2(*reinterpret_cast<const __half2 *>(&hi[c]));
                lo[c] = tc::pack_h2_sat(v0 - hf.x, v1 - hf.y);
            }
            d += __shfl_xor_sync(0xffffffffu, d, 1);
            if (!c16) d += __shfl_xor_sync(0xffffffffu, d, 2);
            if (m0 + i < M && !(ep.dbg & 1)) {
                // one head: Ada-Temp, fp16 hi/lo image row (chunk j -> position j ^ (n & 7))
                const int64_t row = (bi * ep.H + h) * ep.npad + n;
                uint8_t *dst = ep.img + row * XPI_ROW_BYTES;
                const int sw = (int)(n & 7);
                *reinterpret_cast<uint4 *>(dst + ((lc ^ sw) << 4)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4 *>(dst + (((4 + lc) ^ sw) << 4)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                if (c16) {   // channels 16 .. 31 of the padded head do not exist
                    *reinterpret_cast<uint4 *>(dst + (((2 + lc) ^ sw) << 4)) = make_uint4(0, 0, 0, 0);
                    *reinterpret_cast<uint4 *>(dst + (((6 + lc) ^ sw) << 4)) = make_uint4(0, 0, 0, 0);
                }
                if (lc == 0) {
                    const float tau = ep.base_temp + fminf(fmaxf(d + ep.bt, -0.4f), 0.4f);   // Ada-Temp (IrregularMesh.py:168)
                    ep.scale[row] = kLog2e / tau;
                }
            }
            if (++n >= ep.N) { n = 0; ++bi; }
        }
    }
}

// CTA-pair variant (Nn == 256): every x tile is loaded and split once; the pair shares W through cta_group::2 MMAs
int launch_gemm_tc_pack(const float *w, int Nn, int K, uint8_t *img, cudaStream_t st, int ldw = 0);   // W [Nn, K] (row stride ldw) -> [half][panel][hi | lo] image
bool gemm2_tc_supported(int K, int Nn);
int launch_gemm2_tc(int epi, const float *x, const uint8_t *wimg, const float *bias, float *y, int64_t M, int K, int Nn,
                    const GemmEpi &ep, cudaStream_t st);

}  // namespace haet
