// Pass 1 on tcgen05: per (b, h, chunk of points) the running column softmax state over the points
//   m2[g] ~ max_n z2[n,g],  l[g] = sum_n 2^(z2-m2),  S[g,:] = sum_n 2^(z2-m2) x_proj[n,:]
// with z2 = ((x_proj . Ws_g + bs_g - log(-log eps)) / tau_n) log2 e     (IrregularMesh.py:168-183).
//
// Flash-attention-shaped, with the SLICES on the TMEM lanes so that the softmax over the points
// is a per-thread reduction:
//   MMA1 (kind::f16, 3-term bf16 split = fp32-class logits; precision_study.py):
//        D[g, n] = Ws[g,:] . x_proj[n,:]          M = 128 slices, N = 64 points, K' = 96
//   softmax warps (thread = slice row g): t = (D + bc_g) * s_n - m_run, P = 2^t, l += P, P -> TMEM in place.
//        m_run is a LAZY running max: it is only raised when a tile exceeds it by 2^32 (P stays finite in fp32 /
//        tf32), so the TMEM accumulator rescale is off the common path.
//   MMA2 (kind::tf32, A = P from TMEM, B = x_proj tile as MN-major fp32):
//        S[g, c] += sum_n P[g, n] x_proj[n, c]     M = 128, N = 32, K = 64
// Warp roles: 2 softmax warpgroups (slice blocks j = w, w+2, ..), 4 producer warps (x_proj tile -> bf16 hi/lo
// K-major rows, tf32-rounded fp32 rows, s_n = log2e / tau_n), 1 MMA-issuing warp.  All hand-offs are mbarriers.
#include <stdlib.h>

#include "common.cuh"
#include "tc.cuh"

namespace haet {
using namespace tc;

constexpr int ST_C = 32;            // dim_head handled by this kernel
constexpr int ST_TILE = 64;         // points per tile
constexpr int ST_STAGES = 4;
constexpr int ST_THREADS = 416;     // 8 softmax warps + 4 producer warps + 1 MMA warp
constexpr int ST_GB = 128;          // slices per block (UMMA M)
constexpr int ST_MAX_GB = 4;        // up to 512 slices
constexpr int ST_TILE_BYTES = ST_TILE * 128;   // one operand tile: 64 rows x 128 B
constexpr float ST_RAISE = 32.0f;   // lazy-max threshold (log2 units)
constexpr float ST_TF32_BIAS = 1.0f + 1.0f / 4096.0f;   // centres the tensor core's tf32 truncation of P

struct SliceCtl {
    uint64_t b_full[ST_STAGES], b_empty[ST_STAGES];
    uint64_t d_full[2][2], p_full[2][2];
    uint64_t mma2_done[2];
    uint64_t all_done;
    uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t st_dbuf_col(int wg, int buf) { return 128u + (uint32_t)(wg * 2 + buf) * ST_TILE; }

__global__ void __launch_bounds__(ST_THREADS, 1)
slice_tc_kernel(const float *__restrict__ xp, const float *__restrict__ ws, const float *__restrict__ bs,
                const float *__restrict__ wt, float bt, float base_temp, float shift, int64_t N, int H, int G,
                int64_t chunk_pts, float raise_thr, float *__restrict__ parts) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int n_gb = (G + ST_GB - 1) / ST_GB;
    uint8_t *sW = smem;                                              // [n_gb][2][16 KB]: [hi|hi] rows, [lo|0] rows
    uint8_t *sB = sW + (size_t)n_gb * 2 * (ST_GB * 128);             // [stages][B1 8 KB | B2 8 KB]
    float *sS = reinterpret_cast<float *>(sB + (size_t)ST_STAGES * 2 * ST_TILE_BYTES);   // [stages][64] s_n
    SliceCtl *ctl = reinterpret_cast<SliceCtl *>(sS + ST_STAGES * ST_TILE);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bh = blockIdx.y, b = bh / H, h = bh - b * H;
    const int inner = H * ST_C;
    const int64_t n_begin = (int64_t)blockIdx.x * chunk_pts;
    const int64_t n_end = min(N, n_begin + chunk_pts);
    const int n_tiles = (int)((n_end - n_begin + ST_TILE - 1) / ST_TILE);
    const int64_t BH = gridDim.y;

    // ---- one-time setup: barriers, TMEM, the Ws operand images ----
    if (warp == 12) {
        if (lane == 0) {
            for (int s = 0; s < ST_STAGES; ++s) { mbar_init(&ctl->b_full[s], 128); mbar_init(&ctl->b_empty[s], 1); }
            for (int w = 0; w < 2; ++w) {
                for (int k = 0; k < 2; ++k) { mbar_init(&ctl->d_full[w][k], 1); mbar_init(&ctl->p_full[w][k], 128); }
                mbar_init(&ctl->mma2_done[w], 1);
            }
            mbar_init(&ctl->all_done, 1);
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc<512>(&ctl->tmem_base);
    }
    for (int idx = threadIdx.x; idx < n_gb * ST_GB * 4; idx += ST_THREADS) {   // one 16-B chunk (8 k) per step
        const int row = idx >> 2, ck = idx & 3, g = row, j = row / ST_GB, r = row % ST_GB;
        __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float v = g < G ? ws[(size_t)g * ST_C + ck * 8 + i] : 0.f;
            split_bf16(v, hi[i], lo[i]);
        }
        uint8_t *w1 = sW + (size_t)j * 2 * (ST_GB * 128), *w2 = w1 + ST_GB * 128;
        *reinterpret_cast<uint4 *>(w1 + sw128_off(r, ck)) = *reinterpret_cast<const uint4 *>(hi);
        *reinterpret_cast<uint4 *>(w1 + sw128_off(r, ck + 4)) = *reinterpret_cast<const uint4 *>(hi);
        *reinterpret_cast<uint4 *>(w2 + sw128_off(r, ck)) = *reinterpret_cast<const uint4 *>(lo);
        *reinterpret_cast<uint4 *>(w2 + sw128_off(r, ck + 4)) = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp >= 8 && warp < 12) {
        // =========================== producers ===========================
        const int tid = threadIdx.x - 256, r = tid >> 1, hf = tid & 1;
        float wtr[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) wtr[i] = wt[hf * 16 + i];
        const float *src = xp + ((int64_t)b * N) * inner + h * ST_C + hf * 16;
        float4 cur[4], nxt[4];
        auto load = [&](int t, float4 (&v)[4]) {
            const int64_t n = n_begin + (int64_t)t * ST_TILE + r;
            if (t < n_tiles && n < n_end) {
                const float4 *p = reinterpret_cast<const float4 *>(src + n * inner);
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = __ldg(p + i);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
        load(0, cur);
        int stage = 0, phase = 0;
        for (int t = 0; t < n_tiles; ++t) {
            load(t + 1, nxt);
            mbar_wait(&ctl->b_empty[stage], phase ^ 1);
            const bool valid = n_begin + (int64_t)t * ST_TILE + r < n_end;
            const float *f = reinterpret_cast<const float *>(cur);
            float d = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) d = fmaf(f[i], wtr[i], d);
            d += __shfl_xor_sync(0xffffffffu, d, 1);
            if (hf == 0) {
                const float tau = base_temp + fminf(fmaxf(d + bt, -0.4f), 0.4f);   // Ada-Temp (:168)
                sS[stage * ST_TILE + r] = valid ? kLog2e / tau : 0.f;
            }
            uint8_t *b1 = sB + (size_t)stage * 2 * ST_TILE_BYTES, *b2 = b1 + ST_TILE_BYTES;
            __align__(16) __nv_bfloat16 hi[16], lo[16];
            __align__(16) float tf[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                split_bf16(f[i], hi[i], lo[i]);
                uint32_t u;
                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(f[i]));
                tf[i] = __uint_as_float(u);
            }
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                *reinterpret_cast<uint4 *>(b1 + sw128_off(r, hf * 2 + c)) = reinterpret_cast<const uint4 *>(hi)[c];
                *reinterpret_cast<uint4 *>(b1 + sw128_off(r, 4 + hf * 2 + c)) = reinterpret_cast<const uint4 *>(lo)[c];
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)   // MN-major tf32 tile: SWIZZLE_128B_BASE32B
                *reinterpret_cast<uint4 *>(b2 + mn32_off(r, hf * 2 + (c >> 1)) + (c & 1) * 16) = reinterpret_cast<const uint4 *>(tf)[c];
            fence_proxy_async();
            mbar_arrive(&ctl->b_full[stage]);
            if (++stage == ST_STAGES) { stage = 0; phase ^= 1; }
#pragma unroll
            for (int i = 0; i < 4; ++i) cur[i] = nxt[i];
        }
    } else if (warp == 12) {
        // =========================== MMA issuer ===========================
        if (lane == 0) {
            const uint32_t id1 = instr_desc(FMT_BF16, ST_GB, ST_TILE, false, false);
            const uint32_t id2 = instr_desc(FMT_TF32, ST_GB, ST_C, false, true);
            const int Q = n_tiles * n_gb;
            int cnt1[2] = {0, 0}, cnt2[2] = {0, 0};
            for (int q = 0; q < Q + 2; ++q) {
                if (q >= 2) {   // ---- MMA2 of item q-2: S_j += P . x_proj
                    const int qq = q - 2, t = qq / n_gb, j = qq - t * n_gb, w = j & 1, k = cnt2[w]++;
                    const int buf = k & 1, stage = t % ST_STAGES;
                    mbar_wait(&ctl->p_full[w][buf], (uint32_t)((k >> 1) & 1));
                    tc_fence_after();
                    const uint32_t b2 = smem_u32(sB + (size_t)stage * 2 * ST_TILE_BYTES + ST_TILE_BYTES);
                    const uint32_t pa = tmem_base + st_dbuf_col(w, buf), sd = tmem_base + (uint32_t)j * ST_C;
#pragma unroll
                    for (int ks = 0; ks < ST_TILE / 8; ++ks)
                        mma_tf32_ts(sd, pa + ks * 8, smem_desc_mn32(b2 + ks * 1024, 512, 512), id2, (t | ks) != 0);
                    tc_commit(&ctl->mma2_done[w]);
                    if (j == n_gb - 1) tc_commit(&ctl->b_empty[stage]);
                }
                if (q < Q) {    // ---- MMA1 of item q: D = Ws_j . x_proj^T (hi.hi + hi.lo + lo.hi)
                    const int t = q / n_gb, j = q - t * n_gb, w = j & 1, k = cnt1[w]++;
                    const int buf = k & 1, stage = t % ST_STAGES;
                    if (j == 0) { mbar_wait(&ctl->b_full[stage], (uint32_t)((t / ST_STAGES) & 1)); tc_fence_after(); }
                    const uint32_t b1 = smem_u32(sB + (size_t)stage * 2 * ST_TILE_BYTES);
                    const uint32_t a1 = smem_u32(sW + (size_t)j * 2 * (ST_GB * 128)), a2 = a1 + ST_GB * 128;
                    const uint32_t d = tmem_base + st_dbuf_col(w, buf);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        mma_bf16_ss(d, smem_desc_sw128(a1 + ks * 32, 16, 1024), smem_desc_sw128(b1 + ks * 32, 16, 1024), id1, ks != 0);
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks)
                        mma_bf16_ss(d, smem_desc_sw128(a2 + ks * 32, 16, 1024), smem_desc_sw128(b1 + ks * 32, 16, 1024), id1, true);
                    tc_commit(&ctl->d_full[w][buf]);
                }
            }
            tc_commit(&ctl->all_done);   // every MMA of this CTA has completed when this phase flips
        }
        __syncwarp();
    } else {
        // =========================== softmax warpgroups ===========================
        const int w = warp >> 2, q4 = warp & 3, row = q4 * 32 + lane;   // TMEM lane = slice row within its block
        const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
        float m_run[2], l_run[2], bc[2];
        int n_my = 0;
        for (int j = w; j < n_gb; j += 2, ++n_my) {
            const int g = j * ST_GB + row;
            m_run[n_my] = 0.f; l_run[n_my] = 0.f;
            bc[n_my] = g < G ? bs[g] - shift : 0.f;
        }
        int k = 0;   // item counter of this warpgroup
        for (int t = 0; t < n_tiles && n_my > 0; ++t) {
            const int stage = t % ST_STAGES;
            mbar_wait(&ctl->b_full[stage], (uint32_t)((t / ST_STAGES) & 1));
            const float4 *s4 = reinterpret_cast<const float4 *>(sS + stage * ST_TILE);
#pragma unroll 1
            for (int jj = 0; jj < n_my; ++jj, ++k) {
                const int j = w + 2 * jj, buf = k & 1;
                mbar_wait(&ctl->d_full[w][buf], (uint32_t)((k >> 1) & 1));
                tc_fence_after();
                const uint32_t dcol = tmem_base + lane_base + st_dbuf_col(w, buf);
                uint32_t r0[32], r1[32];
                tmem_ld32(dcol, r0);
                tmem_ld32(dcol + 32, r1);
                tmem_wait_ld();
                const float mr = m_run[jj], bcj = bc[jj];
                float tmax = -INFINITY;
#pragma unroll
                for (int c = 0; c < 32; c += 4) {
                    const float4 sa = s4[c >> 2], sb = s4[8 + (c >> 2)];
                    float v;
                    v = fmaf(__uint_as_float(r0[c]) + bcj, sa.x, -mr); r0[c] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                    v = fmaf(__uint_as_float(r0[c + 1]) + bcj, sa.y, -mr); r0[c + 1] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                    v = fmaf(__uint_as_float(r0[c + 2]) + bcj, sa.z, -mr); r0[c + 2] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                    v = fmaf(__uint_as_float(r0[c + 3]) + bcj, sa.w, -mr); r0[c + 3] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                    v = fmaf(__uint_as_float(r1[c]) + bcj, sb.x, -mr); r1[c] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                    v = fmaf(__uint_as_float(r1[c + 1]) + bcj, sb.y, -mr); r1[c + 1] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                    v = fmaf(__uint_as_float(r1[c + 2]) + bcj, sb.z, -mr); r1[c + 2] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                    v = fmaf(__uint_as_float(r1[c + 3]) + bcj, sb.w, -mr); r1[c + 3] = __float_as_uint(v); tmax = fmaxf(tmax, v);
                }
                // lazy running max: the first tile fixes it, later tiles raise it only past 2^ST_RAISE
                const bool first = (t == 0);
                const bool need = first || tmax > raise_thr;
                if (__any_sync(0xffffffffu, need)) {
                    const float delta = need ? tmax : 0.f;   // m_new = m_run + delta
                    const float sc = first ? 0.f : fast_ex2(-delta);
                    m_run[jj] = mr + delta;
                    l_run[jj] *= sc;
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        r0[c] = __float_as_uint(__uint_as_float(r0[c]) - delta);
                        r1[c] = __float_as_uint(__uint_as_float(r1[c]) - delta);
                    }
                    if (!first) {   // rescale the TMEM accumulator of this slice block (all earlier MMA2s are done)
                        mbar_wait(&ctl->mma2_done[w], (uint32_t)((k - 1) & 1));
                        tc_fence_after();
                        uint32_t sr[32];
                        const uint32_t scol = tmem_base + lane_base + (uint32_t)j * ST_C;
                        tmem_ld32(scol, sr);
                        tmem_wait_ld();
#pragma unroll
                        for (int c = 0; c < 32; ++c) sr[c] = __float_as_uint(__uint_as_float(sr[c]) * sc);
                        tmem_st32(scol, sr);
                    }
                }
                float l = l_run[jj];
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const float p = fast_ex2(__uint_as_float(r0[c]));
                    l += p; r0[c] = __float_as_uint(p);
                }
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const float p = fast_ex2(__uint_as_float(r1[c]));
                    l += p; r1[c] = __float_as_uint(p);
                }
                l_run[jj] = l;
                tmem_st32(dcol, r0);
                tmem_st32(dcol + 32, r1);
                tmem_wait_st();
                tc_fence_before();
                mbar_arrive(&ctl->p_full[w][buf]);
            }
        }
        // ---- write this CTA's partial (m2, l, S[C]) per slice ----
        if (n_my > 0 && n_tiles > 0) {
            // (mma2_done may be two phases behind here, so its parity is ambiguous: use the dedicated barrier)
            mbar_wait(&ctl->all_done, 0);
            tc_fence_after();
            const int64_t n_pad = (int64_t)n_tiles * ST_TILE - (n_end - n_begin);   // masked columns of the last tile
            for (int jj = 0; jj < n_my; ++jj) {
                const int j = w + 2 * jj, g = j * ST_GB + row;
                uint32_t sr[32];
                tmem_ld32(tmem_base + lane_base + (uint32_t)j * ST_C, sr);
                tmem_wait_ld();
                if (g < G) {
                    // masked points entered with s_n = 0 and zero rows: each added exactly 2^(-m_run) to l, 0 to S
                    const float l = l_run[jj] - (float)n_pad * fast_ex2(-m_run[jj]);
                    float *o = parts + (((int64_t)blockIdx.x * BH + bh) * G + g) * (ST_C + 2);
                    o[0] = m_run[jj]; o[1] = l;
#pragma unroll
                    for (int c = 0; c < 32; ++c) o[2 + c] = __uint_as_float(sr[c]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 12) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

bool slice_tc_supported(const haet_layer_shape *s) { return s->dim_head == ST_C && s->slice_num <= ST_MAX_GB * ST_GB; }

static int64_t slice_tc_chunk_pts(const haet_layer_shape *s) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t BH = s->B * s->heads;
    int64_t n_chunks = sms / BH;
    if (n_chunks < 1) n_chunks = 1;
    const int64_t max_chunks = cdiv(s->N, ST_TILE);
    if (n_chunks > max_chunks) n_chunks = max_chunks;
    return align_up(cdiv(s->N, n_chunks), ST_TILE);
}
int64_t slice_tc_chunks(const haet_layer_shape *s) { return s->N > 0 ? cdiv(s->N, slice_tc_chunk_pts(s)) : 0; }

int launch_slice_tc(const haet_layer_shape *s, const float *xp, const float *ws, const float *bs, const float *wt,
                    float bt, float *chunk_parts, cudaStream_t st) {
    const int n_gb = (s->slice_num + ST_GB - 1) / ST_GB;
    const size_t smem = 1024 + (size_t)n_gb * 2 * (ST_GB * 128) + (size_t)ST_STAGES * 2 * ST_TILE_BYTES +
                        sizeof(float) * ST_STAGES * ST_TILE + sizeof(SliceCtl);
    static size_t smem_set = 0;
    if (smem > smem_set) {
        HAET_CUDA(cudaFuncSetAttribute(slice_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    const int64_t chunk_pts = slice_tc_chunk_pts(s);
    const char *env = getenv("HAET_ST_RAISE");   // test hook: force the accumulator-rescale path
    const float raise_thr = env ? (float)atof(env) : ST_RAISE;
    dim3 grid((unsigned)cdiv(s->N, chunk_pts), (unsigned)(s->B * s->heads));
    slice_tc_kernel<<<grid, ST_THREADS, smem, st>>>(xp, ws, bs, wt, bt, s->base_temp, rep_slice_shift(s->epsilon), s->N,
                                                     s->heads, s->slice_num, chunk_pts, raise_thr, chunk_parts);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet
