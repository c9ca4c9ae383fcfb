// Pass 1 on tcgen05: per (b, h, chunk of points) the running column softmax state over the points
//   m2[g] ~ max_n z2[n,g],  l[g] = sum_n 2^(z2-m2),  S[g,:] = sum_n 2^(z2-m2) x_proj[n,:]
// with z2 = ((x_proj . Ws_g + bs_g - log(-log eps)) / tau_n) log2 e     (IrregularMesh.py:168-183).
//
// Flash-attention-shaped, with the SLICES on the TMEM lanes so that the softmax over the points
// is a per-thread reduction.  The projected points arrive as the ready-made operand image of the spill
// (common.cuh: rows of [32 fp16 hi | 32 fp16 lo], pre-swizzled), one bulk copy per 64-point tile, and the SAME
// shared-memory tile feeds both MMAs:
//   MMA1 (kind::f16, 3-term fp16 split = fp32-class logits; precision_study.py), tile read K-major:
//        D[g, n] = Ws[g,:] . x_proj[n,:]          M = 128 slices, N = 64 points, K' = 96
//   softmax warps (thread = slice row g): t = D * s_n - m_run (the bias bc_g enters D through one extra K step), P = 2^t, l += P, P -> TMEM as packed fp16.
//        m_run is a LAZY running reference: it starts ST_HEAD = 7 log2 units BELOW the first tile's max and is only raised
//        when a tile exceeds it by 2^15 (P stays inside fp16), so the TMEM accumulator rescale is off the common path.
//        The 7 units of head room put the fp16 flush-to-zero point of P at 2^-31 of the largest weight seen: whatever
//        l (fp32, unrounded p) still counts and S (fp16 P) has lost is then < N_chunk 2^-31 of the column's mass.
//   MMA2 (kind::f16, A = P from TMEM, B = the hi half of the tile read MN-major: 32 columns):
//        S[g, c] += sum_n P[g,n] hi[n,c]      M = 128, N = 32, K = 64   (hi carries x_proj to 2^-12: ample for S)
// dim_head 64 ("pair mode"): a head is run as TWO pseudo-heads of 32 channels (image rows of pseudo-head 2h + sub = channels
// [32 sub, 32 sub + 32) of head h).  Both CTAs of a head compute the SAME logits -- MMA1 runs over both sub-row tiles, always
// sub-row 0 first, so D, the lazy reference and l are bit-identical in the two -- and each accumulates S for its own 32 channels
// (MMA2 on its own tile).  The exponentials are computed twice; everything else is shared with the dim_head-32 code.
// Warp roles: 4 softmax warpgroups (4 warps per scheduler hide the TMEM / MUFU latencies), 1 loader warp (bulk copies),
// 2 MMA-issuing warps (a single issuing thread was the bottleneck of the first version).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc.cuh"

namespace haet {
using namespace tc;

#ifdef HAET_TRACE   // development build only (scripts/trace_pass2.py): cycle counters per warp role, never in the shipped library
__device__ unsigned long long g_st_trace[64];
#define ST_T(var) const long long var = clock64()
#define ST_DECL unsigned long long st_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define ST_ACC(slot, v) st_acc[(slot) & 7] += (unsigned long long)(v)
#define ST_FLUSH(base) do { if (lane == 0) for (int i_ = 0; i_ < 8; ++i_) atomicAdd(&g_st_trace[(base) + i_], st_acc[i_]); } while (0)
#else
#define ST_T(var)
#define ST_DECL
#define ST_ACC(slot, v)
#define ST_FLUSH(base)
#endif

constexpr int ST_C = 32;            // dim_head handled by this kernel
constexpr int ST_TILE = 64;         // points per tile
constexpr int ST_STAGES = 10;
constexpr int ST_THREADS = 608;     // 16 softmax warps + loader warp + 2 MMA-issuing warps
constexpr int ST_GB = 128;          // slices per block (UMMA M)
constexpr int ST_MAX_GB = 4;        // up to 512 slices
constexpr int ST_TILE_BYTES = ST_TILE * 128;   // one operand tile: 64 rows x 128 B
constexpr int ST_SCOLS = 32;        // TMEM columns of one warpgroup's accumulator (P . hi; the fp16 hi half carries x_proj to 2^-12)
constexpr int ST_NBUF = 3;          // D / P buffers per issue stream
constexpr float ST_RAISE = 15.0f;   // lazy-max threshold (log2 units): P <= 2^15 fits fp16
constexpr float ST_HEAD = 7.0f;     // a (re)set reference sits this far below the max that caused it

struct SliceCtl {
    uint64_t b_full[ST_STAGES], b_empty[ST_STAGES];
    uint64_t d_full[2][ST_NBUF], p_full[2][ST_NBUF];
    uint64_t mma2_done[2][2];
    uint64_t all_done[2];
    uint32_t tmem_base;
};

// TMEM columns: accumulators of the 4 softmax warpgroups at 0..127, D / P buffers [stream][3] x 64 at 128..511
__device__ __forceinline__ uint32_t st_dbuf_col(int w, int buf) { return 128u + (uint32_t)(w * ST_NBUF + buf) * ST_TILE; }

// POLY: eighths (0..4) of every thread's logits whose 2^t comes from poly_ex2_pair (FMA / ALU pipes) instead of MUFU.EX2 (tc.cuh)
template <bool F2, int POLY>
__global__ void __launch_bounds__(ST_THREADS, 1)
slice_tc_kernel(const uint8_t *__restrict__ img, const float *__restrict__ scale, const float *__restrict__ ws,
                const float *__restrict__ bs, float shift, int64_t N, int64_t npad, int G, int C, int64_t chunk_pts,
                float raise_thr, float *__restrict__ parts) {   // C = dim_head: 32, 16 (zero-padded to 32 in the image) or 64 (pair mode)
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS / STS)
    const int n_gb = (G + ST_GB - 1) / ST_GB;
    const bool pair = C == 64;
    const int Cw = pair ? 32 : C;                                    // channels of this CTA's image rows
    const int n_stages = pair ? ST_STAGES / 2 : ST_STAGES;           // pair mode: a stage holds the tiles of both sub-rows
    const uint32_t stage_bytes = pair ? 2 * ST_TILE_BYTES : ST_TILE_BYTES;
    const uint32_t blk_bytes = (pair ? 4u : 2u) * (ST_GB * 128);     // Ws images of one slice block
    uint8_t *sW = smem;                                              // [n_gb][sub-row][2][16 KB]: [hi|hi] rows, [lo|bias] rows
    uint8_t *sOne = sW + (size_t)n_gb * blk_bytes;                   // [64 points][128 B]: k = 0, 1 are 1.0 (bias operand)
    uint8_t *sB = sOne + ST_TILE_BYTES;                              // [stages][8 KB (x 2)] point tiles
    float *sS = reinterpret_cast<float *>(sB + (size_t)ST_STAGES * ST_TILE_BYTES);   // [stages][64] s_n
    SliceCtl *ctl = reinterpret_cast<SliceCtl *>(sS + ST_STAGES * ST_TILE);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform role index
    const int bh = blockIdx.y;                                       // (b, h), or (b, h, sub-row) in pair mode
    const int srow = pair ? (bh & 1) : 0;                            // sub-row (channels 32 srow ..) this CTA accumulates
    const int64_t n_begin = (int64_t)blockIdx.x * chunk_pts;
    const int64_t n_end = min(N, n_begin + chunk_pts);
    const int n_tiles = (int)((n_end - n_begin + ST_TILE - 1) / ST_TILE);
    const int64_t BH = gridDim.y;

    // ---- one-time setup: barriers, TMEM, the Ws operand images ----
    if (warp == 17) {
        if (lane == 0) {
            for (int s = 0; s < ST_STAGES; ++s) { mbar_init(&ctl->b_full[s], 1); mbar_init(&ctl->b_empty[s], n_gb > 1 ? 2 : 1); }
            for (int w = 0; w < 2; ++w) {
                for (int k = 0; k < ST_NBUF; ++k) { mbar_init(&ctl->d_full[w][k], 1); mbar_init(&ctl->p_full[w][k], 128); }
                for (int k = 0; k < 2; ++k) mbar_init(&ctl->mma2_done[w][k], 1);
            }
            mbar_init(&ctl->all_done[0], 1);
            mbar_init(&ctl->all_done[1], 1);
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc<512>(&ctl->tmem_base);
    }
    for (int idx = threadIdx.x; idx < n_gb * ST_GB * 4 * (pair ? 2 : 1); idx += ST_THREADS) {   // one 16-B chunk (8 k) per step
        const int hf = idx / (n_gb * ST_GB * 4), id2 = idx - hf * (n_gb * ST_GB * 4);   // sub-row (pair mode)
        const int row = id2 >> 2, ck = id2 & 3, g = row, j = row / ST_GB, r = row % ST_GB;
        __align__(16) __half hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float v = (g < G && ck * 8 + i < Cw) ? ws[(size_t)g * C + hf * 32 + ck * 8 + i] : 0.f;
            hi[i] = __float2half_rn(v);
            lo[i] = __float2half_rn(v - __half2float(hi[i]));
        }
        uint8_t *w1 = sW + (size_t)j * blk_bytes + (size_t)hf * 2 * (ST_GB * 128), *w2 = w1 + ST_GB * 128;
        *reinterpret_cast<uint4 *>(w1 + sw128_off(r, ck)) = *reinterpret_cast<const uint4 *>(hi);
        *reinterpret_cast<uint4 *>(w1 + sw128_off(r, ck + 4)) = *reinterpret_cast<const uint4 *>(hi);
        *reinterpret_cast<uint4 *>(w2 + sw128_off(r, ck)) = *reinterpret_cast<const uint4 *>(lo);
        // k = 32, 33 of the lo rows carry the logit bias bc = bs_g - log(-log eps) as fp16 hi + lo: one more K step of
        // MMA1 against the constant tile sOne adds it to D, so the softmax warps do not spend an FADD per element on it
        uint4 bias4 = make_uint4(0, 0, 0, 0);
        if (ck == 0 && g < G && hf == 0) {
            const float bc = bs[g] - shift;
            const __half bh = __float2half_rn(bc), bl = __float2half_rn(bc - __half2float(bh));
            bias4.x = (uint32_t)__half_as_ushort(bh) | ((uint32_t)__half_as_ushort(bl) << 16);
        }
        *reinterpret_cast<uint4 *>(w2 + sw128_off(r, ck + 4)) = bias4;
    }
    for (int idx = threadIdx.x; idx < ST_TILE * 8; idx += ST_THREADS) {
        const int n = idx >> 3, c = idx & 7;
        *reinterpret_cast<uint4 *>(sOne + sw128_off(n, c)) = make_uint4(c == 0 ? 0x3c003c00u : 0u, 0, 0, 0);   // fp16 1.0, 1.0
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == 16) {
        // =========================== loader: one bulk copy per tile (+ its s_n) ===========================
        if (lane == 0) {
            const int64_t bh0 = pair ? (bh & ~1) : bh;   // pair mode: sub-row 0 of this head, sub-row 1 follows npad rows later
            const uint8_t *src = img + ((int64_t)bh0 * npad + n_begin) * XPI_ROW_BYTES;
            const float *ssrc = scale + (int64_t)bh * npad + n_begin;
            int stage = 0, phase = 0;
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait(&ctl->b_empty[stage], phase ^ 1);
                mbar_arrive_expect_tx(&ctl->b_full[stage], stage_bytes + ST_TILE * 4);
                bulk_g2s(sB + (size_t)stage * stage_bytes, src + (size_t)t * ST_TILE_BYTES, ST_TILE_BYTES, &ctl->b_full[stage]);
                if (pair)
                    bulk_g2s(sB + (size_t)stage * stage_bytes + ST_TILE_BYTES, src + ((size_t)npad + (size_t)t * ST_TILE) * XPI_ROW_BYTES,
                             ST_TILE_BYTES, &ctl->b_full[stage]);
                bulk_g2s(sS + stage * ST_TILE, ssrc + (size_t)t * ST_TILE, ST_TILE * 4, &ctl->b_full[stage]);
                if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp >= 17) {
        // =========================== MMA issuers ===========================
        // warp 17 issues stream 0 (slice blocks j = 0, 2), warp 18 stream 1 (j = 1, 3): two instruction streams, each
        // MMA1 three items ahead of MMA2; item k of a stream uses D / P buffer k % 3 and belongs to softmax warpgroup (w, k & 1).  The whole warp runs the (warp-uniform)
        // loop and one elected lane issues, so operands stay in uniform registers.
        const int w = warp - 17;
        const int pw = (n_gb + 1 - w) >> 1;   // slice blocks of this warpgroup
        const int Qw = n_tiles * pw;
        const uint32_t id1 = instr_desc(FMT_F16, ST_GB, ST_TILE, false, false);
        const uint32_t id2 = instr_desc(FMT_F16, ST_GB, ST_SCOLS, false, true);
        // Low descriptor words (tc.cuh: desc_join): the issuing warp was ~70 % busy per item rebuilding and masking
        // descriptors, so they are kept as running sums -- one add per MMA.
        constexpr uint32_t HI = DESC_HI_SW128_SBO1024;
        const uint32_t STG = stage_bytes >> 4;                              // one stage, in 16-byte units
        constexpr uint32_t TIL = ST_TILE_BYTES >> 4;                        // one tile
        const uint32_t sB_a = smem_u32(sB) >> 4;
        const uint32_t b1_lo0 = sB_a | (1u << 16);                          // point tile, K-major (LBO 16 B)
        const uint32_t b2_lo0 = (sB_a + (uint32_t)srow * TIL) | (64u << 16);  // this CTA's own tile MN-major (LBO 1024 B)
        const uint32_t one_lo = (smem_u32(sOne) >> 4) | (1u << 16);
        const uint32_t a_lo0 = ((smem_u32(sW) >> 4) + (uint32_t)w * (blk_bytes >> 4)) | (1u << 16);   // block j = w + 2 jj
        const uint32_t A_JJ = 2 * (blk_bytes >> 4);                                        // next block of the stream
        constexpr uint32_t A_LO = ST_GB * 128 >> 4;                                        // the [lo | bias] rows
        int jj1 = 0, st1 = 0, ph1 = 0, buf1 = 0;           // MMA1 cursor: block within the stream, stage (+phase), D buffer
        int jj2 = 0, st2 = 0, buf2 = 0, ph2 = 0;           // MMA2 cursor
        uint32_t b1_lo = b1_lo0, b2_lo = b2_lo0;           // running stage descriptors of the two cursors
        ST_DECL;
        for (int q = 0; q < Qw + ST_NBUF; ++q) {
            if (q >= ST_NBUF) {   // ---- MMA2 of this stream's item k = q - 3: S_wg += P . hi   (wg = k & 1)
                const int k = q - ST_NBUF;
                ST_T(m0);
                mbar_wait(&ctl->p_full[w][buf2], (uint32_t)ph2);
                tc_fence_after();
                ST_T(m1);
                ST_ACC(0, m1 - m0); ST_ACC(7, 1);
#ifdef HAET_TRACE
                const long long m1c = m1;
#endif
                if (elect_one()) {
                    const uint32_t pa = tmem_base + st_dbuf_col(w, buf2), sd = tmem_base + (uint32_t)(w * 2 + (k & 1)) * ST_SCOLS;
                    mma_f16_ts(sd, pa, desc_join(b2_lo, HI), id2, k > 1);   // K = 16 points = two 8-row swizzle atoms per step
                    mma_f16_ts(sd, pa + 8, desc_join(b2_lo + 128, HI), id2, true);
                    mma_f16_ts(sd, pa + 16, desc_join(b2_lo + 256, HI), id2, true);
                    mma_f16_ts(sd, pa + 24, desc_join(b2_lo + 384, HI), id2, true);
                    tc_commit(&ctl->mma2_done[w][k & 1]);
                    if (jj2 == pw - 1) tc_commit(&ctl->b_empty[st2]);
                }
                __syncwarp();
#ifdef HAET_TRACE
                ST_ACC(2, clock64() - m1c);
#endif
                if (++buf2 == ST_NBUF) { buf2 = 0; ph2 ^= 1; }
                if (++jj2 == pw) {
                    jj2 = 0;
                    b2_lo += STG;
                    if (++st2 == n_stages) { st2 = 0; b2_lo = b2_lo0; }
                }
            }
            if (q < Qw) {    // ---- MMA1 of item q: D = Ws_j . x_proj^T (hi.hi + hi.lo + lo.hi) + bc_j
                ST_T(m2);
                if (jj1 == 0) { mbar_wait(&ctl->b_full[st1], (uint32_t)ph1); tc_fence_after(); }
                ST_T(m3);
                ST_ACC(1, m3 - m2);
                if (elect_one()) {
                    const uint32_t al = a_lo0 + (uint32_t)jj1 * A_JJ;
                    const uint32_t d = tmem_base + st_dbuf_col(w, buf1);
                    mma_bf16_ss(d, desc_join(al, HI), desc_join(b1_lo, HI), id1, false);          // [hi | hi] . [x hi | x lo], k 0..63
                    mma_bf16_ss(d, desc_join(al + 2, HI), desc_join(b1_lo + 2, HI), id1, true);
                    mma_bf16_ss(d, desc_join(al + 4, HI), desc_join(b1_lo + 4, HI), id1, true);
                    mma_bf16_ss(d, desc_join(al + 6, HI), desc_join(b1_lo + 6, HI), id1, true);
                    mma_bf16_ss(d, desc_join(al + A_LO, HI), desc_join(b1_lo, HI), id1, true);     // lo . x hi, k 0..31
                    mma_bf16_ss(d, desc_join(al + A_LO + 2, HI), desc_join(b1_lo + 2, HI), id1, true);
                    mma_bf16_ss(d, desc_join(al + A_LO + 4, HI), desc_join(one_lo, HI), id1, true);   // + bc_g (k = 32..47 of the lo rows)
                    if (pair) {   // channels 32..63: the second sub-row's Ws images against the second tile of the stage
                        const uint32_t a1 = al + 2 * A_LO, t1 = b1_lo + TIL;
                        mma_bf16_ss(d, desc_join(a1, HI), desc_join(t1, HI), id1, true);
                        mma_bf16_ss(d, desc_join(a1 + 2, HI), desc_join(t1 + 2, HI), id1, true);
                        mma_bf16_ss(d, desc_join(a1 + 4, HI), desc_join(t1 + 4, HI), id1, true);
                        mma_bf16_ss(d, desc_join(a1 + 6, HI), desc_join(t1 + 6, HI), id1, true);
                        mma_bf16_ss(d, desc_join(a1 + A_LO, HI), desc_join(t1, HI), id1, true);
                        mma_bf16_ss(d, desc_join(a1 + A_LO + 2, HI), desc_join(t1 + 2, HI), id1, true);
                    }
                    tc_commit(&ctl->d_full[w][buf1]);
                }
                __syncwarp();
#ifdef HAET_TRACE
                ST_ACC(3, clock64() - m3);
#endif
                if (++buf1 == ST_NBUF) buf1 = 0;
                if (++jj1 == pw) {
                    jj1 = 0;
                    b1_lo += STG;
                    if (++st1 == n_stages) { st1 = 0; ph1 ^= 1; b1_lo = b1_lo0; }
                }
            }
        }
        ST_FLUSH(16);
        if (elect_one()) tc_commit(&ctl->all_done[w]);   // every MMA of this stream has completed when this phase flips
        __syncwarp();
    } else {
        // =========================== softmax warpgroups ===========================
        // 4 warpgroups: w = wg & 1 is the issue stream it serves, sub = wg >> 1 the D / P buffer of that stream it owns
        // (every other item).  With 2 blocks per stream that is always the same slice block; with 1 block the two
        // warpgroups split its tiles.  Either way a warpgroup owns ONE block j, one running (m, l) per thread (slice
        // row) and one TMEM accumulator, and writes one chunk partial.  The 64 columns of an item are processed as two
        // halves of 32 (registers), each with its own lazy-max check.
        const int wg = warp >> 2, w = wg & 1, sub = wg >> 1, q4 = warp & 3, row = q4 * 32 + lane;
        const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
        const int pw = (n_gb + 1 - w) >> 1, Qw = n_tiles * pw;
        const int j = w + (pw == 2 ? 2 * sub : 0), g = j * ST_GB + row;
        const uint32_t scol = tmem_base + lane_base + (uint32_t)(w * 2 + sub) * ST_SCOLS;
        int buf = sub, bph = 0;   // D / P buffer k % 3 of the current item and its phase (k / 3) & 1
        float m_run = 0.f, l_run = 0.f;
        const int last_valid = (int)(n_end - n_begin) - (n_tiles - 1) * ST_TILE;   // points in the last tile
        int t = 0, jj = sub, i = 0;   // item k = t * pw + jj of the stream; i = items this warpgroup has done
        while (pw > 0 && jj >= pw) { jj -= pw; ++t; }
        int stage = t % n_stages, sph = (t / n_stages) & 1;   // running stage + parity of tile t (no division per item)
        ST_DECL;
        ST_T(l_start);
        for (int k = sub; k < Qw; k += 2, ++i) {
            ST_T(t0);
            mbar_wait(&ctl->b_full[stage], (uint32_t)sph);
            ST_T(t1);
            const float4 *s4 = reinterpret_cast<const float4 *>(sS + stage * ST_TILE);
            const bool ragged = (t == n_tiles - 1) && last_valid < ST_TILE;
            mbar_wait(&ctl->d_full[w][buf], (uint32_t)bph);
            tc_fence_after();
            ST_T(t2);
            ST_ACC(0, t1 - t0); ST_ACC(1, t2 - t1); ST_ACC(7, 1);
            const uint32_t dcol = tmem_base + lane_base + st_dbuf_col(w, buf);
#pragma unroll 1
            for (int hh = 0; hh < 2; ++hh) {
                uint32_t r[32];
                tmem_ld32(dcol + hh * 32, r);
                tmem_wait_ld();
                // t, the half's max AND P = 2^t in one loop: P is computed speculatively at the current running max so
                // that the MUFU / FMA / conversion instructions of a warp interleave (a separate "max, then exp" phase
                // left the MUFU pipe idle whenever the scheduler's warps were in the same phase); the rare raise
                // below recomputes P from the kept t.
                float tmax = -INFINITY, l0 = 0.f, l1 = 0.f;
                uint32_t pk[16];   // P as packed fp16: word c = points (2c, 2c+1) of this half
                if (F2) {   // the same arithmetic on packed fp32 pairs: 2 FFMA2 + 2 FADD2 instead of 4 FFMA + 4 FADD per 4 logits
                    const uint64_t nm = f2_pack(-m_run, -m_run);
                    uint64_t l01 = f2_pack(0.f, 0.f);
#pragma unroll
                    for (int c = 0; c < 32; c += 4) {
                        const float4 sa = s4[hh * 8 + (c >> 2)];
                        const uint64_t va = f2_fma(f2_pack(__uint_as_float(r[c]), __uint_as_float(r[c + 1])), f2_pack(sa.x, sa.y), nm);
                        const uint64_t vb = f2_fma(f2_pack(__uint_as_float(r[c + 2]), __uint_as_float(r[c + 3])), f2_pack(sa.z, sa.w), nm);
                        float v0, v1, v2, v3;
                        f2_unpack(va, v0, v1); f2_unpack(vb, v2, v3);
                        r[c] = __float_as_uint(v0); r[c + 1] = __float_as_uint(v1); r[c + 2] = __float_as_uint(v2); r[c + 3] = __float_as_uint(v3);
                        tmax = fmaxf(fmaxf(tmax, v0), v1);   // FMNMX3
                        tmax = fmaxf(fmaxf(tmax, v2), v3);
                        float p0, p1, p2, p3;   // pair j of the half goes to the polynomial when poly_pair<POLY>(j) (tc.cuh)
                        if (poly_pair<POLY>(c >> 1)) poly_ex2_pair(v0, v1, p0, p1); else { p0 = fast_ex2(v0); p1 = fast_ex2(v1); }
                        if (poly_pair<POLY>((c >> 1) + 1)) poly_ex2_pair(v2, v3, p2, p3); else { p2 = fast_ex2(v2); p3 = fast_ex2(v3); }
                        l01 = f2_add(l01, f2_pack(p0, p1));
                        l01 = f2_add(l01, f2_pack(p2, p3));
                        const __half2 ha = __floats2half2_rn(p0, p1), hb = __floats2half2_rn(p2, p3);
                        pk[c >> 1] = *reinterpret_cast<const uint32_t *>(&ha);
                        pk[(c >> 1) + 1] = *reinterpret_cast<const uint32_t *>(&hb);
                    }
                    f2_unpack(l01, l0, l1);
                } else {
#pragma unroll
                for (int c = 0; c < 32; c += 4) {
                    const float4 sa = s4[hh * 8 + (c >> 2)];
                    const float v0 = fmaf(__uint_as_float(r[c]), sa.x, -m_run), v1 = fmaf(__uint_as_float(r[c + 1]), sa.y, -m_run);
                    const float v2 = fmaf(__uint_as_float(r[c + 2]), sa.z, -m_run), v3 = fmaf(__uint_as_float(r[c + 3]), sa.w, -m_run);
                    r[c] = __float_as_uint(v0); r[c + 1] = __float_as_uint(v1); r[c + 2] = __float_as_uint(v2); r[c + 3] = __float_as_uint(v3);
                    tmax = fmaxf(fmaxf(tmax, v0), v1);   // FMNMX3
                    tmax = fmaxf(fmaxf(tmax, v2), v3);
                    const float p0 = fast_ex2(v0), p1 = fast_ex2(v1), p2 = fast_ex2(v2), p3 = fast_ex2(v3);
                    l0 += p0; l1 += p1; l0 += p2; l1 += p3;
                    const __half2 ha = __floats2half2_rn(p0, p1), hb = __floats2half2_rn(p2, p3);
                    pk[c >> 1] = *reinterpret_cast<const uint32_t *>(&ha);
                    pk[(c >> 1) + 1] = *reinterpret_cast<const uint32_t *>(&hb);
                }
                }
                if (ragged) {   // columns past the chunk's last point: weight 0 (their image rows are zero)
                    tmax = -INFINITY;
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        if (hh * 32 + c >= last_valid) r[c] = __float_as_uint(-INFINITY);
                        tmax = fmaxf(tmax, __uint_as_float(r[c]));
                    }
                }
                // lazy running reference: the very first half fixes it (max - ST_HEAD), later ones raise it only past 2^raise_thr
                const bool first = (i == 0 && hh == 0);
                const bool need = first || tmax > raise_thr;
                if (ragged || __any_sync(0xffffffffu, need)) {   // warp-uniform (ragged is per tile)
                    const float delta = need ? tmax - ST_HEAD : 0.f;   // m_new = m_run + delta: the new max maps to 2^ST_HEAD
                    const float sc = first ? 0.f : fast_ex2(-delta);
                    m_run += delta;
                    l_run *= sc;
                    l0 = 0.f; l1 = 0.f;
#pragma unroll
                    for (int c = 0; c < 32; c += 2) {   // P again, at the raised max (and with the ragged columns masked)
                        const float p0 = fast_ex2(__uint_as_float(r[c]) - delta), p1 = fast_ex2(__uint_as_float(r[c + 1]) - delta);
                        l0 += p0; l1 += p1;
                        const __half2 h2 = __floats2half2_rn(p0, p1);
                        pk[c >> 1] = *reinterpret_cast<const uint32_t *>(&h2);
                    }
                    if (i > 0) {   // rescale this warpgroup's TMEM accumulator (its earlier MMA2s are done)
                        mbar_wait(&ctl->mma2_done[w][sub], (uint32_t)((i - 1) & 1));
                        tc_fence_after();
                        uint32_t sr[32];
                        tmem_ld32(scol, sr);
                        tmem_wait_ld();
#pragma unroll
                        for (int c = 0; c < 32; ++c) sr[c] = __float_as_uint(__uint_as_float(sr[c]) * sc);
                        tmem_st32(scol, sr);
                    }
                    if (hh == 1) {   // the first half's P of this item is already in TMEM at the old scale
                        uint32_t ph[16];
                        tmem_ld16(dcol, ph);
                        tmem_wait_ld();
#pragma unroll
                        for (int c = 0; c < 16; ++c) {
                            const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&ph[c]));
                            const __half2 h2 = __floats2half2_rn(f.x * sc, f.y * sc);
                            ph[c] = *reinterpret_cast<const uint32_t *>(&h2);
                        }
                        tmem_st16(dcol, ph);
                    }
                }
                l_run += l0 + l1;
                tmem_st16(dcol + hh * 16, pk);
            }
            ST_T(t3);
            tmem_wait_st();
            tc_fence_before();
            mbar_arrive(&ctl->p_full[w][buf]);
            ST_T(t4);
            ST_ACC(3, t3 - t2); ST_ACC(5, t4 - t3); ST_ACC(6, t4 - t0);
            buf += 2;
            if (buf >= ST_NBUF) { buf -= ST_NBUF; bph ^= 1; }
            jj += 2;
            while (jj >= pw) {
                jj -= pw;
                ++t;
                if (++stage == n_stages) { stage = 0; sph ^= 1; }
            }
        }
        ST_FLUSH(0);
#ifdef HAET_TRACE
        { ST_T(l_end); if (lane == 0) { atomicAdd(&g_st_trace[9], (unsigned long long)(l_end - l_start)); atomicAdd(&g_st_trace[10], 1ull); } }
#endif
        // ---- write this warpgroup's partial (m2, l, S[C]) per slice of its block: chunk slot 2 * blockIdx.x + c ----
        if (pw > 0) {
            const int slot = pw == 1 ? sub : 0;
            // pair mode: both sub-row CTAs of a head hold the same (m, l); sub-row 0 writes them, each writes its 32 channels of S
            const int64_t BHr = pair ? BH / 2 : BH, bhr = pair ? (bh >> 1) : bh;
            const int64_t slot_stride = BHr * (int64_t)G * (C + 2);
            float *o = parts + ((((int64_t)blockIdx.x * 2 + slot) * BHr + bhr) * G + g) * (C + 2);
            float *oS = o + 2 + srow * 32;
            const bool head_cols = srow == 0;
            if (i > 0) {
                mbar_wait(&ctl->all_done[w], 0);
                tc_fence_after();
                uint32_t s0[32];
                tmem_ld32(scol, s0);
                tmem_wait_ld();
                if (g < G) {
                    if (head_cols) { o[0] = m_run; o[1] = l_run; }
#pragma unroll
                    for (int c = 0; c < 32; ++c) if (c < Cw) oS[c] = __uint_as_float(s0[c]);
                }
            } else if (g < G) {   // no tile for this warpgroup: neutral partial
                if (head_cols) { o[0] = -INFINITY; o[1] = 0.f; }
#pragma unroll
                for (int c = 0; c < 32; ++c) if (c < Cw) oS[c] = 0.f;
            }
            if (pw == 2 && g < G) {   // the other slot of this block's rows is nobody's: neutral
                float *o2 = o + slot_stride;
                if (head_cols) { o2[0] = -INFINITY; o2[1] = 0.f; }
#pragma unroll
                for (int c = 0; c < 32; ++c) if (c < Cw) o2[2 + srow * 32 + c] = 0.f;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 17) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

bool slice_tc_supported(const haet_layer_shape *s) {   // dim_head 16 runs as zero-padded dim_head 32 (the kernels are bound by the N H G exponentials, not by C)
    if (s->dim_head == 64) return s->slice_num <= 2 * ST_GB;   // pair mode: the Ws images of both sub-rows must fit shared memory
    return (s->dim_head == ST_C || s->dim_head == 16) && s->slice_num <= ST_MAX_GB * ST_GB;
}

static int64_t slice_tc_chunk_pts(const haet_layer_shape *s) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t BH = s->B * s->heads * (s->dim_head == 64 ? 2 : 1);   // CTAs per chunk (pair mode: two sub-rows per head)
    int64_t n_chunks = sms / BH;
    if (n_chunks < 1) n_chunks = 1;
    const int64_t max_chunks = cdiv(s->N, ST_TILE);
    if (n_chunks > max_chunks) n_chunks = max_chunks;
    return align_up(cdiv(s->N, n_chunks), ST_TILE);
}
int64_t slice_tc_chunks(const haet_layer_shape *s) { return s->N > 0 ? 2 * cdiv(s->N, slice_tc_chunk_pts(s)) : 0; }   // 2 partial slots per CTA

int launch_slice_tc(const haet_layer_shape *s, const uint8_t *img, const float *scale, const float *ws,
                    const float *bs, float *chunk_parts, cudaStream_t st) {
    const int n_gb = (s->slice_num + ST_GB - 1) / ST_GB;
    const bool pair = s->dim_head == 64;
    const size_t smem = 1024 + (size_t)n_gb * (pair ? 4 : 2) * (ST_GB * 128) + (size_t)(ST_STAGES + 1) * ST_TILE_BYTES +
                        sizeof(float) * ST_STAGES * ST_TILE + sizeof(SliceCtl);
    HAET_DYN_SMEM((slice_tc_kernel<true, 0>), smem);
    HAET_DYN_SMEM((slice_tc_kernel<true, 1>), smem);
    HAET_DYN_SMEM((slice_tc_kernel<true, 2>), smem);
    HAET_DYN_SMEM((slice_tc_kernel<true, 3>), smem);
    HAET_DYN_SMEM((slice_tc_kernel<true, 4>), smem);
    HAET_DYN_SMEM((slice_tc_kernel<false, 0>), smem);
    const int64_t chunk_pts = slice_tc_chunk_pts(s);
    const float raise_thr = hooks().st_raise > 0.f ? hooks().st_raise : ST_RAISE;   // haet_set_raise_threshold: tests force the accumulator-rescale path
    dim3 grid((unsigned)cdiv(s->N, chunk_pts), (unsigned)(s->B * s->heads * (pair ? 2 : 1)));
    const int poly = options().no_f32x2 ? 0 : options().poly_p1;
    auto kern = options().no_f32x2 ? slice_tc_kernel<false, 0>
                : poly == 1 ? slice_tc_kernel<true, 1> : poly == 2 ? slice_tc_kernel<true, 2> : poly == 3 ? slice_tc_kernel<true, 3>
                : poly == 4 ? slice_tc_kernel<true, 4> : slice_tc_kernel<true, 0>;
    kern<<<grid, ST_THREADS, smem, st>>>(img, scale, ws, bs, rep_slice_shift(s->epsilon), s->N, xpi_npad(s->N), s->slice_num, s->dim_head,
                                          chunk_pts, raise_thr, chunk_parts);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet

#ifdef HAET_TRACE
extern "C" int haet_debug_st_trace(unsigned long long *out64, int reset) {
    if (out64 && cudaMemcpyFromSymbol(out64, haet::g_st_trace, sizeof(unsigned long long) * 64) != cudaSuccess) return -2;
    if (reset) { unsigned long long z[64] = {0}; if (cudaMemcpyToSymbol(haet::g_st_trace, z, sizeof(z)) != cudaSuccess) return -2; }
    return 0;
}
#endif
