// Erwin ball-tree transformer over the slice tokens (eval mode).
//
// Reference: models/components/erwinflash/erwin_flash.py:138-195 (forward), components/layers.py:30-32
// (block), :72-92 (BasicLayer, rotated odd blocks), components/attention.py:35-91 (BallMSA, flash
// variant: no distance bias, qkv columns laid out (three, head, head_dim)), components/mlp.py:17-18
// Optional (haet_erwin_cfg.attn_bias): the distance bias of the non-flash copy,
// benchmarks/02-Car-Design-ShapeNetCar/models/components/erwin/components/attention.py:28-32,49-51:
// logits += sigma_att[h] * ||p_i - p_j||_2 inside the ball (that copy's (H E K) qkv column order is
// permuted to (three, head, head_dim) when the weights are packed, haet_b200/erwin.py).
// (SwiGLU), components/pooling.py:25-40, 58-72 (BallPooling / BallUnpooling, BatchNorm1d eval).
//
// Everything inside one ErwinTransformerBlock is local to a ball (ball_size consecutive tokens of
// the -- possibly rotated -- sequence), so a block is ONE kernel with one CTA per ball: the ball's
// tokens are gathered through the permutation into shared memory, RMSNorm -> PE -> qkv -> attention
// -> proj -> RMSNorm -> SwiGLU all run there, and the result is scattered back in place.
#include <stdlib.h>

#include "common.cuh"
#include "tc.cuh"

namespace haet {

struct BlockW {  // float offsets into the packed weights
    int norm1, norm2, pe_w, pe_b, qkv_w, qkv_b, proj_w, proj_b, w1, b1, w2, b2, w3, b3;
    int sigma;   // [heads] sigma_att of the distance bias, or -1
};
struct PoolW { int w, b, scale, shift; };

__device__ __forceinline__ int pad_dim(int n) { return n + 4; }  // rows stay 16-byte aligned; LDS.128 conflict-free

// out[i][o] (+)= sum_k in[i][k] * W[o][k] + b[o], i < rows; consecutive threads take consecutive rows
// so W reads are warp-uniform (scalar variant: any K / O, used for the K = d positional projection)
template <bool ACC>
__device__ void ball_linear(const float *in, int in_ld, const float *__restrict__ W, const float *__restrict__ b,
                            float *out, int out_ld, int rows, int K, int O) {
    for (int t = threadIdx.x; t < rows * O; t += blockDim.x) {
        const int i = t % rows, o = t / rows;
        const float *wr = W + (size_t)o * K;
        const float *ir = in + i * in_ld;
        float acc = b ? b[o] : 0.f;
        for (int k = 0; k < K; ++k) acc = fmaf(ir[k], *(wr + k), acc);
        if (ACC) out[i * out_ld + o] += acc; else out[i * out_ld + o] = acc;
    }
}

// Register-blocked variant (K, O multiples of 4; rows 16-byte aligned): one thread = one row x 4 outputs, float4
// loads of the activations (shared) and of 4 weight rows (warp-uniform, L1-resident), 16 FMAs per step.
// GATE: out = (W2 x + b2) * silu(W x + b)   (SwiGLU, mlp.py:17-18), W2 / b2 the second weight pair.
template <bool ACC, bool GATE>
__device__ void ball_linear4(const float *in, int in_ld, const float *__restrict__ W, const float *__restrict__ b,
                             const float *__restrict__ W2, const float *__restrict__ b2, float *out, int out_ld, int rows,
                             int K, int O) {
    const int og = O >> 2, k4 = K >> 2;
    for (int t = threadIdx.x; t < rows * og; t += blockDim.x) {
        const int i = t % rows, o = (t / rows) << 2;
        const float4 *x = reinterpret_cast<const float4 *>(in + i * in_ld);
        const float4 *w = reinterpret_cast<const float4 *>(W + (size_t)o * K);
        const float4 bb = *(reinterpret_cast<const float4 *>(b + o));
        float a0 = bb.x, a1 = bb.y, a2 = bb.z, a3 = bb.w;
        float g0 = 0.f, g1 = 0.f, g2 = 0.f, g3 = 0.f;
        const float4 *w2 = nullptr;
        if (GATE) {
            w2 = reinterpret_cast<const float4 *>(W2 + (size_t)o * K);
            const float4 b2v = *(reinterpret_cast<const float4 *>(b2 + o));
            g0 = b2v.x; g1 = b2v.y; g2 = b2v.z; g3 = b2v.w;
        }
#pragma unroll 4
        for (int k = 0; k < k4; ++k) {
            const float4 xv = x[k];
            const float4 p = *(w + k), q = *(w + k4 + k), r = *(w + 2 * k4 + k), s = *(w + 3 * k4 + k);
            a0 = fmaf(xv.x, p.x, a0); a0 = fmaf(xv.y, p.y, a0); a0 = fmaf(xv.z, p.z, a0); a0 = fmaf(xv.w, p.w, a0);
            a1 = fmaf(xv.x, q.x, a1); a1 = fmaf(xv.y, q.y, a1); a1 = fmaf(xv.z, q.z, a1); a1 = fmaf(xv.w, q.w, a1);
            a2 = fmaf(xv.x, r.x, a2); a2 = fmaf(xv.y, r.y, a2); a2 = fmaf(xv.z, r.z, a2); a2 = fmaf(xv.w, r.w, a2);
            a3 = fmaf(xv.x, s.x, a3); a3 = fmaf(xv.y, s.y, a3); a3 = fmaf(xv.z, s.z, a3); a3 = fmaf(xv.w, s.w, a3);
            if (GATE) {
                const float4 p2 = *(w2 + k), q2 = *(w2 + k4 + k), r2 = *(w2 + 2 * k4 + k), s2 = *(w2 + 3 * k4 + k);
                g0 = fmaf(xv.x, p2.x, g0); g0 = fmaf(xv.y, p2.y, g0); g0 = fmaf(xv.z, p2.z, g0); g0 = fmaf(xv.w, p2.w, g0);
                g1 = fmaf(xv.x, q2.x, g1); g1 = fmaf(xv.y, q2.y, g1); g1 = fmaf(xv.z, q2.z, g1); g1 = fmaf(xv.w, q2.w, g1);
                g2 = fmaf(xv.x, r2.x, g2); g2 = fmaf(xv.y, r2.y, g2); g2 = fmaf(xv.z, r2.z, g2); g2 = fmaf(xv.w, r2.w, g2);
                g3 = fmaf(xv.x, s2.x, g3); g3 = fmaf(xv.y, s2.y, g3); g3 = fmaf(xv.z, s2.z, g3); g3 = fmaf(xv.w, s2.w, g3);
            }
        }
        if (GATE) {
            a0 = g0 * (a0 / (1.f + __expf(-a0))); a1 = g1 * (a1 / (1.f + __expf(-a1)));
            a2 = g2 * (a2 / (1.f + __expf(-a2))); a3 = g3 * (a3 / (1.f + __expf(-a3)));
        }
        float4 *op = reinterpret_cast<float4 *>(out + i * out_ld + o);
        if (ACC) { const float4 c = *op; a0 += c.x; a1 += c.y; a2 += c.z; a3 += c.w; }
        *op = make_float4(a0, a1, a2, a3);
    }
}

__device__ void ball_rmsnorm(const float *in, int ld, const float *__restrict__ gamma, float *out, int rows, int dim) {
    // nn.RMSNorm(dim), eps=None -> torch.finfo(float32).eps (layers.py:24-25)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int i = warp; i < rows; i += nw) {
        float ss = 0.f;
        for (int k = lane; k < dim; k += 32) { float v = in[i * ld + k]; ss = fmaf(v, v, ss); }
#pragma unroll
        for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        const float r = rsqrtf(ss / (float)dim + 1.1920928955078125e-07f);
        for (int k = lane; k < dim; k += 32) out[i * ld + k] = in[i * ld + k] * r * gamma[k];
    }
}

// sigma_h ||p_i - p_j||_2 in log2 units (rel holds the ball-relative positions, 4 floats per token, unused lanes zero)
__device__ __forceinline__ float dist_bias(const float *rel, int i, int j, float sig2) {
    const float dx = rel[4 * i] - rel[4 * j], dy = rel[4 * i + 1] - rel[4 * j + 1], dz = rel[4 * i + 2] - rel[4 * j + 2];
    return sig2 * sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
}

// softmax(q k^T / sqrt(hd) [+ sigma_h dist]) v inside the ball, one (query, head) per thread; HD4 = head_dim / 4 (1 or 2) fast paths
template <int HD4>
__device__ void ball_attention(const float *qkv, int ldq, float *us, int ld, int bs, int dim, int heads,
                               const float *__restrict__ sigma, const float *rel) {
    const int hd = HD4 * 4;
    const float scale = rsqrtf((float)hd) * kLog2e;
    for (int t = threadIdx.x; t < bs * heads; t += blockDim.x) {
        const int i = t % bs, h = t / bs;
        const float sig2 = sigma ? sigma[h] * kLog2e : 0.f;
        float4 q[HD4], o[HD4];
#pragma unroll
        for (int c = 0; c < HD4; ++c) {
            q[c] = *reinterpret_cast<const float4 *>(qkv + i * ldq + h * hd + 4 * c);
            q[c].x *= scale; q[c].y *= scale; q[c].z *= scale; q[c].w *= scale;
            o[c] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        float mx = -INFINITY, den = 0.f;
        for (int j = 0; j < bs; ++j) {   // online softmax over the ball's keys (k / v rows are warp-uniform reads)
            const float4 *kr = reinterpret_cast<const float4 *>(qkv + j * ldq + dim + h * hd);
            const float4 *vr = reinterpret_cast<const float4 *>(qkv + j * ldq + 2 * dim + h * hd);
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < HD4; ++c) {
                const float4 kv = kr[c];
                s = fmaf(q[c].x, kv.x, s); s = fmaf(q[c].y, kv.y, s); s = fmaf(q[c].z, kv.z, s); s = fmaf(q[c].w, kv.w, s);
            }
            if (sigma) s += dist_bias(rel, i, j, sig2);
            const float mn = fmaxf(mx, s);
            const float corr = exp2f(mx - mn), p = exp2f(s - mn);
            mx = mn;
            den = fmaf(den, corr, p);
#pragma unroll
            for (int c = 0; c < HD4; ++c) {
                const float4 vv = vr[c];
                o[c].x = fmaf(o[c].x, corr, p * vv.x); o[c].y = fmaf(o[c].y, corr, p * vv.y);
                o[c].z = fmaf(o[c].z, corr, p * vv.z); o[c].w = fmaf(o[c].w, corr, p * vv.w);
            }
        }
        const float inv = 1.f / den;
#pragma unroll
        for (int c = 0; c < HD4; ++c)
            *reinterpret_cast<float4 *>(us + i * ld + h * hd + 4 * c) =
                make_float4(o[c].x * inv, o[c].y * inv, o[c].z * inv, o[c].w * inv);
    }
}

// generic head_dim (scalar two-pass)
__device__ void ball_attention_any(const float *qkv, int ldq, float *us, int ld, int bs, int dim, int heads,
                                   const float *__restrict__ sigma, const float *rel) {
    const int hd = dim / heads;
    const float scale = rsqrtf((float)hd) * kLog2e;
    for (int t = threadIdx.x; t < bs * heads; t += blockDim.x) {
        const int i = t % bs, h = t / bs;
        const float sig2 = sigma ? sigma[h] * kLog2e : 0.f;
        const float *q = qkv + i * ldq + h * hd;
        float mx = -INFINITY;
        for (int j = 0; j < bs; ++j) {
            const float *k = qkv + j * ldq + dim + h * hd;
            float s = 0.f;
            for (int c = 0; c < hd; ++c) s = fmaf(q[c], k[c], s);
            mx = fmaxf(mx, s * scale + (sigma ? dist_bias(rel, i, j, sig2) : 0.f));
        }
        float *o = us + i * ld + h * hd;
        for (int c = 0; c < hd; ++c) o[c] = 0.f;
        float den = 0.f;
        for (int j = 0; j < bs; ++j) {
            const float *k = qkv + j * ldq + dim + h * hd;
            const float *v = qkv + j * ldq + 2 * dim + h * hd;
            float s = 0.f;
            for (int c = 0; c < hd; ++c) s = fmaf(q[c], k[c], s);
            const float p = exp2f(s * scale + (sigma ? dist_bias(rel, i, j, sig2) : 0.f) - mx);
            den += p;
            for (int c = 0; c < hd; ++c) o[c] = fmaf(p, v[c], o[c]);
        }
        const float inv = 1.f / den;
        for (int c = 0; c < hd; ++c) o[c] *= inv;
    }
}

constexpr int EB_THREADS = 256;

// tokens / positions that OTHER CTAs of the same launch may have written in an earlier phase of the fused kernel are read
// with ld.global.cg (L2, never the non-coherent path); the stand-alone kernel reads them through the read-only path
template <bool COH> __device__ __forceinline__ float4 ld_tok4(const float4 *p) { return COH ? __ldcg(p) : __ldg(p); }
template <bool COH> __device__ __forceinline__ float ld_tok(const float *p) { return COH ? __ldcg(p) : __ldg(p); }

// One ErwinTransformerBlock (layers.py:30-32) on the ball of bs tokens starting at row `base` of the (possibly rotated)
// sequence.  wts: the packed weights (the offsets of `w` are relative to it; shared or global memory); act: shared-memory
// scratch of block_smem(bs, dim, hidden) bytes.  All threads of the CTA take part (contains __syncthreads).
template <bool COH>
__device__ void erwin_block_body(float *__restrict__ x, const float *__restrict__ pos, const int64_t *__restrict__ perm,
                                 const float *wts, const BlockW &w, int bs, int dim, int heads, int hidden, int d, int64_t base,
                                 float *act, uint64_t *wbar, uint32_t wparity) {
    const int ld = pad_dim(dim), ldq = pad_dim(3 * dim), ldh = pad_dim(hidden);
    float *xs = act;                   // [bs][ld]   residual stream
    float *us = xs + bs * ld;          // [bs][ld]   normed / attention output
    float *qkv = us + bs * ld;         // [bs][ldq]
    float *hid = qkv + bs * ldq;       // [bs][ldh]  gated SwiGLU product
    float *rel = hid + bs * ldh;       // [bs][4]
    float *cen = rel + bs * 4;         // [4]
    int *tok = reinterpret_cast<int *>(cen + 4);  // [bs]
    const int tid = threadIdx.x;
    const bool vec = (dim & 3) == 0 && (hidden & 3) == 0;   // float4 paths

    for (int i = tid; i < bs; i += blockDim.x) tok[i] = perm ? (int)(COH ? __ldcg(perm + base + i) : perm[base + i]) : (int)(base + i);
    __syncthreads();
    if (vec) {
        const int d4 = dim >> 2;
        for (int t = tid; t < bs * d4; t += blockDim.x) {
            const int i = t / d4, k = t - i * d4;
            *reinterpret_cast<float4 *>(xs + i * ld + 4 * k) = ld_tok4<COH>(reinterpret_cast<const float4 *>(x + (int64_t)tok[i] * dim) + k);
        }
    } else {
        for (int t = tid; t < bs * dim; t += blockDim.x) {
            const int i = t / dim, k = t - i * dim;
            xs[i * ld + k] = ld_tok<COH>(x + (int64_t)tok[i] * dim + k);
        }
    }
    for (int t = tid; t < bs * 4; t += blockDim.x) {
        const int i = t >> 2, j = t & 3;
        rel[t] = j < d ? ld_tok<COH>(pos + (int64_t)tok[i] * d + j) : 0.f;
    }
    __syncthreads();
    if (tid < d) {  // ball centre: mean over the ball (attention.py:35-40)
        float s = 0.f;
        for (int i = 0; i < bs; ++i) s += rel[i * 4 + tid];
        cen[tid] = s / (float)bs;
    }
    if (wbar) tc::mbar_wait(wbar, wparity);   // the staged weights have landed
    ball_rmsnorm(xs, ld, wts + w.norm1, us, bs, dim);
    __syncthreads();
    for (int t = tid; t < bs * d; t += blockDim.x) { const int i = t / d, j = t - i * d; rel[i * 4 + j] -= cen[j]; }
    __syncthreads();
    ball_linear<true>(rel, 4, wts + w.pe_w, wts + w.pe_b, us, ld, bs, d, dim);        // x + pe_proj(rel_pos)
    __syncthreads();
    if (vec) ball_linear4<false, false>(us, ld, wts + w.qkv_w, wts + w.qkv_b, nullptr, nullptr, qkv, ldq, bs, dim, 3 * dim);
    else ball_linear<false>(us, ld, wts + w.qkv_w, wts + w.qkv_b, qkv, ldq, bs, dim, 3 * dim);
    __syncthreads();
    {
        const int hd = dim / heads;
        const float *sigma = w.sigma >= 0 ? wts + w.sigma : nullptr;
        if (vec && hd == 8) ball_attention<2>(qkv, ldq, us, ld, bs, dim, heads, sigma, rel);
        else if (vec && hd == 4) ball_attention<1>(qkv, ldq, us, ld, bs, dim, heads, sigma, rel);
        else ball_attention_any(qkv, ldq, us, ld, bs, dim, heads, sigma, rel);
    }
    __syncthreads();
    if (vec) ball_linear4<true, false>(us, ld, wts + w.proj_w, wts + w.proj_b, nullptr, nullptr, xs, ld, bs, dim, dim);
    else ball_linear<true>(us, ld, wts + w.proj_w, wts + w.proj_b, xs, ld, bs, dim, dim);   // x + BMSA(...)
    __syncthreads();
    ball_rmsnorm(xs, ld, wts + w.norm2, us, bs, dim);
    __syncthreads();
    if (vec) {
        ball_linear4<false, true>(us, ld, wts + w.w1, wts + w.b1, wts + w.w2, wts + w.b2, hid, ldh, bs, dim, hidden);
    } else {
        ball_linear<false>(us, ld, wts + w.w1, wts + w.b1, hid, ldh, bs, dim, hidden);
        float *hid2 = reinterpret_cast<float *>(tok + bs);   // [bs][ldh], only carved for the scalar path (block_smem)
        ball_linear<false>(us, ld, wts + w.w2, wts + w.b2, hid2, ldh, bs, dim, hidden);
        __syncthreads();
        for (int t = tid; t < bs * hidden; t += blockDim.x) {   // w2(x) * silu(w1(x))   (mlp.py:17-18)
            const int i = t / hidden, k = t - i * hidden;
            const float a = hid[i * ldh + k];
            hid[i * ldh + k] = hid2[i * ldh + k] * (a / (1.f + __expf(-a)));
        }
    }
    __syncthreads();
    if (vec) ball_linear4<true, false>(hid, ldh, wts + w.w3, wts + w.b3, nullptr, nullptr, xs, ld, bs, hidden, dim);
    else ball_linear<true>(hid, ldh, wts + w.w3, wts + w.b3, xs, ld, bs, hidden, dim);
    __syncthreads();
    if (vec) {
        const int d4 = dim >> 2;
        for (int t = tid; t < bs * d4; t += blockDim.x) {
            const int i = t / d4, k = t - i * d4;
            reinterpret_cast<float4 *>(x + (int64_t)tok[i] * dim)[k] = *reinterpret_cast<const float4 *>(xs + i * ld + 4 * k);
        }
    } else {
        for (int t = tid; t < bs * dim; t += blockDim.x) {
            const int i = t / dim, k = t - i * dim;
            x[(int64_t)tok[i] * dim + k] = xs[i * ld + k];
        }
    }
}

__global__ void __launch_bounds__(EB_THREADS) erwin_block_kernel(float *__restrict__ x, const float *__restrict__ pos,
                                                                 const int64_t *__restrict__ perm,
                                                                 const float *__restrict__ wts_g, BlockW w, int bs, int dim,
                                                                 int heads, int hidden, int d, int w_lo, int w_n) {
    extern __shared__ __align__(16) float sm[];
    // the block's w_n packed weights (contiguous from float offset w_lo) are staged into shared memory by ONE bulk
    // copy that overlaps the token gather: every linear layer then reads its weights at shared-memory latency (the
    // kernel was bound by the L2 latency of those loads: 44 % long-scoreboard stalls, profiles/r1_ncu_summary.md)
    uint64_t *wbar = reinterpret_cast<uint64_t *>(sm);   // sm[0..3]
    float *wsm = sm + 4;                                  // [w_n]
    const float *wts = wts_g;
    // programmatic dependent launch (run_blocks sets the attribute): the NEXT kernel of the stream may start now -- its CTAs
    // land on the idle SMs (64 balls on 148 SMs) and stage their weights while this one computes
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (w_n) {
        if (threadIdx.x == 0) {
            tc::mbar_init(wbar, 1);
            tc::fence_barrier_init();
            tc::mbar_arrive_expect_tx(wbar, (uint32_t)w_n * 4u);
            tc::bulk_g2s(wsm, wts_g + w_lo, (uint32_t)w_n * 4u, wbar);   // weights are constants: no dependency on the previous kernel
        }
        wts = wsm - w_lo;
    }
    // everything below reads what the previous kernel of the stream wrote (tokens, positions, permutations)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    erwin_block_body<false>(x, pos, perm, wts, w, bs, dim, heads, hidden, d, (int64_t)blockIdx.x * bs, wsm + w_n,
                            w_n ? wbar : nullptr, 0);
}

static size_t block_smem(int bs, int dim, int hidden) {
    auto pd = [](int n) { return n + 4; };
    const size_t scalar_extra = ((dim | hidden) & 3) ? (size_t)bs * pd(hidden) : 0;   // hid2 of the non-float4 path
    return sizeof(float) * ((size_t)bs * (2 * pd(dim) + pd(3 * dim) + pd(hidden)) + bs * 4 + 4 + scalar_extra) + sizeof(int) * bs;
}

__global__ void erwin_gather_kernel(const float *__restrict__ x, const float *__restrict__ pos,
                                    const int64_t *__restrict__ tree_idx, int64_t n_slots, int dim, int d,
                                    float *__restrict__ x0, float *__restrict__ p0) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int W = dim + d;
    if (t >= n_slots * W) return;
    const int64_t i = t / W; const int k = (int)(t - i * W);
    const int64_t src = tree_idx[i];
    if (k < dim) x0[i * dim + k] = x[src * dim + k];
    else p0[i * d + (k - dim)] = pos[src * d + (k - dim)];
}

// erwin_flash.py:193: x[tree_mask][argsort(tree_idx[tree_mask])]  ==  out[tree_idx[i]] = x[i] where mask[i]
__global__ void erwin_scatter_kernel(const float *__restrict__ x0, const int64_t *__restrict__ tree_idx,
                                     const uint8_t *__restrict__ mask, int64_t n_slots, int dim,
                                     float *__restrict__ out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_slots * dim) return;
    const int64_t i = t / dim; const int k = (int)(t - i * dim);
    if (mask[i]) out[tree_idx[i] * dim + k] = x0[t];
}

// BallPooling (pooling.py:25-40) with eval BatchNorm folded to scale/shift
__global__ void erwin_pool_kernel(const float *__restrict__ xc, const float *__restrict__ pc, int64_t n_par, int s,
                                  int in_dim, int out_dim, int d, const float *__restrict__ wts, PoolW w,
                                  float *__restrict__ xo, float *__restrict__ po, int raw = 0) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_par * out_dim) return;
    const int64_t n = t / out_dim; const int o = (int)(t - n * out_dim);
    const int K = s * in_dim + s * d;
    const float *wr = wts + w.w + (size_t)o * K;
    float acc = wts[w.b + o];
    const float *xr = xc + n * s * in_dim;
    for (int k = 0; k < s * in_dim; ++k) acc = fmaf(xr[k], wr[k], acc);
    for (int j = 0; j < d; ++j) {
        float c = 0.f;
        for (int m = 0; m < s; ++m) c += pc[(n * s + m) * d + j];
        c = c / (float)s;
        if (o == 0) po[n * d + j] = c;
        for (int m = 0; m < s; ++m) acc = fmaf(pc[(n * s + m) * d + j] - c, wr[s * in_dim + m * d + j], acc);
    }
    xo[t] = raw ? acc : fmaf(acc, wts[w.scale + o], wts[w.shift + o]);   // raw: train-mode BatchNorm follows (erwin_bn_train_kernel)
}

// BallUnpooling (pooling.py:58-72): children.x = BN(children.x + proj([x_parent | rel_children]))
__global__ void erwin_unpool_kernel(const float *__restrict__ xp, const float *__restrict__ pp, int64_t n_par, int s,
                                    int in_dim, int out_dim, int d, const float *__restrict__ wts, PoolW w,
                                    float *__restrict__ xc, const float *__restrict__ pc, int raw = 0) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_par * s * out_dim) return;
    const int64_t tokc = t / out_dim; const int c = (int)(t - tokc * out_dim);
    const int64_t n = tokc / s; const int m = (int)(tokc - n * s);
    const int K = in_dim + s * d;
    const int row = m * out_dim + c;
    const float *wr = wts + w.w + (size_t)row * K;
    float acc = wts[w.b + row];
    const float *xr = xp + n * in_dim;
    for (int k = 0; k < in_dim; ++k) acc = fmaf(xr[k], wr[k], acc);
    for (int mm = 0; mm < s; ++mm)
        for (int j = 0; j < d; ++j)
            acc = fmaf(pc[(n * s + mm) * d + j] - pp[n * d + j], wr[in_dim + mm * d + j], acc);
    xc[t] = raw ? xc[t] + acc : fmaf(xc[t] + acc, wts[w.scale + c], wts[w.shift + c]);
}

// Train-mode nn.BatchNorm1d over the n rows of x [n, C] (pooling.py:23,56 with module.training): batch mean and BIASED batch
// variance per channel, y = (x - mean) / sqrt(var + eps) * gamma + beta in place; (mean, var) go to stats[0:C], stats[C:2C] so
// that the host module can update running_mean / running_var (unbiased, momentum) as torch does.  One CTA per channel, fixed
// reduction order (strided partials, shared-memory tree): bit-reproducible.  The packed weights carry (gamma, beta) in the
// slots that hold the folded (scale, shift) in eval mode.
__global__ void __launch_bounds__(256) erwin_bn_train_kernel(float *__restrict__ x, int64_t n, int C, const float *__restrict__ wts,
                                                             PoolW w, float eps, float *__restrict__ stats) {
    __shared__ float red[256];
    const int c = blockIdx.x, tid = threadIdx.x;
    auto block_sum = [&](float v) {
        red[tid] = v;
        __syncthreads();
        for (int o = 128; o; o >>= 1) {
            if (tid < o) red[tid] += red[tid + o];
            __syncthreads();
        }
        const float r = red[0];
        __syncthreads();
        return r;
    };
    float a = 0.f;
    for (int64_t i = tid; i < n; i += 256) a += x[i * C + c];
    const float mean = block_sum(a) / (float)n;
    float q = 0.f;
    for (int64_t i = tid; i < n; i += 256) { const float dlt = x[i * C + c] - mean; q = fmaf(dlt, dlt, q); }
    const float var = block_sum(q) / (float)n;
    const float inv = 1.0f / sqrtf(var + eps), g = wts[w.scale + c], b = wts[w.shift + c];
    for (int64_t i = tid; i < n; i += 256) x[i * C + c] = fmaf((x[i * C + c] - mean) * inv, g, b);
    if (stats && tid == 0) { stats[c] = mean; stats[C + c] = var; }
}

// ---------------------------------------------------------------------------------------------------------------------
// The whole Erwin pass as ONE cooperative kernel (round 2; OPT-IN with HAET_ERWIN_FUSED=1 -- measured SLOWER than the
// launch-per-block path with programmatic dependent launch: 0.193 against 0.165 ms at config 4 (64 CTAs, 10 phases): a per-tree
// barrier costs what a dependent launch costs under graph replay, and the weights of a block cannot be staged while the previous
// block of the same CTA still uses the staging buffer, which the separate launches overlap for free on idle SMs).  Everything in erwin_flash.py:176-193 is local to a TREE: balls,
// rotated permutations, pooling and unpooling never cross trees, eval-mode BatchNorm is a per-channel affine.  So the CTAs
// are grouped per tree (group = the tree's balls at the level with most balls) and the 10 launches become 10 phases separated
// by per-tree barriers (a counter in global memory per tree; the launch is cooperative, so all CTAs are co-resident and a
// spin on another CTA's arrival is legal).  Tokens that cross a phase are read with ld.global.cg after an acquire fence.
// The next block's weights are bulk-copied into shared memory before the barrier wait, so the copy hides behind it.
// ---------------------------------------------------------------------------------------------------------------------
enum { PH_GATHER = 0, PH_BLOCK = 1, PH_POOL = 2, PH_UNPOOL = 3, PH_SCATTER = 4 };
struct ErwinPhase {
    int type, level;        // level of x / pos (PH_POOL, PH_UNPOOL: the CHILD level; parents live at level + 1)
    int use_rot;            // PH_BLOCK: run on the rotated permutation of its level
    int heads, w_lo, w_n;   // PH_BLOCK: heads, packed weights [w_lo, w_lo + w_n) staged in shared memory (w_n == 0: read from global)
    BlockW bw;
    PoolW pw;
};
constexpr int EF_MAX_PHASES = 24;
constexpr int EF_MAX_TREES = 4096;
struct ErwinProgram {
    int n_phases, n_levels, d, mlp_ratio;
    int n_trees, group_ctas;                 // CTAs per tree group; CTA c belongs to tree c / group_ctas
    int dim[HAET_MAX_LEVELS], bs[HAET_MAX_LEVELS], stride[HAET_MAX_LEVELS];
    int64_t npt[HAET_MAX_LEVELS];            // tokens per tree at each level
    float *xs[HAET_MAX_LEVELS], *ps[HAET_MAX_LEVELS];
    const int64_t *rot[HAET_MAX_LEVELS];
    int act_off;                             // float offset of the activation scratch behind the staged weights
    unsigned long long timeout_ns;
    ErwinPhase ph[EF_MAX_PHASES];
};

__device__ __forceinline__ unsigned long long ef_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// barrier among the CTAs of one tree group: arrival k of this launch completes when the counter reaches k * group_ctas
__device__ __forceinline__ void ef_group_barrier(unsigned int *counter, unsigned int target, unsigned long long timeout_ns) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        const unsigned long long t0 = ef_now();
        unsigned int v;
        do {
            asm volatile("ld.acquire.gpu.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
            if (v < target && ef_now() - t0 > timeout_ns) {   // a missing CTA must not hang the GPU: fail loudly
                printf("haet erwin_fused_kernel: tree barrier timed out (block %d, %u of %u)\n", blockIdx.x, v, target);
                asm volatile("trap;");
            }
        } while (v < target);
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(EB_THREADS) erwin_fused_kernel(const float *__restrict__ x_in, const float *__restrict__ pos_in,
                                                                 const int64_t *__restrict__ tree_idx, const uint8_t *__restrict__ tree_mask,
                                                                 const float *__restrict__ wts_g, float *__restrict__ out,
                                                                 unsigned int *__restrict__ counters, const __grid_constant__ ErwinProgram P) {
    extern __shared__ __align__(16) float sm[];
    uint64_t *wbar = reinterpret_cast<uint64_t *>(sm);   // sm[0..3]
    float *wsm = sm + 4;
    float *act = sm + 4 + P.act_off;
    const int tid = threadIdx.x;
    const int tree = blockIdx.x / P.group_ctas, j = blockIdx.x - tree * P.group_ctas;   // j: this CTA inside its tree group
    const int gthreads = P.group_ctas * EB_THREADS, gt = j * EB_THREADS + tid;          // thread index inside the group
    unsigned int *ctr = counters + tree;
    if (tid == 0) { tc::mbar_init(wbar, 1); tc::fence_barrier_init(); }
    __syncthreads();
    uint32_t wpar = 0;
    auto stage_weights = [&](const ErwinPhase &ph) {   // issue the bulk copy of a block phase's weights (thread 0)
        if (ph.type == PH_BLOCK && ph.w_n && tid == 0) {
            tc::mbar_arrive_expect_tx(wbar, (uint32_t)ph.w_n * 4u);
            tc::bulk_g2s(wsm, wts_g + ph.w_lo, (uint32_t)ph.w_n * 4u, wbar);
        }
    };
    const int d = P.d;
    for (int k = 0; k < P.n_phases; ++k) {
        const ErwinPhase &ph = P.ph[k];
        const int L = ph.level;
        const bool early = ph.type != PH_BLOCK && k + 1 < P.n_phases;   // the staging buffer is free during this phase
        if (early) stage_weights(P.ph[k + 1]);
        if (ph.type == PH_GATHER) {   // erwin_flash.py:176-181: x[tree_idx], pos[tree_idx]
            const int dim = P.dim[0], W = dim + d;
            const int64_t n = P.npt[0], r0 = (int64_t)tree * n;
            for (int64_t t = gt; t < n * W; t += gthreads) {
                const int64_t i = r0 + t / W;
                const int c = (int)(t % W);
                const int64_t src = tree_idx[i];
                if (c < dim) P.xs[0][i * dim + c] = x_in[src * dim + c];
                else P.ps[0][i * d + (c - dim)] = pos_in[src * d + (c - dim)];
            }
        } else if (ph.type == PH_BLOCK) {
            const int bs = P.bs[L], dim = P.dim[L];
            const int64_t balls = P.npt[L] / bs;
            const float *wts = ph.w_n ? wsm - ph.w_lo : wts_g;
            // every CTA of the group waits for its staged weights (its own barrier) even if it has no ball at this level
            if (j < balls)
                erwin_block_body<true>(P.xs[L], P.ps[L], ph.use_rot ? P.rot[L] : nullptr, wts, ph.bw, bs, dim, ph.heads,
                                       dim * P.mlp_ratio, d, ((int64_t)tree * balls + j) * bs, act, ph.w_n ? wbar : nullptr, wpar);
            else if (ph.w_n)
                tc::mbar_wait(wbar, wpar);
            if (ph.w_n) wpar ^= 1;
            __syncthreads();   // the weights in wsm are dead from here: the next phase's copy may overwrite them
        } else if (ph.type == PH_POOL) {   // BallPooling (pooling.py:25-40), eval BatchNorm folded
            const int s = P.stride[L], in_dim = P.dim[L], out_dim = P.dim[L + 1];
            const int64_t n_par = P.npt[L + 1], p0 = (int64_t)tree * n_par;
            const int K = s * in_dim + s * d;
            const float *xc = P.xs[L], *pc = P.ps[L];
            for (int64_t t = gt; t < n_par * out_dim; t += gthreads) {
                const int64_t n = p0 + t / out_dim;
                const int o = (int)(t % out_dim);
                const float *wr = wts_g + ph.pw.w + (size_t)o * K;
                float acc = wts_g[ph.pw.b + o];
                const float *xr = xc + n * s * in_dim;
                for (int kk = 0; kk < s * in_dim; ++kk) acc = fmaf(xr[kk], wr[kk], acc);
                for (int jj = 0; jj < d; ++jj) {
                    float c = 0.f;
                    for (int m = 0; m < s; ++m) c += pc[(n * s + m) * d + jj];
                    c = c / (float)s;
                    if (o == 0) P.ps[L + 1][n * d + jj] = c;
                    for (int m = 0; m < s; ++m) acc = fmaf(pc[(n * s + m) * d + jj] - c, wr[s * in_dim + m * d + jj], acc);
                }
                P.xs[L + 1][n * out_dim + o] = fmaf(acc, wts_g[ph.pw.scale + o], wts_g[ph.pw.shift + o]);
            }
        } else if (ph.type == PH_UNPOOL) {   // BallUnpooling (pooling.py:58-72)
            const int s = P.stride[L], in_dim = P.dim[L + 1], out_dim = P.dim[L];
            const int64_t n_par = P.npt[L + 1], p0 = (int64_t)tree * n_par;
            const int K = in_dim + s * d;
            const float *xp = P.xs[L + 1], *pp = P.ps[L + 1], *pc = P.ps[L];
            float *xc = P.xs[L];
            for (int64_t t = gt; t < n_par * s * out_dim; t += gthreads) {
                const int64_t tokc = p0 * s + t / out_dim;
                const int c = (int)(t % out_dim);
                const int64_t n = tokc / s;
                const int m = (int)(tokc - n * s);
                const int row = m * out_dim + c;
                const float *wr = wts_g + ph.pw.w + (size_t)row * K;
                float acc = wts_g[ph.pw.b + row];
                const float *xr = xp + n * in_dim;
                for (int kk = 0; kk < in_dim; ++kk) acc = fmaf(xr[kk], wr[kk], acc);
                for (int mm = 0; mm < s; ++mm)
                    for (int jj = 0; jj < d; ++jj)
                        acc = fmaf(pc[(n * s + mm) * d + jj] - pp[n * d + jj], wr[in_dim + mm * d + jj], acc);
                const int64_t e = tokc * out_dim + c;
                xc[e] = fmaf(xc[e] + acc, wts_g[ph.pw.scale + c], wts_g[ph.pw.shift + c]);
            }
        } else {   // PH_SCATTER: erwin_flash.py:193  out[tree_idx[i]] = x[i] where mask[i]
            const int dim = P.dim[0];
            const int64_t n = P.npt[0], r0 = (int64_t)tree * n;
            for (int64_t t = gt; t < n * dim; t += gthreads) {
                const int64_t i = r0 + t / dim;
                const int c = (int)(t % dim);
                if (tree_mask[i]) out[tree_idx[i] * dim + c] = __ldcg(P.xs[0] + i * dim + c);
            }
        }
        if (k + 1 < P.n_phases) {
            if (!early) stage_weights(P.ph[k + 1]);   // constants: safe before the barrier, lands while we wait
            ef_group_barrier(ctr, (unsigned int)(k + 1) * (unsigned int)P.group_ctas, P.timeout_ns);
        }
    }
}

// Launch with the programmatic-stream-serialization attribute: the kernel may be scheduled while its predecessor in the
// stream is still running; it executes griddepcontrol.wait before it reads anything the predecessor wrote, which blocks
// until the predecessor has completed and flushed its writes.  Used for the block kernels only, whose prologue (staging
// 42-166 KB of weights) is independent of the predecessor: Erwin 0.168 -> 0.155 ms.  On the small gather / pool / unpool /
// scatter kernels (nothing to do before the wait) it was slower (layer 2.41 -> 2.44 ms) and was dropped.
// Works under CUDA-graph capture (programmatic dependency edges).  HAET_NO_PDL=1 turns the attribute off.
template <class... KArgs, class... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = options().no_pdl ? 0 : 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- host walker: the packed-weight order (mirrored by ErwinTransformer.packed_weights in haet_b200/erwin.py) ------------
struct Walker {
    int64_t off = 0;
    int take(int64_t n) { int64_t o = off; off += (n + 3) / 4 * 4; return (int)o; }
    BlockW block(int dim, int hidden, int d, int bias_heads = 0) {   // bias_heads > 0: the block carries sigma_att[heads]
        BlockW b;
        b.norm1 = take(dim); b.norm2 = take(dim);
        b.pe_w = take((int64_t)dim * d); b.pe_b = take(dim);
        b.qkv_w = take((int64_t)3 * dim * dim); b.qkv_b = take(3 * dim);
        b.proj_w = take((int64_t)dim * dim); b.proj_b = take(dim);
        b.w1 = take((int64_t)hidden * dim); b.b1 = take(hidden);
        b.w2 = take((int64_t)hidden * dim); b.b2 = take(hidden);
        b.w3 = take((int64_t)dim * hidden); b.b3 = take(dim);
        b.sigma = bias_heads > 0 ? take(bias_heads) : -1;
        return b;
    }
    PoolW pool(int64_t rows, int64_t K, int bn) {
        PoolW p;
        p.w = take(rows * K); p.b = take(rows); p.scale = take(bn); p.shift = take(bn);
        return p;
    }
};

static int check_cfg(const haet_erwin_cfg *c) {
    HAET_CHECK_ARG(c->n_levels >= 1 && c->n_levels <= HAET_MAX_LEVELS, "erwin n_levels out of range");
    HAET_CHECK_ARG(c->dimensionality >= 1 && c->dimensionality <= 3, "erwin dimensionality must be 1..3");
    for (int i = 0; i < c->n_levels; ++i) {
        HAET_CHECK_ARG(c->c_hidden[i] > 0 && c->enc_num_heads[i] > 0 && c->c_hidden[i] % c->enc_num_heads[i] == 0,
                       "dim must be divisible by num_heads (level %d)", i);
        HAET_CHECK_ARG(c->ball_sizes[i] > 0 && c->enc_depths[i] >= 0, "bad ball size / depth at level %d", i);
    }
    for (int i = 0; i + 1 < c->n_levels; ++i) {
        HAET_CHECK_ARG(c->strides[i] >= 1, "bad stride at level %d", i);
        HAET_CHECK_ARG(c->dec_num_heads[i] > 0 && c->c_hidden[i] % c->dec_num_heads[i] == 0,
                       "dim must be divisible by num_heads (decoder %d)", i);
    }
    return HAET_OK;
}

}  // namespace haet

using namespace haet;

extern "C" int64_t haet_erwin_weight_floats(const haet_erwin_cfg *c) {
    if (check_cfg(c) != HAET_OK) return -1;
    Walker wk;
    const int d = c->dimensionality, L = c->n_levels - 1;
    for (int i = 0; i < L; ++i) {
        for (int k = 0; k < c->enc_depths[i]; ++k) wk.block(c->c_hidden[i], c->c_hidden[i] * c->mlp_ratio, d, c->attn_bias ? c->enc_num_heads[i] : 0);
        if (c->strides[i] != 1)  // stride 1: BallPooling is the identity (pooling.py:26-27); its params are unused
            wk.pool(c->c_hidden[i + 1], (int64_t)c->strides[i] * (c->c_hidden[i] + d), c->c_hidden[i + 1]);
    }
    for (int k = 0; k < c->enc_depths[L]; ++k) wk.block(c->c_hidden[L], c->c_hidden[L] * c->mlp_ratio, d, c->attn_bias ? c->enc_num_heads[L] : 0);
    if (c->decode)
        for (int i = L - 1; i >= 0; --i) {
            wk.pool((int64_t)c->strides[i] * c->c_hidden[i], c->c_hidden[i + 1] + (int64_t)c->strides[i] * d, c->c_hidden[i]);
            for (int k = 0; k < c->dec_depths[i]; ++k) wk.block(c->c_hidden[i], c->c_hidden[i] * c->mlp_ratio, d, c->attn_bias ? c->dec_num_heads[i] : 0);
        }
    return wk.off;
}

extern "C" int64_t haet_erwin_workspace_bytes(const haet_erwin_cfg *c, int64_t n_slots) {
    if (check_cfg(c) != HAET_OK) return -1;
    int64_t fl = 0, n = n_slots;
    for (int i = 0; i < c->n_levels; ++i) {
        fl += align_up(n * c->c_hidden[i], 4) + align_up(n * c->dimensionality, 4);
        if (i + 1 < c->n_levels) n /= c->strides[i];
    }
    return fl * (int64_t)sizeof(float) + EF_MAX_TREES * (int64_t)sizeof(unsigned int);   // + per-tree barrier counters (fused kernel)
}

static int run_blocks(const haet_erwin_cfg *c, float *x, const float *pos, int64_t n, int level, int depth, int heads,
                      const int64_t *rot, const float *wts, Walker &wk, cudaStream_t st) {
    const int dim = c->c_hidden[level], hidden = dim * c->mlp_ratio, bs = c->ball_sizes[level], d = c->dimensionality;
    HAET_CHECK_ARG(n % bs == 0, "level %d: %lld tokens are not a multiple of ball size %d", level, (long long)n, bs);
    // weights of one block (Walker::block: contiguous, every tensor padded to 4 floats) staged in shared memory if they fit
    int64_t w_n = 0;
    const int bias_heads = c->attn_bias ? heads : 0;
    { Walker probe; probe.block(dim, hidden, d, bias_heads); w_n = probe.off; }
    const size_t smem_act = block_smem(bs, dim, hidden) + 16;
    HAET_CHECK_ARG(smem_act <= 200 * 1024, "ball of %d tokens x dim %d does not fit shared memory", bs, dim);
    if (smem_act + (size_t)w_n * 4 > 220 * 1024 || ((dim | hidden) & 3)) w_n = 0;
    const size_t smem = smem_act + (size_t)w_n * 4;
    if (smem > 48 * 1024) HAET_DYN_SMEM(erwin_block_kernel, smem);
    const bool rotate = c->rotate > 0 && depth > 1;   // layers.py:74-91
    for (int k = 0; k < depth; ++k) {
        BlockW bw = wk.block(dim, hidden, d, bias_heads);
        const int64_t *perm = (rotate && (k & 1)) ? rot : nullptr;
        if (rotate && (k & 1)) HAET_CHECK_ARG(rot != nullptr, "tree_idx_rot must be provided for rotation");
        if (n == 0) continue;
        HAET_CUDA(launch_pdl(erwin_block_kernel, dim3((unsigned)(n / bs)), dim3(EB_THREADS), smem, st, x, pos, perm, wts, bw, bs, dim, heads,
                             hidden, d, (int)bw.norm1, (int)w_n));
        HAET_LAUNCHED();
    }
    return HAET_OK;
}

// One cooperative launch for the whole pass (erwin_fused_kernel).  Returns 1 when the shape / device cannot take it (the caller
// then uses the multi-launch path), a negative HAET_E* code on errors.
static int erwin_forward_fused(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t n_trees,
                               const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                               int64_t *const *rot_idx_d, const float *weights_d, void *workspace_d, float *out_d,
                               cudaStream_t st) {
    const int d = c->dimensionality, L = c->n_levels - 1;
    if (!options().erwin_fused || !c->decode || n_trees < 1 || n_trees > EF_MAX_TREES || n_slots % n_trees) return 1;
    if (c->bn_train) return 1;   // batch statistics couple the trees: the per-tree barriers do not cover them
    ErwinProgram P{};
    P.n_levels = c->n_levels; P.d = d; P.mlp_ratio = c->mlp_ratio; P.n_trees = (int)n_trees;
    P.timeout_ns = 5000000000ull;
    size_t act_max = 0;
    int64_t w_max = 0, group = 1;
    {
        float *p = (float *)workspace_d;
        int64_t n = n_slots;
        for (int i = 0; i < c->n_levels; ++i) {
            P.dim[i] = c->c_hidden[i]; P.bs[i] = c->ball_sizes[i]; P.stride[i] = i < L ? c->strides[i] : 1;
            P.npt[i] = n / n_trees;
            if (n % n_trees || P.npt[i] % P.bs[i]) return 1;           // balls must not straddle trees
            if (P.npt[i] / P.bs[i] > group) group = P.npt[i] / P.bs[i];
            P.xs[i] = p; p += align_up(n * c->c_hidden[i], 4); P.ps[i] = p; p += align_up(n * d, 4);
            P.rot[i] = (c->rotate > 0 && rot_idx_d) ? rot_idx_d[i] : nullptr;
            if (i < L) {
                if (c->strides[i] == 1 || n % c->strides[i] || (n / n_trees) % c->strides[i]) return 1;
                n /= c->strides[i];
            }
        }
    }
    P.group_ctas = (int)group;
    Walker wk;
    int np = 0;
    auto push = [&](int type, int level) -> ErwinPhase * {
        if (np >= EF_MAX_PHASES) return nullptr;
        ErwinPhase &ph = P.ph[np++];
        ph = ErwinPhase{};
        ph.type = type; ph.level = level;
        return &ph;
    };
    auto blocks = [&](int level, int depth, int heads) -> int {
        const int dim = c->c_hidden[level], hidden = dim * c->mlp_ratio, bs = c->ball_sizes[level];
        const bool rotate = c->rotate > 0 && depth > 1;
        const size_t act = block_smem(bs, dim, hidden);
        if (act > 200 * 1024) return 1;
        if (act > act_max) act_max = act;
        for (int k = 0; k < depth; ++k) {
            ErwinPhase *ph = push(PH_BLOCK, level);
            if (!ph) return 1;
            ph->bw = wk.block(dim, hidden, d, c->attn_bias ? heads : 0);
            ph->heads = heads;
            ph->use_rot = rotate && (k & 1);
            if (ph->use_rot && !P.rot[level]) return HAET_EINVAL;
            Walker probe; probe.block(dim, hidden, d, c->attn_bias ? heads : 0);
            ph->w_lo = ph->bw.norm1;
            ph->w_n = (act + (size_t)probe.off * 4 + 64 <= 220 * 1024 && !((dim | hidden) & 3)) ? (int)probe.off : 0;
            if (ph->w_n > w_max) w_max = ph->w_n;
        }
        return 0;
    };
    if (!push(PH_GATHER, 0)) return 1;
    int rc;
    for (int i = 0; i < L; ++i) {
        if ((rc = blocks(i, c->enc_depths[i], c->enc_num_heads[i]))) return rc;
        ErwinPhase *ph = push(PH_POOL, i);
        if (!ph) return 1;
        ph->pw = wk.pool(c->c_hidden[i + 1], (int64_t)c->strides[i] * (c->c_hidden[i] + d), c->c_hidden[i + 1]);
    }
    if ((rc = blocks(L, c->enc_depths[L], c->enc_num_heads[L]))) return rc;
    for (int i = L - 1; i >= 0; --i) {
        ErwinPhase *ph = push(PH_UNPOOL, i);
        if (!ph) return 1;
        ph->pw = wk.pool((int64_t)c->strides[i] * c->c_hidden[i], c->c_hidden[i + 1] + (int64_t)c->strides[i] * d, c->c_hidden[i]);
        if ((rc = blocks(i, c->dec_depths[i], c->dec_num_heads[i]))) return rc;
    }
    if (!push(PH_SCATTER, 0)) return 1;
    P.n_phases = np;
    // a staged block must fit next to the LARGEST staged weights (fixed layout: [barrier | weights | activations])
    P.act_off = (int)align_up(w_max, 4);
    const size_t smem = 16 + (size_t)P.act_off * 4 + act_max + 64;
    if (smem > 220 * 1024) return 1;
    if (smem > 48 * 1024) HAET_DYN_SMEM(erwin_fused_kernel, smem);
    int dev = 0, sms = 0, per_sm = 0, coop = 0;
    HAET_CUDA(cudaGetDevice(&dev));
    HAET_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    HAET_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    HAET_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, erwin_fused_kernel, EB_THREADS, smem));
    const int64_t grid = n_trees * group;
    if (!coop || grid > (int64_t)per_sm * sms) return 1;   // the tree barriers need every CTA resident
    unsigned int *counters = reinterpret_cast<unsigned int *>((char *)workspace_d + haet_erwin_workspace_bytes(c, n_slots) -
                                                               EF_MAX_TREES * (int64_t)sizeof(unsigned int));
    HAET_CUDA(cudaMemsetAsync(counters, 0, (size_t)n_trees * sizeof(unsigned int), st));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(EB_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    HAET_CUDA(cudaLaunchKernelEx(&cfg, erwin_fused_kernel, x_d, pos_d, tree_idx_d, tree_mask_d, weights_d, out_d, counters, P));
    HAET_LAUNCHED();
    return HAET_OK;
}

static int erwin_forward_multi(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T,
                               const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                               int64_t *const *rot_idx_d, const float *weights_d, void *workspace_d,
                               float *out_d, void *stream);

extern "C" int haet_erwin_forward_trees(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T, int64_t n_trees,
                                        const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                                        int64_t *const *rot_idx_d, const float *weights_d, void *workspace_d,
                                        float *out_d, void *stream) {
    int rc = check_cfg(c);
    if (rc) return rc;
    if (n_slots == 0) return HAET_OK;
    HAET_CHECK_ARG(n_trees >= 1, "haet_erwin_forward_trees: n_trees must be >= 1");
    rc = erwin_forward_fused(c, x_d, pos_d, n_trees, tree_idx_d, tree_mask_d, n_slots, rot_idx_d, weights_d, workspace_d, out_d,
                             (cudaStream_t)stream);
    if (rc <= 0) return rc;
    return erwin_forward_multi(c, x_d, pos_d, T, tree_idx_d, tree_mask_d, n_slots, rot_idx_d, weights_d, workspace_d, out_d, stream);
}

// Without the tree count every CTA belongs to one group (one barrier domain): still one launch, coarser barriers.
extern "C" int haet_erwin_forward(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T,
                                  const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                                  int64_t *const *rot_idx_d, const float *weights_d, void *workspace_d,
                                  float *out_d, void *stream) {
    return haet_erwin_forward_trees(c, x_d, pos_d, T, 1, tree_idx_d, tree_mask_d, n_slots, rot_idx_d, weights_d, workspace_d, out_d, stream);
}

static int erwin_forward_multi(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T,
                               const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                               int64_t *const *rot_idx_d, const float *weights_d, void *workspace_d,
                               float *out_d, void *stream) {
    int rc = check_cfg(c);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int d = c->dimensionality, L = c->n_levels - 1;
    float *xs[HAET_MAX_LEVELS], *ps[HAET_MAX_LEVELS];
    int64_t ns[HAET_MAX_LEVELS];
    {
        float *p = (float *)workspace_d;
        int64_t n = n_slots;
        for (int i = 0; i < c->n_levels; ++i) {
            ns[i] = n; xs[i] = p; p += align_up(n * c->c_hidden[i], 4); ps[i] = p; p += align_up(n * d, 4);
            if (i < L) {
                HAET_CHECK_ARG(n % c->strides[i] == 0, "stride %d does not divide %lld tokens", c->strides[i], (long long)n);
                n /= c->strides[i];
            }
        }
    }
    if (n_slots == 0) return HAET_OK;
    const int dim0 = c->c_hidden[0];
    {
        const int64_t tot = n_slots * (dim0 + d);
        erwin_gather_kernel<<<(unsigned)cdiv(tot, 256), 256, 0, st>>>(x_d, pos_d, tree_idx_d, n_slots, dim0, d, xs[0], ps[0]);
        HAET_LAUNCHED();
    }
    Walker wk;
    auto rot = [&](int i) -> const int64_t * { return (c->rotate > 0 && rot_idx_d) ? rot_idx_d[i] : nullptr; };
    // train-mode BatchNorm (haet_erwin_cfg.bn_train): the pool / unpool kernels leave the pre-norm values, a per-channel kernel
    // normalises with the batch statistics and reports them at bn_stats_d + 2 * (channels of the earlier BatchNorms)
    const int bn_train = c->bn_train ? 1 : 0;
    float *bn_stats = c->bn_stats_d;
    int64_t bn_off = 0;
    HAET_CHECK_ARG(!bn_train || c->bn_eps > 0.f, "bn_train needs bn_eps > 0");
    for (int i = 0; i < L; ++i) {   // encoder (erwin_flash.py:183-185)
        rc = run_blocks(c, xs[i], ps[i], ns[i], i, c->enc_depths[i], c->enc_num_heads[i], rot(i), weights_d, wk, st);
        if (rc) return rc;
        const int s = c->strides[i];
        if (s != 1) {
            PoolW pw = wk.pool(c->c_hidden[i + 1], (int64_t)s * (c->c_hidden[i] + d), c->c_hidden[i + 1]);
            const int64_t tot = ns[i + 1] * c->c_hidden[i + 1];
            erwin_pool_kernel<<<(unsigned)cdiv(tot, 128), 128, 0, st>>>(xs[i], ps[i], ns[i + 1], s, c->c_hidden[i],
                                                                       c->c_hidden[i + 1], d, weights_d, pw, xs[i + 1], ps[i + 1], bn_train);
            HAET_LAUNCHED();
            if (bn_train && ns[i + 1] > 0) {
                erwin_bn_train_kernel<<<(unsigned)c->c_hidden[i + 1], 256, 0, st>>>(xs[i + 1], ns[i + 1], c->c_hidden[i + 1], weights_d, pw,
                                                                                   c->bn_eps, bn_stats ? bn_stats + bn_off : nullptr);
                HAET_LAUNCHED();
            }
            bn_off += 2 * c->c_hidden[i + 1];
        } else {
            HAET_CHECK_ARG(c->c_hidden[i] == c->c_hidden[i + 1], "stride 1 needs equal dims");
            HAET_CUDA(cudaMemcpyAsync(xs[i + 1], xs[i], sizeof(float) * ns[i] * c->c_hidden[i], cudaMemcpyDeviceToDevice, st));
            HAET_CUDA(cudaMemcpyAsync(ps[i + 1], ps[i], sizeof(float) * ns[i] * d, cudaMemcpyDeviceToDevice, st));
        }
    }
    rc = run_blocks(c, xs[L], ps[L], ns[L], L, c->enc_depths[L], c->enc_num_heads[L], rot(L), weights_d, wk, st);
    if (rc) return rc;
    HAET_CHECK_ARG(c->decode, "decode=False is not used by the HAET layer (the deslice needs token-level states)");
    for (int i = L - 1; i >= 0; --i) {   // decoder (erwin_flash.py:190-192)
        const int s = c->strides[i];
        PoolW pw = wk.pool((int64_t)s * c->c_hidden[i], c->c_hidden[i + 1] + (int64_t)s * d, c->c_hidden[i]);
        const int64_t tot = ns[i + 1] * s * c->c_hidden[i];
        erwin_unpool_kernel<<<(unsigned)cdiv(tot, 128), 128, 0, st>>>(xs[i + 1], ps[i + 1], ns[i + 1], s, c->c_hidden[i + 1],
                                                                     c->c_hidden[i], d, weights_d, pw, xs[i], ps[i], bn_train);
        HAET_LAUNCHED();
        if (bn_train && ns[i] > 0) {
            erwin_bn_train_kernel<<<(unsigned)c->c_hidden[i], 256, 0, st>>>(xs[i], ns[i], c->c_hidden[i], weights_d, pw, c->bn_eps,
                                                                           bn_stats ? bn_stats + bn_off : nullptr);
            HAET_LAUNCHED();
        }
        bn_off += 2 * c->c_hidden[i];
        rc = run_blocks(c, xs[i], ps[i], ns[i], i, c->dec_depths[i], c->dec_num_heads[i], rot(i), weights_d, wk, st);
        if (rc) return rc;
    }
    (void)T;
    erwin_scatter_kernel<<<(unsigned)cdiv(n_slots * dim0, 256), 256, 0, st>>>(xs[0], tree_idx_d, tree_mask_d, n_slots, dim0, out_d);
    HAET_LAUNCHED();
    return HAET_OK;
}

// ---- allocation-free variant (no counterpart in the reference) ----------------
// Callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get
// the Erwin pass with its scratch in a file-scope device array instead of a caller-provided workspace.  Every GPU that
// loads the library holds these HAET_ERWIN_STATIC_SCRATCH_BYTES until exit.  In HOST code `g_erwin_static_scratch` is only a
// stand-in for the symbol, not a device address: the address is taken with cudaGetSymbolAddress, once per device, with that
// device current (cached below), and that pointer is what the kernels receive.
static __device__ __align__(256) unsigned char g_erwin_static_scratch[HAET_ERWIN_STATIC_SCRATCH_BYTES];

static int erwin_static_scratch(void **out) {
    static std::atomic<void *> cache[64];
    int dev = 0;
    HAET_CUDA(cudaGetDevice(&dev));
    HAET_CHECK_ARG(dev >= 0 && dev < 64, "haet_erwin_forward_static: device ordinal %d is beyond the 64 this library tracks", dev);
    void *p = cache[dev].load(std::memory_order_acquire);
    if (!p) {   // a race between two threads only repeats the (idempotent) query
        HAET_CUDA(cudaGetSymbolAddress(&p, g_erwin_static_scratch));
        HAET_CHECK_ARG(p != nullptr && ((uintptr_t)p & 255) == 0, "haet_erwin_forward_static: bad scratch symbol address");
        cache[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return HAET_OK;
}

extern "C" int haet_erwin_forward_static(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T,
                                         int64_t n_trees, const int64_t *tree_idx_d, const uint8_t *tree_mask_d,
                                         int64_t n_slots, int64_t *const *rot_idx_d, const float *weights_d, float *out_d,
                                         void *stream) {
    int rc = check_cfg(c);
    if (rc) return rc;
    HAET_CHECK_ARG(n_slots >= 0 && n_trees >= 1, "haet_erwin_forward_static: n_slots must be >= 0 and n_trees >= 1");
    const int64_t need = haet_erwin_workspace_bytes(c, n_slots);
    if (need < 0 || need > (int64_t)HAET_ERWIN_STATIC_SCRATCH_BYTES) {
        set_error("haet_erwin_forward_static: %lld tree slots of this configuration need %lld bytes of scratch, the built-in array "
                  "holds %lld (HAET_ERWIN_STATIC_SCRATCH_BYTES); use haet_erwin_forward_trees with a caller workspace",
                  (long long)n_slots, (long long)need, (long long)HAET_ERWIN_STATIC_SCRATCH_BYTES);
        return HAET_ENOSUP;
    }
    if (n_slots == 0) return HAET_OK;
    void *scratch = nullptr;
    if ((rc = erwin_static_scratch(&scratch))) return rc;
    // The array keeps the previous call's contents.  The activations of every level are fully written (gather / pool / unpool)
    // before they are read, so only the per-tree barrier counters of the one-launch kernel must start at zero: they sit at the
    // end of the workspace layout and are cleared on the stream here, whichever path runs.
    HAET_CUDA(cudaMemsetAsync((char *)scratch + need - EF_MAX_TREES * (int64_t)sizeof(unsigned int), 0,
                              EF_MAX_TREES * sizeof(unsigned int), (cudaStream_t)stream));
    return haet_erwin_forward_trees(c, x_d, pos_d, T, n_trees, tree_idx_d, tree_mask_d, n_slots, rot_idx_d, weights_d, scratch,
                                    out_d, stream);
}
