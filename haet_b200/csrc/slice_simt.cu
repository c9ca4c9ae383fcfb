// fp32 SIMT slicing / deslicing kernels (exact-fp32 path; see DESIGN.md "kernel inventory").
//
// The Rep-Slice softmax runs over the N points (IrregularMesh.py:174), so the layer is
// two passes over the points:
//   pass 1  slice_partial: per (chunk of points, b, h): running column max m2, l = sum 2^(z2-m2),
//           S = sum 2^(z2-m2) x_proj in one sweep (online softmax along n, thread = slice g)
//           slice_merge:   fixed-order log-sum-exp merge of the chunk (and rank) partials
//   pass 2  deslice: W[n,g] = 2^(z2 - lse2[g]) recomputed from x_proj, out_h = W . S'_h
// z2 = ((x_proj . Ws_g + bs_g - log(-log eps)) / tau_n) * log2(e)   (IrregularMesh.py:168-174)
#include "common.cuh"

namespace haet {

constexpr int kChunk = 2048;  // points per pass-1 CTA
constexpr int kTile = 32;     // points staged in shared memory at a time
constexpr int kSub = 4;       // points per online-softmax update (bounds register use)

int64_t slice_chunks(const haet_layer_shape *s) { return cdiv(s->N, kChunk); }

template <int C>
__global__ void __launch_bounds__(256) slice_partial_kernel(
    const float *__restrict__ xp, const float *__restrict__ ws, const float *__restrict__ bs,
    const float *__restrict__ wt, float bt, float base_temp, float shift, int64_t N, int H, int G,
    float *__restrict__ parts) {
    __shared__ __align__(16) float xs[kTile][C];
    __shared__ float its[kTile];
    __shared__ float wts[C];
    const int tid = threadIdx.x;
    const int bh = blockIdx.y, b = bh / H, h = bh - b * H;
    const int inner = H * C;
    const int64_t n_begin = (int64_t)blockIdx.x * kChunk;
    const int64_t n_end = min(N, n_begin + (int64_t)kChunk);
    const int64_t BH = gridDim.y;
    if (tid < C) wts[tid] = wt[tid];

    for (int gg = 0; gg < G; gg += 256) {
        const int g = gg + tid;
        const bool active = g < G;
        float w[C], S[C];
#pragma unroll
        for (int k = 0; k < C; ++k) { w[k] = active ? ws[(int64_t)g * C + k] : 0.f; S[k] = 0.f; }
        const float bc = active ? bs[g] - shift : 0.f;
        float m = -INFINITY, l = 0.f;

        for (int64_t n0 = n_begin; n0 < n_end; n0 += kTile) {
            __syncthreads();
            for (int idx = tid; idx < kTile * C; idx += 256) {
                int p = idx / C, k = idx - p * C;
                int64_t n = n0 + p;
                xs[p][k] = n < n_end ? xp[((int64_t)b * N + n) * inner + h * C + k] : 0.f;
            }
            __syncthreads();
            if (tid < kTile) {
                float d = bt;
#pragma unroll
                for (int k = 0; k < C; ++k) d = fmaf(xs[tid][k], wts[k], d);
                float tau = base_temp + fminf(fmaxf(d, -0.4f), 0.4f);   // Ada-Temp, :168
                its[tid] = kLog2e / tau;
            }
            __syncthreads();
            if (!active) continue;
            for (int p0 = 0; p0 < kTile; p0 += kSub) {   // online softmax along n, kSub points at a time
                float z[kSub];
                float tmax = -INFINITY;
#pragma unroll
                for (int q = 0; q < kSub; ++q) {
                    const int p = p0 + q;
                    float d = 0.f;
#pragma unroll
                    for (int k = 0; k < C; k += 4) {
                        float4 v = *reinterpret_cast<const float4 *>(&xs[p][k]);
                        d = fmaf(v.x, w[k], d); d = fmaf(v.y, w[k + 1], d);
                        d = fmaf(v.z, w[k + 2], d); d = fmaf(v.w, w[k + 3], d);
                    }
                    z[q] = (n0 + p < n_end) ? (d + bc) * its[p] : -INFINITY;
                    tmax = fmaxf(tmax, z[q]);
                }
                if (tmax > m) {
                    float sc = exp2f(m - tmax);   // m = -inf on the first tile -> 0
                    l *= sc;
#pragma unroll
                    for (int k = 0; k < C; ++k) S[k] *= sc;
                    m = tmax;
                }
#pragma unroll
                for (int q = 0; q < kSub; ++q) {
                    const int p = p0 + q;
                    float e = exp2f(z[q] - m);
                    l += e;
#pragma unroll
                    for (int k = 0; k < C; k += 4) {
                        float4 v = *reinterpret_cast<const float4 *>(&xs[p][k]);
                        S[k] = fmaf(e, v.x, S[k]); S[k + 1] = fmaf(e, v.y, S[k + 1]);
                        S[k + 2] = fmaf(e, v.z, S[k + 2]); S[k + 3] = fmaf(e, v.w, S[k + 3]);
                    }
                }
            }
        }
        if (active) {
            float *o = parts + (((int64_t)blockIdx.x * BH + bh) * G + g) * (C + 2);
            o[0] = m; o[1] = l;
#pragma unroll
            for (int k = 0; k < C; ++k) o[2 + k] = S[k];
        }
    }
}

int launch_slice_partial(const haet_layer_shape *s, const float *xp, const float *ws, const float *bs,
                         const float *wt, float bt, float *chunk_parts, cudaStream_t st) {
    dim3 grid((unsigned)slice_chunks(s), (unsigned)(s->B * s->heads));
    const float shift = rep_slice_shift(s->epsilon);
#define HAET_SP(CC)                                                                                   \
    slice_partial_kernel<CC><<<grid, 256, 0, st>>>(xp, ws, bs, wt, bt, s->base_temp, shift, s->N,     \
                                                   s->heads, s->slice_num, chunk_parts)
    switch (s->dim_head) {
        case 4: HAET_SP(4); break;
        case 8: HAET_SP(8); break;
        case 16: HAET_SP(16); break;
        case 32: HAET_SP(32); break;
        case 64: HAET_SP(64); break;
        default: set_error("dim_head %d not supported (4, 8, 16, 32, 64)", s->dim_head); return HAET_ENOSUP;
    }
#undef HAET_SP
    HAET_LAUNCHED();
    return HAET_OK;
}

// One warp per slice row (b,h,g).  parts: [n_parts][rows][C+2].  Fixed part order => the result is
// bit-identical on every rank and from run to run.
__global__ void slice_merge_kernel(const float *__restrict__ parts, int64_t n_parts, int64_t rows, int C,
                                   float *__restrict__ packed_out, float *__restrict__ eidetic,
                                   float *__restrict__ lse2) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int W = C + 2;
    float m = -INFINITY;
    for (int64_t p = lane; p < n_parts; p += 32) m = fmaxf(m, parts[(p * rows + row) * W]);
#pragma unroll
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float l = 0.f, s0 = 0.f, s1 = 0.f;
    for (int64_t p = 0; p < n_parts; ++p) {
        const float *q = parts + (p * rows + row) * W;
        float mp = q[0];
        float sc = (mp == -INFINITY) ? 0.f : exp2f(mp - m);
        l = fmaf(q[1], sc, l);
        if (lane < C) s0 = fmaf(q[2 + lane], sc, s0);
        if (lane + 32 < C) s1 = fmaf(q[2 + lane + 32], sc, s1);
    }
    if (packed_out) {
        float *o = packed_out + row * W;
        if (lane == 0) { o[0] = m; o[1] = l; }
        if (lane < C) o[2 + lane] = s0;
        if (lane + 32 < C) o[2 + lane + 32] = s1;
    }
    if (eidetic) {
        // S / (sum_n W + 1e-5) with sum_n W == l/l == 1 (IrregularMesh.py:177-188)
        const float inv = 1.0f / (l * (1.0f + 1e-5f));
        if (lane < C) eidetic[row * C + lane] = s0 * inv;
        if (lane + 32 < C) eidetic[row * C + lane + 32] = s1 * inv;
        if (lane == 0) lse2[row] = m + log2f(l);
    }
}

int launch_slice_merge(const haet_layer_shape *s, const float *parts, int64_t n_parts, float *packed_out,
                       float *eidetic, float *lse2, cudaStream_t st) {
    const int64_t rows = s->B * s->heads * s->slice_num;
    HAET_CHECK_ARG(s->dim_head <= 64, "dim_head > 64 not supported");
    slice_merge_kernel<<<(unsigned)cdiv(rows, 8), 256, 0, st>>>(parts, n_parts, rows, s->dim_head, packed_out,
                                                                 eidetic, lse2);
    HAET_LAUNCHED();
    return HAET_OK;
}

// pos[t, j] = (S[t, j] - min_t S[t, j]) / (max_t S[t, j] - min_t S[t, j] + 1e-8), j < d
// (IrregularMesh.py:204-209; min/max over ALL B*H*G rows jointly).  One CTA: T <= a few 10^4 rows.
__global__ void __launch_bounds__(1024) slice_pos_kernel(const float *__restrict__ eidetic, int64_t T, int C,
                                                         int d, float *__restrict__ pos) {
    __shared__ float smin[3][32], smax[3][32];
    __shared__ float fmin_[3], frng_[3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t t = tid; t < T; t += blockDim.x)
        for (int j = 0; j < d; ++j) {
            float v = eidetic[t * C + j];
            mn[j] = fminf(mn[j], v); mx[j] = fmaxf(mx[j], v);
        }
    for (int j = 0; j < d; ++j) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            mn[j] = fminf(mn[j], __shfl_xor_sync(0xffffffffu, mn[j], o));
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], o));
        }
        if (lane == 0) { smin[j][warp] = mn[j]; smax[j][warp] = mx[j]; }
    }
    __syncthreads();
    if (tid < d) {
        float a = INFINITY, b = -INFINITY;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a = fminf(a, smin[tid][w]); b = fmaxf(b, smax[tid][w]); }
        fmin_[tid] = a;
        frng_[tid] = __fadd_rn(__fsub_rn(b, a), 1e-8f);
    }
    __syncthreads();
    for (int64_t i = tid; i < T * d; i += blockDim.x) {
        int64_t t = i / d; int j = (int)(i - t * d);
        pos[i] = __fdiv_rn(__fsub_rn(eidetic[t * C + j], fmin_[j]), frng_[j]);
    }
}

int launch_slice_pos(const haet_layer_shape *s, const float *eidetic, float *pos, cudaStream_t st) {
    HAET_CHECK_ARG(s->dimensionality >= 1 && s->dimensionality <= 3 && s->dimensionality <= s->dim_head,
                   "dimensionality must be 1..3 and <= dim_head");
    slice_pos_kernel<<<1, 1024, 0, st>>>(eidetic, s->B * s->heads * s->slice_num, s->dim_head, s->dimensionality,
                                         pos);
    HAET_LAUNCHED();
    return HAET_OK;
}

// Pass 2: o[b, n, h*C + c] = sum_g 2^(z2[n,h,g] - lse2[b,h,g]) * S'[b,h,g,c]   (IrregularMesh.py:228-232)
// CTA = 32 points (lane = point), warps stride over the heads; Ws / S' reads are warp-uniform.
template <int C>
__global__ void __launch_bounds__(256) deslice_kernel(
    const float *__restrict__ xp, const float *__restrict__ lse2, const float *__restrict__ states,
    const float *__restrict__ ws, const float *__restrict__ bs, const float *__restrict__ wt, float bt,
    float base_temp, float shift, int64_t N, int H, int G, float *__restrict__ o) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.y;
    const int64_t n = (int64_t)blockIdx.x * 32 + lane;
    const bool valid = n < N;
    const int inner = H * C;
    for (int h = warp; h < H; h += 8) {
        float xv[C], acc[C];
        const float *xr = xp + ((int64_t)b * N + (valid ? n : 0)) * inner + h * C;
        float d = bt;
#pragma unroll
        for (int k = 0; k < C; k += 4) {
            float4 v = *reinterpret_cast<const float4 *>(xr + k);
            xv[k] = v.x; xv[k + 1] = v.y; xv[k + 2] = v.z; xv[k + 3] = v.w;
        }
#pragma unroll
        for (int k = 0; k < C; ++k) { d = fmaf(xv[k], wt[k], d); acc[k] = 0.f; }
        const float it = kLog2e / (base_temp + fminf(fmaxf(d, -0.4f), 0.4f));
        const int64_t row0 = ((int64_t)b * H + h) * G;
        for (int g = 0; g < G; ++g) {
            const float *wr = ws + (int64_t)g * C;
            float dot = 0.f;
#pragma unroll
            for (int k = 0; k < C; k += 4) {
                float4 v = __ldg(reinterpret_cast<const float4 *>(wr + k));
                dot = fmaf(xv[k], v.x, dot); dot = fmaf(xv[k + 1], v.y, dot);
                dot = fmaf(xv[k + 2], v.z, dot); dot = fmaf(xv[k + 3], v.w, dot);
            }
            const float wgt = exp2f((dot + (__ldg(bs + g) - shift)) * it - __ldg(lse2 + row0 + g));
            const float *sr = states + (row0 + g) * C;
#pragma unroll
            for (int k = 0; k < C; k += 4) {
                float4 v = __ldg(reinterpret_cast<const float4 *>(sr + k));
                acc[k] = fmaf(wgt, v.x, acc[k]); acc[k + 1] = fmaf(wgt, v.y, acc[k + 1]);
                acc[k + 2] = fmaf(wgt, v.z, acc[k + 2]); acc[k + 3] = fmaf(wgt, v.w, acc[k + 3]);
            }
        }
        if (valid) {
            float *orow = o + ((int64_t)b * N + n) * inner + h * C;
#pragma unroll
            for (int k = 0; k < C; k += 4)
                *reinterpret_cast<float4 *>(orow + k) = make_float4(acc[k], acc[k + 1], acc[k + 2], acc[k + 3]);
        }
    }
}

int launch_deslice(const haet_layer_shape *s, const float *xp, const float *lse2, const float *states,
                   const float *ws, const float *bs, const float *wt, float bt, float *o, cudaStream_t st) {
    if (s->N == 0 || s->B == 0) return HAET_OK;
    dim3 grid((unsigned)cdiv(s->N, 32), (unsigned)s->B);
    const float shift = rep_slice_shift(s->epsilon);
#define HAET_DS(CC)                                                                                       \
    deslice_kernel<CC><<<grid, 256, 0, st>>>(xp, lse2, states, ws, bs, wt, bt, s->base_temp, shift, s->N, \
                                             s->heads, s->slice_num, o)
    switch (s->dim_head) {
        case 4: HAET_DS(4); break;
        case 8: HAET_DS(8); break;
        case 16: HAET_DS(16); break;
        case 32: HAET_DS(32); break;
        case 64: HAET_DS(64); break;
        default: set_error("dim_head %d not supported (4, 8, 16, 32, 64)", s->dim_head); return HAET_ENOSUP;
    }
#undef HAET_DS
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet
