// fp32 SIMT tiled GEMM: out[M, Nn] = A[M, K] . W[Nn, K]^T + bias.
// A is either a dense row-major matrix (Linear; IrregularMesh.py:157, :235) or the implicit
// im2col view of a channels-last grid (Conv2d/Conv3d k x k(x k), stride 1, pad k/2;
// StructuredMesh2D.py:173-189, StructuredMesh3D.py:177-193).  This is the exact-fp32 path
// (the device-side cross-check of the tensor-core kernels and the path for odd shapes).
#include "common.cuh"

namespace haet {

struct AView {
    const float *x;
    int64_t M;      // rows (B*N)
    int K;          // reduction length (dim or taps*dim)
    int conv_nd;    // 0 dense
    int din;        // channels per tap
    int g0, g1, g2; // grid extents (g2 = 1 for 2D)
    int ks;         // kernel size
    int64_t N;      // points per batch item
};

__device__ __forceinline__ float4 load_a4(const AView &a, int64_t m, int k) {
    float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m >= a.M || k >= a.K) return z;
    if (a.conv_nd == 0) {
        const float *p = a.x + m * (int64_t)a.K + k;
        if (k + 3 < a.K && (a.K & 3) == 0) return *reinterpret_cast<const float4 *>(p);
        z.x = p[0];
        if (k + 1 < a.K) z.y = p[1];
        if (k + 2 < a.K) z.z = p[2];
        if (k + 3 < a.K) z.w = p[3];
        return z;
    }
    // implicit im2col: k = tap*din + ci (din % 4 == 0 is checked by the host)
    int tap = k / a.din, ci = k - tap * a.din;
    int64_t b = m / a.N, n = m - b * a.N;
    int pad = a.ks / 2;
    int64_t nn;
    if (a.conv_nd == 2) {
        int i = (int)(n / a.g1), j = (int)(n - (int64_t)i * a.g1);
        int di = tap / a.ks - pad, dj = tap % a.ks - pad;
        int ii = i + di, jj = j + dj;
        if (ii < 0 || ii >= a.g0 || jj < 0 || jj >= a.g1) return z;
        nn = (int64_t)ii * a.g1 + jj;
    } else {
        int64_t ij = n / a.g2;
        int kz = (int)(n - ij * a.g2);
        int i = (int)(ij / a.g1), j = (int)(ij - (int64_t)i * a.g1);
        int t2 = a.ks * a.ks;
        int di = tap / t2 - pad, dj = (tap / a.ks) % a.ks - pad, dk = tap % a.ks - pad;
        int ii = i + di, jj = j + dj, kk = kz + dk;
        if (ii < 0 || ii >= a.g0 || jj < 0 || jj >= a.g1 || kk < 0 || kk >= a.g2) return z;
        nn = ((int64_t)ii * a.g1 + jj) * a.g2 + kk;
    }
    return *reinterpret_cast<const float4 *>(a.x + (b * a.N + nn) * (int64_t)a.din + ci);
}

constexpr int BM = 64, BN = 64, BK = 16;

__global__ void __launch_bounds__(256) gemm_simt_kernel(AView a, const float *__restrict__ w,
                                                        const float *__restrict__ bias,
                                                        float *__restrict__ out, int Nn) {
    __shared__ float As[BK][BM + 4];
    __shared__ float Ws[BK][BN + 4];
    const int tid = threadIdx.x;
    const int64_t m0 = (int64_t)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;
    const int lr = tid >> 2, lk = (tid & 3) * 4;   // loader: row 0..63, k offset 0,4,8,12
    const int tx = tid & 15, ty = tid >> 4;        // compute: 16 x 16 threads, 4 x 4 each
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < a.K; k0 += BK) {
        float4 av = load_a4(a, m0 + lr, k0 + lk);
        float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
        {
            int o = n0 + lr, k = k0 + lk;
            if (o < Nn && k < a.K) {
                const float *p = w + (int64_t)o * a.K + k;
                if (k + 3 < a.K && (a.K & 3) == 0) wv = *reinterpret_cast<const float4 *>(p);
                else {
                    wv.x = p[0];
                    if (k + 1 < a.K) wv.y = p[1];
                    if (k + 2 < a.K) wv.z = p[2];
                    if (k + 3 < a.K) wv.w = p[3];
                }
            }
        }
        __syncthreads();
        As[lk + 0][lr] = av.x; As[lk + 1][lr] = av.y; As[lk + 2][lr] = av.z; As[lk + 3][lr] = av.w;
        Ws[lk + 0][lr] = wv.x; Ws[lk + 1][lr] = wv.y; Ws[lk + 2][lr] = wv.z; Ws[lk + 3][lr] = wv.w;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float4 af = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
            float4 wf = *reinterpret_cast<const float4 *>(&Ws[k][tx * 4]);
            float ar[4] = {af.x, af.y, af.z, af.w}, wr[4] = {wf.x, wf.y, wf.z, wf.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], wr[j], acc[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int64_t m = m0 + ty * 4 + i;
        if (m >= a.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int o = n0 + tx * 4 + j;
            if (o < Nn) out[m * (int64_t)Nn + o] = acc[i][j] + (bias ? bias[o] : 0.f);
        }
    }
}

static int launch_gemm(const AView &a, const float *w, const float *bias, float *out, int Nn, cudaStream_t st) {
    if (a.M == 0) return HAET_OK;
    dim3 grid((unsigned)cdiv(a.M, BM), (unsigned)cdiv(Nn, BN));
    gemm_simt_kernel<<<grid, 256, 0, st>>>(a, w, bias, out, Nn);
    HAET_LAUNCHED();
    return HAET_OK;
}

int launch_linear(const float *a, const float *w, const float *bias, float *out, int64_t M, int K, int Nn,
                  cudaStream_t st) {
    AView v{a, M, K, 0, K, 1, 1, 1, 1, M};
    return launch_gemm(v, w, bias, out, Nn, st);
}

int launch_in_project(const haet_layer_shape *s, const float *x, const float *wx, const float *bx, float *xp,
                      cudaStream_t st) {
    const int inner = s->heads * s->dim_head;
    AView v;
    v.x = x; v.M = s->B * s->N; v.conv_nd = s->conv_nd; v.din = s->dim; v.N = s->N; v.ks = s->kernel;
    if (s->conv_nd == 0) {
        v.K = s->dim; v.g0 = v.g1 = v.g2 = 1;
    } else {
        HAET_CHECK_ARG(s->conv_nd == 2 || s->conv_nd == 3, "conv_nd must be 0, 2 or 3");
        HAET_CHECK_ARG(s->dim % 4 == 0, "conv front-end needs dim %% 4 == 0 (got %d)", s->dim);
        HAET_CHECK_ARG(s->kernel % 2 == 1, "conv kernel must be odd");
        v.g0 = s->grid[0]; v.g1 = s->grid[1]; v.g2 = s->conv_nd == 3 ? s->grid[2] : 1;
        HAET_CHECK_ARG((int64_t)v.g0 * v.g1 * v.g2 == s->N, "grid %dx%dx%d does not match N=%lld", v.g0, v.g1,
                       v.g2, (long long)s->N);
        int taps = s->kernel * s->kernel * (s->conv_nd == 3 ? s->kernel : 1);
        v.K = taps * s->dim;
    }
    return launch_gemm(v, wx, bx, xp, inner, st);
}

}  // namespace haet
