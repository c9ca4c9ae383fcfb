// The per-point spill of the tensor-core path (layout in common.cuh) and the SIMT packer that builds it from
// fp32 projected points (used behind the conv front-ends; the Linear front-end writes the image straight from
// the in-projection GEMM's epilogue, gemm_tc.cu).
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc.cuh"

namespace haet {

bool spill_needs_xp(const haet_layer_shape *s) {
    const int inner = s->heads * s->dim_head;
    return hooks().path != 0 || s->conv_nd != 0 || !slice_tc_supported(s) || !deslice_tc_supported(s) ||
           !gemm_tc_supported(s->dim, inner) || s->dim_head == 64;   // dim_head 64: the image is packed from fp32 (Ada-Temp spans both sub-rows)
}

static bool spill_has_image(const haet_layer_shape *s) { return slice_tc_supported(s) && deslice_tc_supported(s); }

SpillView spill_view(const haet_layer_shape *s, void *base) {
    SpillView v;
    v.npad = xpi_npad(s->N);
    // dim_head 64: two pseudo-heads of 32 channels per head (slice_tc.cu, "pair mode")
    const int64_t rows = spill_has_image(s) ? s->B * s->heads * (s->dim_head == 64 ? 2 : 1) * v.npad : 0;
    char *p = (char *)base;
    v.img = (uint8_t *)p;
    p += align_up(rows * XPI_ROW_BYTES, 256);
    v.scale = (float *)p;
    p += align_up(rows * (int64_t)sizeof(float), 256);
    v.xp = (float *)p;
    return v;
}

// one thread per (b, h, row): C (16 or 32) projected channels -> hi/lo fp16 image row + s_n; with C == 16 the chunks of the
// missing channels 16 .. 31 are zero (the passes run their dim_head-32 kernels on zero-padded heads).  C == 64: H counts the
// PSEUDO-heads (two per head), row (b, 2h + sub) carries channels [32 sub, 32 sub + 32) of head h and the head's s_n
__global__ void __launch_bounds__(256) xpi_pack_kernel(const float *__restrict__ xp, const float *__restrict__ wt, float bt,
                                                       float base_temp, int64_t N, int64_t npad, int H, int C, int64_t rows,
                                                       uint8_t *__restrict__ img, float *__restrict__ scale, int64_t Nxp) {
    // Nxp: rows per batch item of xp (== N, or more when xp also holds the halo planes of a sharded grid: haet_slice_reduce_halo)
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rows) return;
    const int64_t bh = t / npad, n = t - bh * npad;
    const int b = (int)(bh / H), h = (int)(bh - (int64_t)b * H);
    uint4 out[8];
    float sn = 0.f;
    if (n < N) {
        const int Cr = C == 64 ? 32 : C;   // channels per image row
        const float4 *src = reinterpret_cast<const float4 *>(xp + ((int64_t)b * Nxp + n) * (H * Cr) + h * Cr);
        float d = bt;
        const float *wtr = wt;
        float d64 = 0.f;
        if (C == 64) {   // Ada-Temp needs the whole head; summed in ONE fixed order so that both sub-rows of a head get the same bits
            float dh[2];
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                const float4 *sp = reinterpret_cast<const float4 *>(xp + ((int64_t)b * Nxp + n) * (H * Cr) + ((h & ~1) + hf) * Cr);
                const float *wh = wt + hf * 32;
                float acc = 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 a = __ldg(sp + j);
                    acc = fmaf(a.x, __ldg(wh + 4 * j), acc); acc = fmaf(a.y, __ldg(wh + 4 * j + 1), acc);
                    acc = fmaf(a.z, __ldg(wh + 4 * j + 2), acc); acc = fmaf(a.w, __ldg(wh + 4 * j + 3), acc);
                }
                dh[hf] = acc;
            }
            d64 = (bt + dh[0]) + dh[1];
            wtr = wt + (h & 1) * 32;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {   // chunk j: channels 8j .. 8j+7
            if (8 * j >= Cr) { out[j] = out[4 + j] = make_uint4(0, 0, 0, 0); continue; }
            const float4 a = __ldg(src + 2 * j), c = __ldg(src + 2 * j + 1);
            const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                d = fmaf(v[2 * i], __ldg(wtr + 8 * j + 2 * i), d);
                d = fmaf(v[2 * i + 1], __ldg(wtr + 8 * j + 2 * i + 1), d);
                hi[i] = tc::pack_h2_sat(v[2 * i], v[2 * i + 1]);   // saturating: never (inf, -inf)
                const float2 hf = __half22float2(*reinterpret_cast<const __half2 *>(&hi[i]));
                lo[i] = tc::pack_h2_sat(v[2 * i] - hf.x, v[2 * i + 1] - hf.y);
            }
            out[j] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            out[4 + j] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        if (C == 64) d = d64;
        const float tau = base_temp + fminf(fmaxf(d, -0.4f), 0.4f);   // Ada-Temp (IrregularMesh.py:168)
        sn = kLog2e / tau;
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) out[j] = make_uint4(0, 0, 0, 0);
    }
    uint8_t *row = img + t * XPI_ROW_BYTES;
    const int sw = (int)(n & 7);
#pragma unroll
    for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4 *>(row + ((j ^ sw) << 4)) = out[j];
    scale[t] = sn;
}

int launch_xpi_pack(const haet_layer_shape *s, const float *xp, const float *wt, float bt, uint8_t *img, float *scale,
                    cudaStream_t st, int64_t xp_rows_per_batch) {
    HAET_CHECK_ARG(s->dim_head == 32 || s->dim_head == 16 || s->dim_head == 64, "the operand image needs dim_head 16, 32 or 64");
    const int ph = s->heads * (s->dim_head == 64 ? 2 : 1);
    const int64_t npad = xpi_npad(s->N), rows = s->B * ph * npad;
    if (rows == 0) return HAET_OK;
    xpi_pack_kernel<<<(unsigned)cdiv(rows, 256), 256, 0, st>>>(xp, wt, bt, s->base_temp, s->N, npad, ph, s->dim_head, rows, img, scale,
                                                                  xp_rows_per_batch > 0 ? xp_rows_per_batch : s->N);
    HAET_LAUNCHED();
    return HAET_OK;
}

}  // namespace haet
