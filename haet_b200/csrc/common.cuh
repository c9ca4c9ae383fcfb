// Shared helpers of libhaet_b200 (device + host).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#include <atomic>

#include "../../include/haet_b200.h"

namespace haet {

void set_error(const char *fmt, ...);
extern std::atomic<int64_t> g_launches;   // haet_launch_count()

// Ablation switches ("environment switches"): read from the environment ONCE, when the library is first used,
// and immutable afterwards, so that a *_bytes() query and the launch that follows always agree.
struct Options {
    bool gemm_1cta, no_opi, no_ln_fuse, conv_simt, conv_nh1, no_pdl, erwin_fused;
    int poly_p1, poly_p2;    // HAET_POLY_P1 / HAET_POLY_P2: eighths (0..4) of pass 1's / pass 2's exponentials computed by the FMA-pipe polynomial
    bool mlp_fc2_split;      // HAET_MLP_FC2_SPLIT: block MLP fc2 as one CTA-pair launch per 256 hidden columns (A/B of the streamed single pass)
    bool conv_xp;            // HAET_CONV_XP: conv front-ends write fp32 x_proj + a separate image pack (A/B of the image epilogue)
    bool no_f32x2;           // HAET_NO_F32X2: scalar FFMA / FADD in the softmax loops of the two passes (A/B of the packed pairs)
    int tree_roles;          // 0: one CTA per rotated level
    int gemm_pf;             // HAET_GEMM_PF: L2 prefetch distance of the GEMM producers in (tile, panel) items
    double peer_timeout_s;
};
const Options &options();
// Test / bench hooks (haet_set_path, haet_set_stage_mask, haet_set_raise_threshold): THREAD-LOCAL, so a caller that never
// touches them is never affected by another thread that does -- the entry points themselves keep no mutable state.
struct Hooks {
    int path = 0;            // 0: tensor-core kernels where supported; 1: fp32 SIMT kernels everywhere
    int stage_mask = 15;
    float st_raise = 0.f;    // 0: default lazy-max threshold of pass 1
};
Hooks &hooks();

// Opt a kernel in to `bytes` of dynamic shared memory.  The attribute is per device, so the cache is too (a second GPU in the
// same process gets its own opt-in); a race between two threads only repeats the (idempotent) call.
#define HAET_DYN_SMEM(kernel, bytes)                                                                          \
    do {                                                                                                      \
        static std::atomic<size_t> done_[64];                                                                 \
        int dev_ = 0;                                                                                         \
        HAET_CUDA(cudaGetDevice(&dev_));                                                                      \
        if ((size_t)(bytes) > done_[dev_ & 63].load(std::memory_order_relaxed)) {                             \
            HAET_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            done_[dev_ & 63].store((size_t)(bytes), std::memory_order_relaxed);                               \
        }                                                                                                     \
    } while (0)

#define HAET_CHECK_ARG(cond, ...)            \
    do {                                     \
        if (!(cond)) {                       \
            haet::set_error(__VA_ARGS__);    \
            return HAET_EINVAL;              \
        }                                    \
    } while (0)

#define HAET_CUDA(call)                                                                   \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) {                                                          \
            haet::set_error("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return HAET_ECUDA;                                                            \
        }                                                                                 \
    } while (0)

// after a kernel launch: count it and surface launch-configuration errors
#define HAET_LAUNCHED()                      \
    do {                                     \
        haet::g_launches.fetch_add(1, std::memory_order_relaxed); \
        HAET_CUDA(cudaGetLastError());       \
    } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t align_up(int64_t a, int64_t b) { return cdiv(a, b) * b; }

constexpr float kLog2e = 1.4426950408889634f;

// log(-log(eps)) evaluated like the reference does (fp32 tensor ops, IrregularMesh.py:172)
static inline float rep_slice_shift(float eps) { return logf(-logf(eps)); }

// ---- the per-point spill between pass 1 and pass 2 (opaque to the caller; haet_spill_bytes) ----------
// Tensor-core path: the projected points are kept as a ready-made UMMA operand image
//   img  [B*H][Npad rows][128 B]: row n of head h = 32 fp16 "hi" then 32 fp16 "lo" (x_proj = hi + lo to ~2^-22);
//        the 16-byte chunk j of a row sits at chunk position j ^ (n & 7), i.e. any run of rows starting at a
//        multiple of 8 IS a SWIZZLE_128B shared-memory tile (K-major for the logits MMA, MN-major for P.x) and
//        is brought in by one bulk copy.  Npad = N rounded up to 128; pad rows are zero.
//   scale[B*H][Npad] fp32: s_n = log2(e) / tau_n (Ada-Temp, IrregularMesh.py:168); 0 on pad rows.
//   xp   [B][N][H*C] fp32: only for the fp32 SIMT kernels and the conv front-ends (which pack img from it).
constexpr int XPI_ROW_BYTES = 128;
static inline int64_t xpi_npad(int64_t N) { return align_up(N, 128); }
struct SpillView {
    uint8_t *img;
    float *scale;
    float *xp;
    int64_t npad;
};
bool spill_needs_xp(const haet_layer_shape *s);
SpillView spill_view(const haet_layer_shape *s, void *base);
int launch_xpi_pack(const haet_layer_shape *s, const float *xp, const float *wt, float bt, uint8_t *img, float *scale,
                    cudaStream_t st, int64_t xp_rows_per_batch = 0);

// ---- stage entry points implemented in the .cu files ------------------------------------
int launch_in_project(const haet_layer_shape *s, const float *x, const float *wx, const float *bx,
                      float *xp, cudaStream_t st);
int launch_linear(const float *a, const float *w, const float *bias, float *out, int64_t M, int K,
                  int Nn, cudaStream_t st);
bool gemm_tc_supported(int K, int Nn);
int64_t gemm_tc_image_bytes(int K, int Nn);
// OPI: the to_out GEMM's A operand as pass 2 writes it when both tensor-core kernels run: rows m = b * N + n in tiles of 128,
// [tile][k panel of 64][hi | lo][128 rows x 128 B]: 16-byte chunk c (8 bf16) of row r at r * 128 + ((c ^ (r & 7)) << 4), i.e.
// every (tile, panel) is one 32 KB SWIZZLE_128B K-major stage of the GEMM, loaded by one bulk copy (no conversion warps).
bool gemm_opi_supported(int K, int Nn);
inline int64_t opi_bytes(int64_t M, int K) { return (M + 255) / 256 * 256 * (int64_t)K * 4; }
// resid (nullable) [M, Nn]: added to the output rows in the epilogue (block fusion)
int launch_gemm_tc_opi(const uint8_t *a_img, const float *w, const float *bias, float *y, int64_t M, int K, int Nn, void *img,
                       cudaStream_t st, const float *resid = nullptr);
int launch_gemm_tc(const float *x, const float *w, const float *bias, float *y, int64_t M, int K, int Nn, void *img,
                   cudaStream_t st, const float *resid = nullptr);
// ln_1 in front of the layer (HAETransolver_Irregular_Mesh.py:125), fused into the CTA-pair in-projection when ln != nullptr:
// scratch = ln_fused_scratch_bytes() bytes (folded weights, wsum, per-row statistics)
struct LnArgs {
    const float *gamma, *beta;
    float eps;
    void *scratch;
};
bool ln_fused_supported(const haet_layer_shape *s);
int64_t ln_fused_scratch_bytes(const haet_layer_shape *s);
int launch_inproj_tc(const haet_layer_shape *s, const float *x, const float *wx, const float *bx, const float *wt, float bt,
                     uint8_t *xpi, float *scale, void *img, cudaStream_t st, const LnArgs *ln = nullptr);
// block MLP on the CTA-pair kernel (gemm_tc.cu)
bool mlp_tc_supported(int dim, int hidden);
int64_t mlp_tc_workspace_bytes(int64_t M, int dim, int hidden);
int launch_mlp_tc(const float *x, int64_t M, int dim, int hidden, const float *ln_g, const float *ln_b, float ln_eps,
                  const float *w1, const float *b1, const float *w2, const float *b2, float *out, void *workspace, cudaStream_t st);
// unfused helpers for the shapes / paths the fused kernels do not cover
int launch_layernorm_rows(const float *x, const float *gamma, const float *beta, float eps, float *y, int64_t M, int K, cudaStream_t st);
int launch_add_rows(float *y, const float *resid, int64_t n, cudaStream_t st);
// Conv2d / Conv3d front-end as an implicit GEMM on tcgen05 (conv_tc.cu); wx tap-major [inner, taps, dim]
bool conv_tc_supported(const haet_layer_shape *s);
int64_t conv_tc_image_bytes(const haet_layer_shape *s);
struct GemmEpi;
int launch_conv_tc(const haet_layer_shape *s, const float *x, const float *wx, const float *bx, float *xp, void *img,
                   cudaStream_t st, const GemmEpi *epi = nullptr);
// the same with the operand image + scales of the spill written by the epilogue (dim_head 16 / 32): no fp32 x_proj, no pack kernel
int launch_conv_tc_img(const haet_layer_shape *s, const float *x, const float *wx, const float *bx, const float *wt, float bt,
                       uint8_t *xpi, float *scale, void *img, cudaStream_t st);
int launch_xpi_pad(const haet_layer_shape *s, uint8_t *xpi, float *scale, cudaStream_t st);
bool gemm_stream_opi_supported(int K, int Nn);
int launch_gemm_stream_opi(const uint8_t *opi, const float *w, const float *bias, const float *resid, float *out, int64_t M, int K, int Nn,
                           void *img, cudaStream_t st);   // zero the image rows N .. npad
bool slice_tc_supported(const haet_layer_shape *s);
int64_t slice_tc_chunks(const haet_layer_shape *s);
int launch_slice_tc(const haet_layer_shape *s, const uint8_t *img, const float *scale, const float *ws,
                    const float *bs, float *chunk_parts, cudaStream_t st);
bool deslice_tc_supported(const haet_layer_shape *s);
int launch_deslice_tc(const haet_layer_shape *s, const uint8_t *img, const float *scale, const float *lse2,
                      const float *states, const float *ws, const float *bs, float *o, int o_image, cudaStream_t st);
int64_t slice_chunks(const haet_layer_shape *s);
int launch_slice_partial(const haet_layer_shape *s, const float *xp, const float *ws, const float *bs,
                         const float *wt, float bt, float *chunk_parts, cudaStream_t st);
int launch_slice_merge(const haet_layer_shape *s, const float *parts, int64_t n_parts, float *packed_out,
                       float *eidetic, float *lse2, cudaStream_t st);
int launch_slice_pos(const haet_layer_shape *s, const float *eidetic, float *pos, cudaStream_t st);
int launch_deslice(const haet_layer_shape *s, const float *xp, const float *lse2, const float *states,
                   const float *ws, const float *bs, const float *wt, float bt, float *o, cudaStream_t st);

}  // namespace haet
