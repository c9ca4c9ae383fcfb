// Shared helpers of libhaet_b200 (device + host).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#include "../../include/haet_b200.h"

namespace haet {

void set_error(const char *fmt, ...);
extern int64_t g_launches;

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
        ++haet::g_launches;                  \
        HAET_CUDA(cudaGetLastError());       \
    } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t align_up(int64_t a, int64_t b) { return cdiv(a, b) * b; }

constexpr float kLog2e = 1.4426950408889634f;

// log(-log(eps)) evaluated like the reference does (fp32 tensor ops, IrregularMesh.py:172)
static inline float rep_slice_shift(float eps) { return logf(-logf(eps)); }

// ---- stage entry points implemented in the .cu files ------------------------------------
int launch_in_project(const haet_layer_shape *s, const float *x, const float *wx, const float *bx,
                      float *xp, cudaStream_t st);
int launch_linear(const float *a, const float *w, const float *bias, float *out, int64_t M, int K,
                  int Nn, cudaStream_t st);
bool gemm_tc_supported(int K, int Nn);
int64_t gemm_tc_image_bytes(int K, int Nn);
int launch_gemm_tc(const float *x, const float *w, const float *bias, float *y, int64_t M, int K, int Nn, void *img,
                   cudaStream_t st);
bool slice_tc_supported(const haet_layer_shape *s);
int64_t slice_tc_chunks(const haet_layer_shape *s);
int launch_slice_tc(const haet_layer_shape *s, const float *xp, const float *ws, const float *bs, const float *wt,
                    float bt, float *chunk_parts, cudaStream_t st);
bool deslice_tc_supported(const haet_layer_shape *s);
int launch_deslice_tc(const haet_layer_shape *s, const float *xp, const float *lse2, const float *states,
                      const float *ws, const float *bs, const float *wt, float bt, float *o, cudaStream_t st);
extern int g_path;   // 0: tensor-core kernels where supported; 1: fp32 SIMT kernels everywhere
int64_t slice_chunks(const haet_layer_shape *s);
int launch_slice_partial(const haet_layer_shape *s, const float *xp, const float *ws, const float *bs,
                         const float *wt, float bt, float *chunk_parts, cudaStream_t st);
int launch_slice_merge(const haet_layer_shape *s, const float *parts, int64_t n_parts, float *packed_out,
                       float *eidetic, float *lse2, cudaStream_t st);
int launch_slice_pos(const haet_layer_shape *s, const float *eidetic, float *pos, cudaStream_t st);
int launch_deslice(const haet_layer_shape *s, const float *xp, const float *lse2, const float *states,
                   const float *ws, const float *bs, const float *wt, float bt, float *o, cudaStream_t st);

}  // namespace haet
