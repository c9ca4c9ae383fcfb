// Micro-benchmark: what bounds the softmax loops of the two passes?  Elements (logits) per clock and SM for the instruction
// mixes of the loop body, at the softmax warps' occupancy (4 warps per scheduler) and at full occupancy.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o xu_mix xu_mix.cu && ./xu_mix
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t f2fp(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ uint64_t pk(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(uint64_t v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

// MODE 0: 2 ex2                      1: 2 ex2 + F2FP                 2: FFMA2 + 2 ex2 + F2FP (pass 2 body)
//      3: FFMA2 + 2 ex2 + integer pack (2 x (shl + add) as LEA, 1 PRMT)        4: FFMA2 + 2 ex2 + PRMT only (truncating pack)
//      5: F2FP only                  6: FFMA2 only                   7: FFMA2 + 2 ex2 (no pack: result xor-ed)
template <int MODE>
__global__ void k(uint32_t *out, int iters, float seed) {
    float a[8];
    uint32_t h[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed + 0.001f * (threadIdx.x + i);
    const uint64_t s2 = pk(0.999f, 0.999f), c2 = pk(-0.0001f, -0.0002f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; i += 2) {   // one PAIR of logits per step
            float v0 = a[i], v1 = a[i + 1];
            if (MODE == 2 || MODE == 3 || MODE == 4 || MODE == 6 || MODE == 7) upk(fma2(pk(v0, v1), s2, c2), v0, v1);
            float p0 = v0, p1 = v1;
            if (MODE != 5 && MODE != 6) { p0 = ex2(v0); p1 = ex2(v1); }
            uint32_t w;
            if (MODE == 1 || MODE == 2 || MODE == 5) w = f2fp(p0, p1);
            else if (MODE == 3) {
                const uint32_t b0 = (__float_as_uint(p0) << 3) + 0x8000u, b1 = (__float_as_uint(p1) << 3) + 0x8000u;
                w = __byte_perm(b0, b1, 0x7632);
            } else if (MODE == 4) w = __byte_perm(__float_as_uint(p0), __float_as_uint(p1), 0x7632);
            else w = __float_as_uint(p0) ^ __float_as_uint(p1);
            h[i >> 1] ^= w;
            a[i] = v0 * 0.5f; a[i + 1] = v1 * 0.5f;   // keep the chain alive without another XU op (FMUL)
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = h[0] ^ h[1] ^ h[2] ^ h[3] ^ __float_as_uint(a[0] + a[7]);
}

template <int MODE>
void run(const char *name, int threads) {
    int sms = 148, clk_khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int blocks = sms, iters = 8192;
    uint32_t *out;
    cudaMalloc(&out, sizeof(uint32_t) * threads * blocks);
    k<MODE><<<blocks, threads>>>(out, 16, 0.5f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(out, iters, 0.5f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double elems = (double)threads * blocks * iters * 8, cyc = ms * 1e-3 * clk_khz * 1e3;
    printf("%-58s %4d thr/SM %8.3f ms  %6.2f elements/clk/SM  [%s]\n", name, threads, ms, elems / cyc / sms, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    for (int threads : {512, 1024}) {
        run<0>("2 ex2", threads);
        run<5>("F2FP only", threads);
        run<6>("FFMA2 only", threads);
        run<1>("2 ex2 + F2FP", threads);
        run<7>("FFMA2 + 2 ex2", threads);
        run<2>("FFMA2 + 2 ex2 + F2FP (pass-2 loop body)", threads);
        run<3>("FFMA2 + 2 ex2 + integer pack (2 LEA + PRMT)", threads);
        run<4>("FFMA2 + 2 ex2 + PRMT (truncating pack)", threads);
    }
    return 0;
}
