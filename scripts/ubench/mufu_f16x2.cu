// Micro-benchmark: issue rate of ex2.approx.ftz.f32, ex2.approx.f16x2 and cvt.rn.f16x2.f32 (F2FP) per SM on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_f16x2 mufu_f16x2.cu && ./mufu_f16x2
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

template <int MODE>
__global__ void k(uint32_t *out, int iters, float seed) {
    float a[8];
    uint32_t h[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = seed + 0.001f * (threadIdx.x + i); h[i] = 0x3c003800u + threadIdx.x + i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) {          // fp32 ex2: 1 element per instruction
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            } else if (MODE == 1) {   // packed fp16 ex2: 2 elements per instruction
                asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
            } else if (MODE == 2) {   // F2FP pack only
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(a[i]), "f"(a[(i + 1) & 7]));
                a[i] = __uint_as_float(h[i]);
            } else if (MODE == 3) {   // current pass-1 mix per pair: 2 ex2.f32 + 1 F2FP
                float p0, p1;
                asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(a[i]));
                asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(a[(i + 1) & 7]));
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(p0), "f"(p1));
                a[i] = __uint_as_float(h[i]) * 1e-30f;
            } else if (MODE == 4) {   // candidate mix per pair: 1 F2FP + 1 ex2.f16x2
                uint32_t t;
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(t) : "f"(a[i]), "f"(a[(i + 1) & 7]));
                asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(h[i]) : "r"(t));
                a[i] = __uint_as_float(h[i]) * 1e-30f;
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= h[i] ^ __float_as_uint(a[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
void run(const char *name, int per_iter_pairs) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int threads = 512, blocks = sms * 2, iters = 4096;
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
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double groups = (double)threads * blocks * iters * 8;   // loop bodies executed (thread level)
    const double cyc = ms * 1e-3 * clk_khz * 1e3;
    printf("%-40s %8.3f ms  %7.2f thread-bodies/clk/SM  (%.2f elements/clk/SM)  [%s]\n", name, ms, groups / cyc / sms,
           groups * per_iter_pairs / cyc / sms, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    run<0>("ex2.f32 (1 elem)", 1);
    run<1>("ex2.f16x2 (2 elem)", 2);
    run<2>("F2FP cvt.rn.f16x2.f32 (2 elem)", 2);
    run<3>("2 ex2.f32 + F2FP (2 elem)", 2);
    run<4>("F2FP + ex2.f16x2 (2 elem)", 2);
    return 0;
}
