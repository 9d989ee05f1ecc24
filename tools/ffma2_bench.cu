// Microbenchmark: scalar FFMA vs packed FFMA2 issue/pipe throughput on sm_100a (decides whether the tile sweeps
// should be written on float2 pixel pairs).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 ffma2_bench.cu
#include <cuda_runtime.h>
#include <cstdio>
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
    float2 a[8];
    unsigned u[4] = {threadIdx.x, threadIdx.x * 3u, 7u, blockIdx.x};
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
    const float2 m = make_float2(s, s), c = make_float2(0.25f, 0.125f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0 || MODE == 2) { a[i].x = fmaf(a[i].x, m.x, c.x); a[i].y = fmaf(a[i].y, m.y, c.y); }
                else a[i] = __ffma2_rn(a[i], m, c);
            }
            if (MODE >= 2) {   // interleave ALU-pipe work (16 lane-FMAs : 8 ALU ops)
#pragma unroll
                for (int j = 0; j < 4; j++) { u[j] = (u[j] ^ (u[(j + 1) & 3] >> 3)) + 0x9e37u; u[j] = (u[j] << 5) | (u[j] >> 27); }
            }
        }
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) acc += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + (float)(u[0] ^ u[1] ^ u[2] ^ u[3]);
}
// MODE 4: FFMA2 with three distinct 64-bit register operands (no operand reuse); MODE 5: same as scalar FFMAs;
// MODE 6: FFMA2 (3 regs) + one MUFU.EX2 per 4 FFMA2 (the tile sweeps' mix)
template <int MODE>
__global__ void __launch_bounds__(256) k3(float* out, int iters, float s) {
    float2 a[6], b[6], c[6];
#pragma unroll
    for (int i = 0; i < 6; i++) { a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f); b[i] = make_float2(s, s * 0.999f + i * 1e-4f); c[i] = make_float2(0.25f * i, 0.125f); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 6; i++) {
                const int j = (i + r) % 6, k = (i + 2 * r + 1) % 6;
                if (MODE == 5) { a[i].x = fmaf(b[j].x, c[k].x, a[i].x); a[i].y = fmaf(b[j].y, c[k].y, a[i].y); }
                else a[i] = __ffma2_rn(b[j], c[k], a[i]);
            }
            if (MODE == 6) {
#pragma unroll
                for (int i = 0; i < 2; i++) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(c[i].x)); c[i].x = y; }
            }
        }
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 6; i++) acc += a[i].x + a[i].y + b[i].x + c[i].x;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE> void run3(const char* name, float* d) {
    const int blocks = 148 * 8, iters = 4096;
    k3<MODE><<<blocks, 256>>>(d, 16, 0.999f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k3<MODE><<<blocks, 256>>>(d, iters, 0.999f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fma = (double)blocks * 256 * iters * 4 * 12;
    printf("%-28s %8.3f ms  %7.2f TFLOP/s (FMA=2)  %6.1f lane-FMA/clk/SM @1.965GHz\n", name, ms, 2 * fma / ms / 1e9,
           fma / (ms * 1e-3) / 148 / 1.965e9);
}
template <int MODE> void run(const char* name, float* d) {
    const int blocks = 148 * 8, iters = 4096;
    k<MODE><<<blocks, 256>>>(d, 16, 0.999f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(d, iters, 0.999f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fma = (double)blocks * 256 * iters * 4 * 16;
    printf("%-28s %8.3f ms  %7.2f TFLOP/s (FMA=2)  %6.1f lane-FMA/clk/SM @1.965GHz\n", name, ms, 2 * fma / ms / 1e9,
           fma / (ms * 1e-3) / 148 / 1.965e9);
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("scalar FFMA", d); run<1>("packed FFMA2", d); run<2>("scalar FFMA + ALU mix", d); run<3>("packed FFMA2 + ALU mix", d);
    run3<4>("FFMA2 3 distinct regs", d); run3<5>("FFMA 3 distinct regs", d); run3<6>("FFMA2 3 regs + MUFU 1:12", d);
    cudaError_t e = cudaDeviceSynchronize(); printf("%s\n", cudaGetErrorString(e));
    return 0;
}
