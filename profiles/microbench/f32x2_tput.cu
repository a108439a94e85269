// f32x2_tput.cu — issue-rate microbenchmark: scalar FFMA/FADD vs packed FFMA2/FADD2/FMUL2 on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_tput f32x2_tput.cu
// Prints flops/clk/SM equivalents measured with CUDA events (full chip, 148 x 2 CTAs x 512 threads).
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int CH = 8;  // independent chains per thread

template <int MODE>
__global__ void __launch_bounds__(512) k(float* out, float a, float b) {
    float2 v[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) v[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
    const float2 a2 = make_float2(a, a * 1.0001f), b2 = make_float2(b, b * 0.999f);
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CH; i++) {
            if (MODE == 0) { v[i].x = fmaf(v[i].x, a2.x, b2.x); v[i].y = fmaf(v[i].y, a2.y, b2.y); }
            if (MODE == 1) v[i] = __ffma2_rn(v[i], a2, b2);
            if (MODE == 2) { v[i].x = __fadd_rn(v[i].x, b2.x); v[i].y = __fadd_rn(v[i].y, b2.y); }
            if (MODE == 3) v[i] = __fadd2_rn(v[i], b2);
            if (MODE == 4) { v[i].x = __fmul_rn(v[i].x, a2.x); v[i].y = __fmul_rn(v[i].y, a2.y); }
            if (MODE == 5) v[i] = __fmul2_rn(v[i], a2);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) s += v[i].x + v[i].y;
    if (s == 123.456f) out[0] = s;
}

template <int MODE>
void run(const char* name, float* d) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = 148 * 2;
    k<MODE><<<grid, 512>>>(d, 1.0f, 0.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<grid, 512>>>(d, 1.0f, 0.0f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double elem_ops = (double)grid * 512 * ITERS * CH * 2;   // scalar-equivalent operations
    printf("%-8s %.3f ms  %.1f G elem-op/s  (%.1f elem-op/clk/SM at 1.9 GHz)\n", name, ms, elem_ops / ms * 1e-6,
           elem_ops / (ms * 1e-3) / 148 / 1.9e9);
}

int main() {
    float* d; cudaMalloc(&d, 4);
    run<0>("FFMA", d); run<1>("FFMA2", d); run<2>("FADD", d); run<3>("FADD2", d); run<4>("FMUL", d); run<5>("FMUL2", d);
    return 0;
}
