#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(*reinterpret_cast<unsigned long long*>(&d)) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
    return d;
}
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float* out, int iters, float s, int* iout) {
    float2 a[8];
    int ia[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f - i);
#pragma unroll
    for (int i = 0; i < 4; ++i) ia[i] = threadIdx.x + i;
    float2 m = make_float2(s, s), c = make_float2(0.5f * s, 0.25f * s);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (MODE == 0) {  // scalar: 2 FFMA per element
                    a[i].x = fmaf(a[i].x, m.x, c.x);
                    a[i].y = fmaf(a[i].y, m.y, c.y);
                } else if (MODE == 1) {  // packed: 1 FFMA2
                    a[i] = ffma2(a[i], m, c);
                } else if (MODE == 2) {  // packed + 1 int op per packed
                    a[i] = ffma2(a[i], m, c);
                    ia[i & 3] = (ia[i & 3] ^ it) + r;
                } else if (MODE == 3) {  // scalar (2 FFMA) + 1 int op
                    a[i].x = fmaf(a[i].x, m.x, c.x);
                    a[i].y = fmaf(a[i].y, m.y, c.y);
                    ia[i & 3] = (ia[i & 3] ^ it) + r;
                }
            }
        }
    }
    float acc = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    iout[blockIdx.x * blockDim.x + threadIdx.x] = ia[0] + ia[1] + ia[2] + ia[3];
}
template <int MODE>
void run(const char* name, float* out, int* iout) {
    int iters = 4000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148, 512>>>(out, 10, 0.999f, iout);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<148, 512>>>(out, iters, 0.999f, iout);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fmas = 148.0 * 512 * iters * 4 * 8 * 2;
    printf("%s: %.3f ms, %.2f TFMA/s (x2 = TFLOP/s %.1f), fma lanes/clk/SM @1.965GHz = %.1f\n", name, ms, fmas / ms * 1e-9, 2 * fmas / ms * 1e-9, fmas / (ms * 1e-3) / 148 / 1.965e9);
}
int main() {
    float* out; int* iout;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&iout, 148 * 512 * 4);
    run<0>("scalar FFMA        ", out, iout);
    run<1>("packed FFMA2       ", out, iout);
    run<2>("packed FFMA2 + int ", out, iout);
    run<3>("scalar FFMA + int  ", out, iout);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
