// Shared-memory / shuffle issue costs on sm_100a, measured (the streaming kernel is bound by the
// shared-memory / L1 data pipe; which instruction costs how many cycles of it, and do shuffles run somewhere else?)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/mb_shfl scripts/microbench_shfl_lds.cu && scripts/mb_shfl
// Every kernel: 148 x 4 CTAs of 128 threads (the streaming kernel's shape: 16 warps per SM), ITER iterations of 8 independent
// conflict-free accesses per thread (asm volatile with a memory clobber: neither hoisted nor merged) and / or 32 shuffles in 4 independent chains.
// Prints cycles per loop iteration (clock64 of CTA 0) for each kind, and for LDS.128 + SHFL together; read with ncu.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITER = 4096;
enum { LDS32 = 1, LDS64 = 2, LDS128 = 3, STS32 = 4, STS64 = 5, STS128 = 6, SHFL = 8 };

template <int MODE>
__global__ void __launch_bounds__(128, 4) k(float* out, long long* cycles) {
    __shared__ float4 sm[128 * 8];
    const int tid = threadIdx.x;
    for (int i = tid; i < 128 * 8; i += 128) sm[i] = make_float4(i, 1.f, 2.f, 3.f);
    __syncthreads();
    float acc = 0.f, v0 = static_cast<float>(tid), v1 = v0 + 1.f, v2 = v0 + 2.f, v3 = v0 + 3.f;
    constexpr int M = MODE & 7;
    constexpr int bytes = (M == LDS32 || M == STS32) ? 4 : ((M == LDS64 || M == STS64) ? 8 : 16);
    const unsigned base = static_cast<unsigned>(__cvta_generic_to_shared(sm)) + tid * bytes;
    const long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
        if (M != 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const unsigned addr = base + ((it + j) & 7) * 128 * bytes;  // consecutive lanes, consecutive words: conflict-free
                float4 x = make_float4(v0, v1, v2, v3);
                if (M == LDS32) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x.x) : "r"(addr) : "memory");
                if (M == LDS64) asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x.x), "=f"(x.y) : "r"(addr) : "memory");
                if (M == LDS128) asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "r"(addr) : "memory");
                if (M == STS32) asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(x.x) : "memory");
                if (M == STS64) asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(x.x), "f"(x.y) : "memory");
                if (M == STS128) asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w) : "memory");
                acc += x.x + x.y + x.z + x.w;
            }
        }
        if (MODE & SHFL) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v0 += __shfl_xor_sync(0xffffffffu, v0, j + 1);
                v1 += __shfl_xor_sync(0xffffffffu, v1, j + 2);
                v2 += __shfl_xor_sync(0xffffffffu, v2, j + 3);
                v3 += __shfl_xor_sync(0xffffffffu, v3, j + 4);
            }
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * 128 + tid] = acc + v0 + v1 + v2 + v3;
    if (tid == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
double run(float* out, long long* cyc, int grid) {
    k<MODE><<<grid, 128>>>(out, cyc);
    cudaDeviceSynchronize();
    k<MODE><<<grid, 128>>>(out, cyc);
    cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    return static_cast<double>(h) / ITER;  // cycles per iteration with the SM's 16 warps running
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * 4;
    float* out;
    long long* cyc;
    cudaMalloc(&out, grid * 128 * sizeof(float));
    cudaMalloc(&cyc, grid * sizeof(long long));
    // Cycles of CTA 0 per loop iteration, SM fully loaded.  Take the executed instruction counts from ncu
    // (sm__inst_executed_pipe_lsu.sum: ptxas halves the trip count of the pure load loops), not from the source:
    //   ncu --clock-control none --metrics l1tex__data_pipe_lsu_wavefronts.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,\
    //       sm__inst_executed_pipe_lsu.sum,sm__cycles_elapsed.max scripts/mb_shfl        (profiles/r02e/microbench_lsu.md)
    const double l32 = run<LDS32>(out, cyc, grid), l64 = run<LDS64>(out, cyc, grid), l128 = run<LDS128>(out, cyc, grid);
    const double s32 = run<STS32>(out, cyc, grid), s64 = run<STS64>(out, cyc, grid), s128 = run<STS128>(out, cyc, grid);
    const double sh = run<SHFL>(out, cyc, grid);
    const double both = run<LDS128 | SHFL>(out, cyc, grid);
    printf("{\"sms\": %d, \"cycles_per_iteration\": {\"LDS.32\": %.1f, \"LDS.64\": %.1f, \"LDS.128\": %.1f, \"STS.32\": %.1f, \"STS.64\": %.1f, "
           "\"STS.128\": %.1f, \"SHFL_x512\": %.1f, \"LDS.128_and_SHFL_x512\": %.1f}}\n",
           sms, l32, l64, l128, s32, s64, s128, sh, both);
    return 0;
}
