// Noise / roll augmentation of the reference's dataset (src/dataloaders/audioset_dataset.py:197-199):
//     fbank = fbank + torch.rand(T, F) * np.random.rand() / 10
//     fbank = torch.roll(fbank, np.random.randint(-10, 10), 0)
// for a batch:  out[b, (t + shift_b) mod T, f] = in[b, t, f] + (noise[b, t, f] * scale_b) / 10,   scale_b = fp32(np.random.rand()).
// The two numpy scalars are drawn on the host (in the reference's RNG order) and passed in.  The uniform noise field is
// either passed in as well (a device tensor the host drew with torch.rand: bit-exact parity with the reference) or
// generated here with Philox-4x32-10 (same distribution, U[0,1) with 24 random bits like torch.rand for fp32; it cannot
// reproduce torch's CPU generator).  Off in every recipe of the reference (run.py:77,80), so it is a separate
// HBM-bound pass (one read, one write of the features) instead of a branch in the fused kernel.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace psla {

constexpr int kNoiseClips = 256;  // clips per launch: their (scale, shift) travel in the kernel's parameter space
struct NoiseMeta {
    float scale[kNoiseClips];
    int shift[kNoiseClips];  // normalised to [0, T)
};

__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    constexpr unsigned kM0 = 0xD2511F53u, kM1 = 0xCD9E8D57u, kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(kM0, ctr.x), lo0 = kM0 * ctr.x;
        const unsigned hi1 = __umulhi(kM1, ctr.z), lo1 = kM1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += kW0;
        key.y += kW1;
    }
    return ctr;
}
__device__ __forceinline__ float u01(unsigned x) { return static_cast<float>(x >> 8) * (1.0f / 16777216.0f); }

// one warp per output row; block = 256 threads = 8 rows
template <bool VEC128>
__global__ void __launch_bounds__(256) noise_roll_kernel(const float* __restrict__ in, float* __restrict__ out, const float* __restrict__ noise,
                                                         const __grid_constant__ NoiseMeta meta, int n_clips, int T, int nmel,
                                                         unsigned long long seed, int clip0) {
    const int lane = threadIdx.x & 31;
    const long long row = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
    if (row >= static_cast<long long>(n_clips) * T) return;
    const int b = static_cast<int>(row / T), t = static_cast<int>(row - static_cast<long long>(b) * T);
    const float s = meta.scale[b];
    int t2 = t + meta.shift[b];
    t2 -= (t2 >= T) ? T : 0;
    const size_t src = (static_cast<size_t>(b) * T + t) * nmel, dst = (static_cast<size_t>(b) * T + t2) * nmel;
    const uint2 key = make_uint2(static_cast<unsigned>(seed), static_cast<unsigned>(seed >> 32));
    if (VEC128) {  // nmel == 128: one float4 per lane
        const float4 x = __ldcs(reinterpret_cast<const float4*>(in + src) + lane);
        float4 u;
        if (noise) u = __ldcs(reinterpret_cast<const float4*>(noise + src) + lane);
        else {
            const uint4 r = philox4x32_10(make_uint4(static_cast<unsigned>(t) * 32u + lane, static_cast<unsigned>(clip0 + b), 0u, 0u), key);
            u = make_float4(u01(r.x), u01(r.y), u01(r.z), u01(r.w));
        }
        // the reference evaluates (rand * np_scalar) / 10 elementwise and then adds: same two roundings here
        float4 y;
        y.x = __fadd_rn(x.x, __fdiv_rn(__fmul_rn(u.x, s), 10.0f));
        y.y = __fadd_rn(x.y, __fdiv_rn(__fmul_rn(u.y, s), 10.0f));
        y.z = __fadd_rn(x.z, __fdiv_rn(__fmul_rn(u.z, s), 10.0f));
        y.w = __fadd_rn(x.w, __fdiv_rn(__fmul_rn(u.w, s), 10.0f));
        __stcs(reinterpret_cast<float4*>(out + dst) + lane, y);
    } else {
        for (int f = lane; f < nmel; f += 32) {
            float u;
            if (noise) u = __ldcs(noise + src + f);
            else {
                const uint4 r = philox4x32_10(make_uint4(static_cast<unsigned>(t), static_cast<unsigned>(clip0 + b), static_cast<unsigned>(f), 1u), key);
                u = u01(r.x);
            }
            __stcs(out + dst + f, __fadd_rn(__ldcs(in + src + f), __fdiv_rn(__fmul_rn(u, s), 10.0f)));
        }
    }
}

}  // namespace psla
