// Device code of libpsla_frontend.so: the per-clip statistics kernel and the fused fbank tile kernel.
// (Host side / C ABI: psla_frontend.cu.  Per-lane FFT / mel phases: fbank_tile.cuh.)
#pragma once

#include <cuda_runtime.h>

#include "fbank_tile.cuh"

namespace psla {

// ------------------------------------------------------------------------------------------------
// Device-side data
// ------------------------------------------------------------------------------------------------
struct alignas(16) ClipDesc {  // 64 bytes, one per output clip
    long long off1, off2;  // start sample of the primary / partner waveform inside `wav`
    int n1, n2;            // lengths; n2 == 0 means no mixup
    float l1, l2;          // fp32(lambda), fp32(1 - lambda) (the latter rounded from the float64 difference)
    int f0, f1;            // frequency mask [f0, f1)
    int t0, t1;            // time mask [t0, t1)
    int m_eff;             // min(num_frames(n1), target_length): rows that hold real features
    int aligned;           // bit 0: off1 % 4 == 0, bit 1: off2 % 4 == 0 (128-bit loads allowed)
    int pad0, pad1;
};
static_assert(sizeof(ClipDesc) == 64, "ClipDesc layout");

constexpr int kStatSegs = 8;  // partial sums per clip
constexpr int kStatVals = 3;  // sum(w1), sum(w2), sum(w2[0:min(n1,n2)])

// Flat constant-table buffer (floats), copied verbatim into shared memory by every CTA.
constexpr int kTabWin = 0;                       // [416]  0.5 * window (400 used)
constexpr int kTabTw = 416;                      // [1024] inter-pass twiddles (re, im)
constexpr int kTabMelStart = kTabTw + 1024;      // 128 shorts
constexpr int kTabMelCount = kTabMelStart + 64;  // 128 shorts
constexpr int kTabSlotMax = kTabMelCount + 64;   // 4 ints
constexpr int kTabSlotOff = kTabSlotMax + 4;     // 4 ints
constexpr int kTabMelW = kTabSlotOff + 4;        // packed weights
static_assert(kTabMelW % 4 == 0, "table buffer is copied as float4");

// Shared-memory carve-up (floats)
constexpr int kSmD = 0;
constexpr int kSmXS = kSmD + kStageFloats;  // 48: XS[0..33], clip means at [40..42]
constexpr int kSmBS = kSmXS + 48;           // 352
constexpr int kSmEX = kSmBS + 352;
constexpr int kSmTab = kSmEX + kWarpsPerCta * 2 * kExPairFloats;
static_assert((kSmEX * 4) % 16 == 0 && (kSmTab * 4) % 16 == 0, "16-byte alignment of vector regions");

struct FbankParams {
    const float* wav;
    const ClipDesc* desc;
    const double* partials;
    const float* tab;
    float* out;
    int tab_floats;
    int T;
    int nmel;
    float norm_mean, norm_div, norm_rdiv, fill;  // fill = (0 - mean) / div in IEEE fp32 (0 if skip_norm)
    int skip_norm;
    int fast_log;
    float preemph, c1;
};

__device__ __forceinline__ ClipDesc load_desc(const ClipDesc* __restrict__ p) {
    union {
        ClipDesc d;
        int4 v[4];
    } u;
    const int4* src = reinterpret_cast<const int4*>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) u.v[i] = __ldg(src + i);
    return u.d;
}

// 4 consecutive samples [i, i+4) of a waveform of n samples, zeros past the end
__device__ __forceinline__ float4 load4(const float* __restrict__ w, int i, int n, bool aligned) {
    if (aligned && i + 3 < n) return __ldg(reinterpret_cast<const float4*>(w + i));
    float4 v;
    v.x = (i < n) ? __ldg(w + i) : 0.f;
    v.y = (i + 1 < n) ? __ldg(w + i + 1) : 0.f;
    v.z = (i + 2 < n) ? __ldg(w + i + 2) : 0.f;
    v.w = (i + 3 < n) ? __ldg(w + i + 3) : 0.f;
    return v;
}

// sum of w[lo, hi) by one 256-thread CTA's thread `tid` (strided float4s, 4 loads in flight)
__device__ __forceinline__ float strided_sum(const float* __restrict__ w, int lo, int hi, int tid, bool aligned) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int i = lo + 4 * tid;
    if (aligned) {
        for (; i + 3 * 1024 + 3 < hi; i += 4 * 1024) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(w + i));
            const float4 b = __ldg(reinterpret_cast<const float4*>(w + i + 1024));
            const float4 c = __ldg(reinterpret_cast<const float4*>(w + i + 2048));
            const float4 d = __ldg(reinterpret_cast<const float4*>(w + i + 3072));
            s0 += (a.x + a.y) + (a.z + a.w);
            s1 += (b.x + b.y) + (b.z + b.w);
            s2 += (c.x + c.y) + (c.z + c.w);
            s3 += (d.x + d.y) + (d.z + d.w);
        }
    }
    for (; i < hi; i += 1024) {
        const float4 v = load4(w, i, hi, aligned);
        s0 += (v.x + v.y) + (v.z + v.w);
    }
    return (s0 + s1) + (s2 + s3);
}

// ------------------------------------------------------------------------------------------------
// Kernel 1: per-clip sums.  grid = (kStatSegs, B), block = 256.  Deterministic: fixed partition, fixed
// reduction tree; the tile kernel adds the kStatSegs partials in order.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) clip_stats_kernel(const float* __restrict__ wav, const ClipDesc* __restrict__ desc,
                                                         double* __restrict__ partials) {
    const int b = blockIdx.y, g = blockIdx.x, tid = threadIdx.x;
    const ClipDesc d = load_desc(desc + b);
    float s1, s2 = 0.f, s2c = 0.f;
    {
        const int seg = ((d.n1 + 4 * kStatSegs - 1) / (4 * kStatSegs)) * 4;
        const int lo = min(d.n1, g * seg), hi = min(d.n1, lo + seg);
        s1 = strided_sum(wav + d.off1, lo, hi, tid, d.aligned & 1);
    }
    if (d.n2 > 0) {
        const int nmin = min(d.n1, d.n2);
        const int seg = ((d.n2 + 4 * kStatSegs - 1) / (4 * kStatSegs)) * 4;
        const int lo = min(d.n2, g * seg), hi = min(d.n2, lo + seg);
        const bool al = (d.aligned & 2) != 0;
        s2 = strided_sum(wav + d.off2, lo, hi, tid, al);
        if (nmin >= hi) s2c = s2;                                               // segment entirely kept
        else if (nmin > lo) s2c = strided_sum(wav + d.off2, lo, nmin, tid, al);  // the one straddling segment
    }
    __shared__ double red[8][kStatVals];
    double v0 = s1, v1 = s2, v2 = s2c;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
        v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        v2 += __shfl_xor_sync(0xffffffffu, v2, o);
    }
    if ((tid & 31) == 0) {
        red[tid >> 5][0] = v0;
        red[tid >> 5][1] = v1;
        red[tid >> 5][2] = v2;
    }
    __syncthreads();
    if (tid < kStatVals) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += red[w][tid];
        partials[(static_cast<size_t>(b) * kStatSegs + g) * kStatVals + tid] = s;
    }
}

// ------------------------------------------------------------------------------------------------
// Staging of one tile: x = mixed & mean-removed waveform (audioset_dataset.py:92 / :98-117, same fp32
// op order), D = x[j] - p*x[j-1], 16-sample block sums of x, (1-p)*x at frame starts.
// All global loads are issued before the first use (6-12 128-bit loads in flight per thread).
// FULL: the tile lies entirely inside both waveforms and both are 16-byte aligned (no bounds checks).
// ------------------------------------------------------------------------------------------------
struct MixCoef {
    float mean1, mean2, mixmean, l1, l2;
    int n1, nmin;
    bool mix;
};

template <bool FULL>
__device__ __forceinline__ float mix_sample(const MixCoef& m, float a, float c, int s) {
    if (!FULL && s >= m.n1) return 0.f;
    float v = __fsub_rn(a, m.mean1);
    if (m.mix) {
        const float v2 = (FULL || s < m.nmin) ? __fsub_rn(c, m.mean2) : 0.f;
        v = __fsub_rn(__fadd_rn(__fmul_rn(m.l1, v), __fmul_rn(m.l2, v2)), m.mixmean);
    }
    return v;
}

constexpr int kGroups = kTileSamples / 4;                      // 1340 float4 groups per tile
constexpr int kStageIters = (kGroups + kThreads - 1) / kThreads;  // 6

template <bool FULL>
__device__ __forceinline__ void stage_load(const float* __restrict__ w1, const float* __restrict__ w2, const MixCoef& m,
                                           int S0, int tid, int lane, bool al1, bool al2, float4 (&a)[kStageIters],
                                           float4 (&c)[kStageIters], float (&pa)[kStageIters], float (&pc)[kStageIters]) {
#pragma unroll
    for (int it = 0; it < kStageIters; ++it) {
        const int g = tid + it * kThreads;
        const int s = S0 + 4 * g;
        a[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        c[it] = a[it];
        pa[it] = pc[it] = 0.f;
        if (it < kStageIters - 1 || g < kGroups) {
            if (FULL) {
                a[it] = __ldg(reinterpret_cast<const float4*>(w1 + s));
                if (m.mix) c[it] = __ldg(reinterpret_cast<const float4*>(w2 + s));
            } else {
                a[it] = load4(w1, s, m.n1, al1);
                if (m.mix) c[it] = load4(w2, s, m.nmin, al2);
            }
            if (lane == 0 && g > 0) {  // the sample before this warp's first one (pre-emphasis neighbour)
                if (FULL || s - 1 < m.n1) pa[it] = __ldg(w1 + s - 1);
                if (m.mix && (FULL || s - 1 < m.nmin)) pc[it] = __ldg(w2 + s - 1);
            }
        }
    }
}

template <bool FULL>
__device__ __forceinline__ void stage_store(const MixCoef& m, int S0, int tid, int lane, float pe, float c1,
                                            const float4 (&a)[kStageIters], const float4 (&c)[kStageIters],
                                            const float (&pa)[kStageIters], const float (&pc)[kStageIters],
                                            float* __restrict__ D, float* __restrict__ BS, float* __restrict__ XS) {
#pragma unroll
    for (int it = 0; it < kStageIters; ++it) {
        const int g = tid + it * kThreads;
        const int u = 4 * g, s = S0 + u;
        const bool in = (it < kStageIters - 1) || g < kGroups;
        float4 x;
        x.x = mix_sample<FULL>(m, a[it].x, c[it].x, s);
        x.y = mix_sample<FULL>(m, a[it].y, c[it].y, s + 1);
        x.z = mix_sample<FULL>(m, a[it].z, c[it].z, s + 2);
        x.w = mix_sample<FULL>(m, a[it].w, c[it].w, s + 3);
        float prev = __shfl_up_sync(0xffffffffu, x.w, 1);
        if (lane == 0) prev = (g > 0) ? mix_sample<FULL>(m, pa[it], pc[it], s - 1) : 0.f;
        float s4 = (x.x + x.y) + (x.z + x.w);
        s4 += __shfl_xor_sync(0xffffffffu, s4, 1);
        s4 += __shfl_xor_sync(0xffffffffu, s4, 2);
        if (in) {
            float4 dd;
            dd.x = fmaf(-pe, prev, x.x);
            dd.y = fmaf(-pe, x.x, x.y);
            dd.z = fmaf(-pe, x.y, x.z);
            dd.w = fmaf(-pe, x.z, x.w);
            *reinterpret_cast<float4*>(D + dpad(u)) = dd;
            if ((lane & 3) == 0) BS[g >> 2] = s4;
            if (u % kFrameShift == 0 && u / kFrameShift < kFramesPerCta) XS[u / kFrameShift] = c1 * x.x;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Kernel 2: the fused tile kernel.  grid = (ceil(T / 32), B), block = 256, 2 CTAs / SM.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void store_fill_rows(float* __restrict__ out, int fw0, int T, int nmel, int lane, float fill) {
#pragma unroll
    for (int q = 0; q < kFramesPerWarp; ++q) {
        if (fw0 + q >= T) break;
        float* row = out + static_cast<size_t>(fw0 + q) * nmel;
#pragma unroll
        for (int i = 0; i < kMelSlots; ++i)
            if (lane + 32 * i < nmel) __stcs(row + lane + 32 * i, fill);
    }
}

__global__ void __launch_bounds__(kThreads, 2) fbank_tile_kernel(const FbankParams p) {
    extern __shared__ __align__(16) float smem[];
    float* D = smem + kSmD;
    float* XS = smem + kSmXS;
    float* BS = smem + kSmBS;
    float* EX = smem + kSmEX;
    float* TAB = smem + kSmTab;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y;
    const int F0 = blockIdx.x * kFramesPerCta;
    const ClipDesc d = load_desc(p.desc + b);
    float* out = p.out + static_cast<size_t>(b) * p.T * p.nmel;
    const int fw0 = F0 + kFramesPerWarp * warp;

    // rows of this tile that carry computed features: inside the clip and outside the time mask
    const int tile_hi = min(F0 + kFramesPerCta, d.m_eff);
    const int masked = max(0, min(tile_hi, d.t1) - max(F0, d.t0));
    if (tile_hi - F0 - masked <= 0) {  // padding / fully time-masked tile: constant rows only (CTA-uniform)
        store_fill_rows(out, fw0, p.T, p.nmel, lane, p.fill);
        return;
    }
    unsigned live = 0;
#pragma unroll
    for (int q = 0; q < kFramesPerWarp; ++q) {
        const int fa = fw0 + q;
        if (fa < d.m_eff && !(fa >= d.t0 && fa < d.t1)) live |= 1u << q;
    }

    // ---- issue every global load of the tile first ----
    MixCoef m;
    m.mix = d.n2 > 0;
    m.n1 = d.n1;
    m.nmin = min(d.n1, d.n2);
    m.l1 = d.l1;
    m.l2 = d.l2;
    const float* w1 = p.wav + d.off1;
    const float* w2 = p.wav + d.off2;
    const int S0 = F0 * kFrameShift;
    const bool al1 = d.aligned & 1, al2 = (d.aligned & 2) != 0;
    const bool full = al1 && (S0 + kTileSamples <= d.n1) && (!m.mix || (al2 && S0 + kTileSamples <= m.nmin));
    float4 ra[kStageIters], rc[kStageIters];
    float pa[kStageIters], pc[kStageIters];
    if (full) stage_load<true>(w1, w2, m, S0, tid, lane, al1, al2, ra, rc, pa, pc);
    else stage_load<false>(w1, w2, m, S0, tid, lane, al1, al2, ra, rc, pa, pc);

    // ---- constant tables -> shared memory; clip means from the stats kernel's partial sums (warp 0) ----
    {
        const float4* src = reinterpret_cast<const float4*>(p.tab);
        float4* dst = reinterpret_cast<float4*>(TAB);
        for (int i = tid; i < p.tab_floats / 4; i += kThreads) dst[i] = __ldg(src + i);
    }
    if (warp == 0) {
        const double* ps = p.partials + static_cast<size_t>(b) * kStatSegs * kStatVals;
        double s1 = 0.0, s2 = 0.0, s2c = 0.0;
        if (lane < kStatSegs) {
            s1 = ps[lane * kStatVals];
            s2 = ps[lane * kStatVals + 1];
            s2c = ps[lane * kStatVals + 2];
        }
#pragma unroll
        for (int o = kStatSegs / 2; o > 0; o >>= 1) {  // fixed tree => deterministic
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            s2c += __shfl_xor_sync(0xffffffffu, s2c, o);
        }
        if (lane == 0) {
            const float mean1 = static_cast<float>(s1 / d.n1);
            float mean2 = 0.f, mixmean = 0.f;
            if (m.mix) {
                mean2 = static_cast<float>(s2 / d.n2);
                // mean over n1 samples of l1*(w1-mean1) + l2*(w2'-mean2), w2' zero-padded / cut to n1
                const double r1 = s1 - static_cast<double>(d.n1) * mean1;
                const double r2 = s2c - static_cast<double>(m.nmin) * mean2;
                mixmean = static_cast<float>((static_cast<double>(d.l1) * r1 + static_cast<double>(d.l2) * r2) / d.n1);
            }
            XS[40] = mean1;
            XS[41] = mean2;
            XS[42] = mixmean;
        }
    }
    __syncthreads();
    m.mean1 = XS[40];
    m.mean2 = XS[41];
    m.mixmean = XS[42];
    if (full) stage_store<true>(m, S0, tid, lane, p.preemph, p.c1, ra, rc, pa, pc, D, BS, XS);
    else stage_store<false>(m, S0, tid, lane, p.preemph, p.c1, ra, rc, pa, pc, D, BS, XS);
    __syncthreads();

    if (!live) {  // every row of this warp is padding or time-masked (warp-uniform)
        store_fill_rows(out, fw0, p.T, p.nmel, lane, p.fill);
        return;
    }

    float acc[kMelSlots][kFramesPerWarp];
    {
        const float* WIN = TAB + kTabWin;
        const float* TW = TAB + kTabTw;
        MelTableView mt;
        mt.start = reinterpret_cast<const short*>(TAB + kTabMelStart);
        mt.count = reinterpret_cast<const short*>(TAB + kTabMelCount);
        mt.slot_max = reinterpret_cast<const int*>(TAB + kTabSlotMax);
        mt.slot_off = reinterpret_cast<const int*>(TAB + kTabSlotOff);
        mt.w = TAB + kTabMelW;

        // per-frame means (kaldi.py:183-186) from the 16-sample block sums; lanes 0..3 own one frame each
        float cmu = 0.f;
        if (lane < kFramesPerWarp) {
            const float* bs = BS + 10 * (kFramesPerWarp * warp + lane);
            float sum = 0.f;
#pragma unroll
            for (int k = 0; k < 25; ++k) sum += bs[k];
            cmu = p.c1 * (sum * (1.0f / kFrameLen));
        }
        const int h = lane >> 4, n2 = lane & 15;
        const float cmu_x = __shfl_sync(0xffffffffu, cmu, 2 * h);
        const float cmu_y = __shfl_sync(0xffffffffu, cmu, 2 * h + 1);

        float* EXw = EX + warp * 2 * kExPairFloats;
        float* EXp = EXw + h * kExPairFloats;
        pair_pass1(n2, kFramesPerWarp * warp + 2 * h, D, XS, WIN, TW, EXp, cmu_x, cmu_y);
        __syncwarp();
        {
            float ar[16], ai[16], br[16], bi[16];
            pair_pass2_fft(n2, EXp, ar, ai, br, bi);
            __syncwarp();  // P aliases EX: all reads done before any write
            pair_power_store(n2, ar, ai, br, bi, EXw + (2 * h) * kPRowFloats, EXw + (2 * h + 1) * kPRowFloats);
        }
        __syncwarp();
        warp_mel(lane, EXw, mt, acc);
    }

    // ---- epilogue: log floor, SpecAugment fill, normalisation, coalesced streaming store ----
    const float eps = 1.1920928955078125e-07f;    // kaldi.py:22
    const float log_eps = -15.9423847198486328f;  // fp32 log(eps): what max(x, eps).log() gives on the floor
    const float mean = p.skip_norm ? 0.f : p.norm_mean;
    const float div = p.skip_norm ? 1.f : p.norm_div;
    const float rdiv = p.skip_norm ? 1.f : p.norm_rdiv;
    bool keep[kMelSlots];  // column survives the frequency mask and exists
#pragma unroll
    for (int i = 0; i < kMelSlots; ++i) keep[i] = !(lane + 32 * i >= d.f0 && lane + 32 * i < d.f1);
    float* orow = out + static_cast<size_t>(fw0) * p.nmel + lane;
    if (p.fast_log) {
#pragma unroll
        for (int q = 0; q < kFramesPerWarp; ++q) {
            const bool rlive = (live >> q) & 1u;
            if (fw0 + q < p.T) {
#pragma unroll
                for (int i = 0; i < kMelSlots; ++i) {
                    const float e = acc[i][q];
                    const float lg = (e > eps) ? __logf(e) : log_eps;
                    // (lg - mean) / div as a Newton-corrected multiply by 1/div: correctly rounded but for rare
                    // last-bit cases; live cells only need 1e-3, the fill constant is the exact IEEE quotient
                    const float x = __fsub_rn(lg, mean);
                    const float q0 = x * rdiv;
                    const float v = fmaf(fmaf(-q0, div, x), rdiv, q0);
                    if (lane + 32 * i < p.nmel) __stcs(orow + 32 * i, (rlive && keep[i]) ? v : p.fill);
                }
            }
            orow += p.nmel;
        }
    } else {
#pragma unroll
        for (int q = 0; q < kFramesPerWarp; ++q) {
            const bool rlive = (live >> q) & 1u;
            if (fw0 + q < p.T) {
#pragma unroll
                for (int i = 0; i < kMelSlots; ++i) {
                    const float lg = logf(fmaxf(acc[i][q], eps));
                    const float v = p.skip_norm ? lg : __fdiv_rn(__fsub_rn(lg, p.norm_mean), p.norm_div);
                    if (lane + 32 * i < p.nmel) __stcs(orow + 32 * i, (rlive && keep[i]) ? v : p.fill);
                }
            }
            orow += p.nmel;
        }
    }
}

}  // namespace psla
