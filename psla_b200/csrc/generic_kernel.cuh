// Geometry-generic frontend kernel: the same recipe as fbank_stream_kernel for any sample rate whose 25 ms window fits an
// FFT of at most 4096 points (torchaudio/compliance/kaldi.py:138-140: window_size = int(sr * 0.025), window_shift =
// int(sr * 0.010), padded_window_size = next power of two) -- 8 kHz (200 / 80 / 256) ... 48 kHz (1200 / 480 / 2048).
// "generalising the kernel beyond 16 kHz".  Every recipe of the reference resamples to 16 kHz
// (egs/fsd50k/prep_fsd.py:17-35), so this path is built for correctness and coverage, not for the roofline: one warp per
// frame, a radix-4 Stockham FFT in shared memory, the clip means always taken from clip_stats_kernel (the reference's expression
// verbatim, audioset_dataset.py:92,98-99,116-117).  The 16 kHz geometry keeps the specialised streaming kernel;
// psla_frontend_set_generic_path(handle, 1) runs it through this kernel too (a second, independent implementation the tests
// compare the streaming kernel with).
#pragma once

#include "stream_kernel.cuh"

namespace psla {

constexpr int kGenWarps = 4;          // warps per CTA
constexpr int kGenFramesPerWarp = 4;  // consecutive output rows per warp
constexpr int kGenFramesPerCta = kGenWarps * kGenFramesPerWarp;
constexpr int kGenMinFft = 64, kGenMaxFft = 4096;

struct GenericParams {
    const void* wav;          // fp32 samples or int16 PCM (Pcm<S>)
    const ClipDesc* desc;     // [n_clips] (device memory: this path never uses parameter-space descriptors)
    const double* partials;   // clip_stats_kernel's sums, exact mode (every clip carries kFlagStats)
    const float* window;      // [L] analysis window
    const float2* tw;         // [M]        exp(-2 pi i j / M),  M = N / 2
    const float2* twn;        // [M/2 + 1]  exp(-2 pi i k / N)
    const int2* mel_range;    // [nmel] (first, one-past-last) spectrum bin with non-zero weight
    const float* mel_w;       // [nmel][M] dense weights (kaldi.py:436-511)
    float* out;               // [n_clips][T][nmel]
    int n_clips, T, nmel;
    int L, S, N, log2M;       // frame length, frame shift, FFT size, log2(N / 2)
    float norm_mean, norm_div, fill, floor_val, preemph;
    int skip_norm;
};

// per-CTA clip constants (audioset_dataset.py:92,98-99,116-117 evaluated from the statistics kernel's sums, like
// enter_clip<EXACT = true> in stream_kernel.cuh)
struct GenClip {
    long long off1, off2;
    int n1, nmin, mix, m_eff;
    int f0, f1, t0, t1;
    float l1, l2, mean1, mean2, mixmean;
};

// shared memory: GenClip, then per warp [N floats: staged samples, later the power spectrum][N floats: M complex points]
__host__ __device__ constexpr size_t generic_smem_bytes(int N) { return 128 + static_cast<size_t>(kGenWarps) * 2 * N * sizeof(float); }

template <class S>
__global__ void __launch_bounds__(kGenWarps * 32) fbank_generic_kernel(const GenericParams p) {
    extern __shared__ __align__(16) unsigned char gen_smem[];
    GenClip& gc = *reinterpret_cast<GenClip*>(gen_smem);
    const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* stage = reinterpret_cast<float*>(gen_smem + 128) + static_cast<size_t>(warp) * 2 * p.N;
    float2* Z = reinterpret_cast<float2*>(stage + p.N);
    const int M = p.N >> 1;

    if (warp == 0) {
        const ClipDesc d = load_desc(p.desc + b);
        const bool mix = d.n2 > 0;
        const int nmin = min(d.n1, d.n2);
        const double* ps = p.partials + static_cast<size_t>(b) * kStatSegs * kStatVals;
        double s1 = 0.0, s2 = 0.0, s2c = 0.0;
        if (lane < kStatSegs) {
            s1 = __ldcg(ps + lane * kStatVals);
            s2 = __ldcg(ps + lane * kStatVals + 1);
            s2c = __ldcg(ps + lane * kStatVals + 2);
        }
#pragma unroll
        for (int o = kStatSegs / 2; o > 0; o >>= 1) {
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            s2c += __shfl_xor_sync(0xffffffffu, s2c, o);
        }
        if (lane == 0) {
            const float mean1 = static_cast<float>(s1 / d.n1);
            float mean2 = 0.f, mixmean = 0.f;
            if (mix) {
                mean2 = static_cast<float>(s2 / d.n2);
                const double r1 = s1 - static_cast<double>(d.n1) * mean1;
                const double r2 = s2c - static_cast<double>(nmin) * mean2;
                mixmean = static_cast<float>((static_cast<double>(d.l1) * r1 + static_cast<double>(d.l2) * r2) / d.n1);
            }
            gc.off1 = d.off1;
            gc.off2 = d.off2;
            gc.n1 = d.n1;
            gc.nmin = nmin;
            gc.mix = mix ? 1 : 0;
            gc.m_eff = d.m_eff;
            gc.f0 = d.f0;
            gc.f1 = d.f1;
            gc.t0 = d.t0;
            gc.t1 = d.t1;
            gc.l1 = d.l1;
            gc.l2 = d.l2;
            gc.mean1 = mean1;
            gc.mean2 = mean2;
            gc.mixmean = mixmean;
        }
    }
    __syncthreads();
    const S* w1 = static_cast<const S*>(p.wav) + gc.off1;
    const S* w2 = static_cast<const S*>(p.wav) + gc.off2;
    const bool mix = gc.mix != 0;
    const float eps = 1.1920928955078125e-07f;  // kaldi.py:22
    float* out_clip = p.out + static_cast<size_t>(b) * p.T * p.nmel;

    for (int i = 0; i < kGenFramesPerWarp; ++i) {
        const int t = blockIdx.x * kGenFramesPerCta + warp * kGenFramesPerWarp + i;
        if (t >= p.T) break;
        float* row = out_clip + static_cast<size_t>(t) * p.nmel;
        if (t >= gc.m_eff || (t >= gc.t0 && t < gc.t1)) {  // padding (audioset_dataset.py:122-128) or time mask (:184-185): 0.0, then normalised
            for (int m = lane; m < p.nmel; m += 32) __stcs(row + m, p.fill);
            continue;
        }
        // ---- the frame's samples: clip DC removal and mixup in the reference's fp32 order (mix_sample_exact) ----
        const long long u0 = static_cast<long long>(t) * p.S;
        float sum = 0.f;
        for (int j = lane; j < p.L; j += 32) {
            const long long u = u0 + j;  // < n1: t < m_eff
            float x = __fsub_rn(Pcm<S>::load1(w1 + u), gc.mean1);
            if (mix) {
                const float v2 = (u < gc.nmin) ? __fsub_rn(Pcm<S>::load1(w2 + u), gc.mean2) : 0.f;
                x = __fsub_rn(__fadd_rn(__fmul_rn(gc.l1, x), __fmul_rn(gc.l2, v2)), gc.mixmean);
            }
            stage[j] = x;
            sum += x;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const float mu = sum / static_cast<float>(p.L);  // per-frame DC removal, kaldi.py:183-186
        __syncwarp();
        // ---- pre-emphasis with replicate padding (:193-198), window (:200), zero padding to N (:204-206); stored as the M
        //      complex points z[c] = (y[2c], y[2c+1]) in natural order ----
        float* Zf = reinterpret_cast<float*>(Z);
        for (int j = lane; j < p.N; j += 32) {
            float y = 0.f;
            if (j < p.L) {
                const float s = stage[j] - mu;
                const float sp = stage[j > 0 ? j - 1 : 0] - mu;
                y = (s - p.preemph * sp) * __ldg(p.window + j);
            }
            Zf[j] = y;
        }
        __syncwarp();
        // ---- M-point complex FFT: Stockham autosort (no bit reversal; every pass reads one buffer with unit stride and
        //      writes the other), radix-4 passes and one radix-2 pass when log2(M) is odd.  Pass with Ns points already
        //      combined, radix R:  v[r] = in[j + r M/R] * W_(Ns R)^(r k),  k = j mod Ns;  out[(j - k) R + k + r Ns] = DFT_R(v)[r]
        float2* src = Z;
        float2* dst = reinterpret_cast<float2*>(stage);  // the staged samples are dead
        int Ns = 1;
        for (; Ns * 4 <= M; Ns *= 4) {
            const int q4 = M >> 2, tstep = M / (4 * Ns);
#pragma unroll 2
            for (int j = lane; j < q4; j += 32) {
                const int k = j & (Ns - 1);
                const float2 a0 = src[j];
                float2 a1 = src[j + q4], a2 = src[j + 2 * q4], a3 = src[j + 3 * q4];
                if (Ns > 1) {
                    const float2 w1 = __ldg(p.tw + k * tstep), w2 = __ldg(p.tw + 2 * k * tstep), w3 = __ldg(p.tw + 3 * k * tstep);
                    a1 = make_float2(a1.x * w1.x - a1.y * w1.y, a1.x * w1.y + a1.y * w1.x);
                    a2 = make_float2(a2.x * w2.x - a2.y * w2.y, a2.x * w2.y + a2.y * w2.x);
                    a3 = make_float2(a3.x * w3.x - a3.y * w3.y, a3.x * w3.y + a3.y * w3.x);
                }
                const float2 t0 = make_float2(a0.x + a2.x, a0.y + a2.y), t1 = make_float2(a0.x - a2.x, a0.y - a2.y);
                const float2 t2 = make_float2(a1.x + a3.x, a1.y + a3.y);
                const float2 t3 = make_float2(a1.y - a3.y, -(a1.x - a3.x));  // -i (a1 - a3)
                const int j0 = ((j - k) << 2) + k;
                dst[j0] = make_float2(t0.x + t2.x, t0.y + t2.y);
                dst[j0 + Ns] = make_float2(t1.x + t3.x, t1.y + t3.y);
                dst[j0 + 2 * Ns] = make_float2(t0.x - t2.x, t0.y - t2.y);
                dst[j0 + 3 * Ns] = make_float2(t1.x - t3.x, t1.y - t3.y);
            }
            __syncwarp();
            float2* t = src;
            src = dst;
            dst = t;
        }
        if (Ns < M) {  // Ns == M / 2: the last radix-2 pass, k = j
            const int q2 = M >> 1;
            for (int j = lane; j < q2; j += 32) {
                const float2 w = __ldg(p.tw + j);
                const float2 a = src[j], c = src[j + q2];
                const float vr = c.x * w.x - c.y * w.y, vi = c.x * w.y + c.y * w.x;
                dst[j] = make_float2(a.x + vr, a.y + vi);
                dst[j + q2] = make_float2(a.x - vr, a.y - vi);
            }
            __syncwarp();
            float2* t = src;
            src = dst;
            dst = t;
        }
        // ---- spectrum of the real frame from the packed transform (now in `src`), power (:616-618), into the other buffer ----
        //   E = (Z[k] + conj(Z[M-k])) / 2,  O = (Z[k] - conj(Z[M-k])) / 2i,  X[k] = E + W_N^k O,  X[M-k] = conj(E - W_N^k O)
        float* P = reinterpret_cast<float*>(dst);
        for (int k = lane; k <= (M >> 1); k += 32) {
            const float2 a = src[k], zb = src[(M - k) & (M - 1)];
            const float er = 0.5f * (a.x + zb.x), ei = 0.5f * (a.y - zb.y);
            const float orr = 0.5f * (a.y + zb.y), oi = -0.5f * (a.x - zb.x);
            const float2 w = __ldg(p.twn + k);
            const float tr = orr * w.x - oi * w.y, ti = orr * w.y + oi * w.x;
            const float xr = er + tr, xi = ei + ti, yr = er - tr, yi = ei - ti;
            P[k] = xr * xr + xi * xi;
            if (k != 0 && k != M - k) P[M - k] = yr * yr + yi * yi;
        }
        __syncwarp();
        // ---- mel projection (:630), log floor (:633), SpecAugment frequency mask, normalisation (audioset_dataset.py:191-192) ----
        for (int m = lane; m < p.nmel; m += 32) {
            const int2 r = __ldg(p.mel_range + m);
            const float* wrow = p.mel_w + static_cast<size_t>(m) * M;
            float acc = 0.f;
            for (int k = r.x; k < r.y; ++k) acc = fmaf(__ldg(wrow + k), P[k], acc);
            float v;
            if (m >= gc.f0 && m < gc.f1) v = p.fill;
            else if (!(acc > eps)) v = p.floor_val;
            else {
                v = logf(acc);
                if (!p.skip_norm) v = __fdiv_rn(__fsub_rn(v, p.norm_mean), p.norm_div);
            }
            __stcs(row + m, v);
        }
        __syncwarp();  // both buffers are rewritten by the next frame
    }
}

}  // namespace psla
