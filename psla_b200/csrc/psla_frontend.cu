// libpsla_frontend.so -- fused Kaldi-compatible log-mel frontend for sm_100a (B200).
//
// Replaces, for a batch of clips on one GPU, what the reference computes per item on CPU dataloader
// workers: AudiosetDataset._wav2fbank + the SpecAugment / normalisation tail of __getitem__
// (src/dataloaders/audioset_dataset.py:88-137, 177-192), whose arithmetic is
// torchaudio.compliance.kaldi.fbank (torchaudio/compliance/kaldi.py:514-645).
//
// Two kernels per forward():
//   clip_stats_kernel  per-clip waveform sums (deterministic two-level reduction) for the whole-clip DC
//                      removal and the mixup mean (audioset_dataset.py:92,98-99,117)
//   fbank_tile_kernel  one CTA = 32 consecutive frames of one clip: stage the (mixed, mean-removed,
//                      pre-emphasised) samples once in shared memory, then every warp turns 4 frames into
//                      4 normalised output rows (see fbank_tile.cuh for the per-lane algorithm).
// No cuFFT, no tensor cores (the mel projection is 504 non-zeros, not a dense contraction), no CPU path.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/psla_frontend.h"
#include "fbank_tile.cuh"
#include "tables.h"

namespace psla {

// ------------------------------------------------------------------------------------------------
// Device-side data
// ------------------------------------------------------------------------------------------------
struct ClipDesc {  // 56 bytes, one per output clip
    long long off1, off2;  // start sample of the primary / partner waveform inside `wav`
    int n1, n2;            // lengths; n2 == 0 means no mixup
    float l1, l2;          // fp32(lambda), fp32(1 - lambda) (the latter rounded from the float64 difference)
    int f0, f1;            // frequency mask [f0, f1)
    int t0, t1;            // time mask [t0, t1)
    int m_eff;             // min(num_frames(n1), target_length): rows that hold real features
    int aligned;           // bit 0: off1 % 4 == 0, bit 1: off2 % 4 == 0 (float4 loads allowed)
};
static_assert(sizeof(ClipDesc) == 56, "ClipDesc layout");

constexpr int kStatSegs = 8;  // partial sums per clip
constexpr int kStatVals = 3;  // sum(w1), sum(w2), sum(w2[0:min(n1,n2)])

// Flat constant-table buffer (floats), copied verbatim into shared memory by every CTA.
constexpr int kTabWin = 0;                      // [416]  0.5 * window (400 used)
constexpr int kTabTw = 416;                     // [1024] inter-pass twiddles (re, im)
constexpr int kTabMelStart = kTabTw + 1024;     // 128 shorts
constexpr int kTabMelCount = kTabMelStart + 64;  // 128 shorts
constexpr int kTabSlotMax = kTabMelCount + 64;  // 4 ints
constexpr int kTabSlotOff = kTabSlotMax + 4;    // 4 ints
constexpr int kTabMelW = kTabSlotOff + 4;       // packed weights
static_assert(kTabMelW % 4 == 0, "table buffer is copied as float4");

// Shared-memory carve-up (floats)
constexpr int kSmD = 0;
constexpr int kSmXS = kSmD + kStageFloats;  // 48: XS[0..33], stats at [40..43]
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
    float norm_mean, norm_div;
    int skip_norm;
    int fast_log;
    float preemph, c1;
};

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

// ------------------------------------------------------------------------------------------------
// Kernel 1: per-clip sums.  grid = (kStatSegs, B), block = 256.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) clip_stats_kernel(const float* __restrict__ wav, const ClipDesc* __restrict__ desc,
                                                         double* __restrict__ partials) {
    const int b = blockIdx.y, g = blockIdx.x, tid = threadIdx.x;
    const ClipDesc d = desc[b];
    float s1 = 0.f, s2 = 0.f, s2c = 0.f;
    {
        const int seg = ((d.n1 + 4 * kStatSegs - 1) / (4 * kStatSegs)) * 4;
        const int lo = g * seg, hi = min(d.n1, lo + seg);
        const float* w = wav + d.off1;
        for (int i = lo + 4 * tid; i < hi; i += 4 * 256) {
            const float4 v = load4(w, i, hi, d.aligned & 1);
            s1 += (v.x + v.y) + (v.z + v.w);
        }
    }
    if (d.n2 > 0) {
        const int nmin = min(d.n1, d.n2);
        const int seg = ((d.n2 + 4 * kStatSegs - 1) / (4 * kStatSegs)) * 4;
        const int lo = g * seg, hi = min(d.n2, lo + seg);
        const float* w = wav + d.off2;
        for (int i = lo + 4 * tid; i < hi; i += 4 * 256) {
            const float4 v = load4(w, i, hi, (d.aligned & 2) != 0);
            s2 += (v.x + v.y) + (v.z + v.w);
            if (i + 3 < nmin) {
                s2c += (v.x + v.y) + (v.z + v.w);
            } else {
                s2c += (i < nmin ? v.x : 0.f) + (i + 1 < nmin ? v.y : 0.f) + (i + 2 < nmin ? v.z : 0.f);
            }
        }
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
        for (int w = 0; w < 8; ++w) s += red[w][tid];
        partials[(static_cast<size_t>(b) * kStatSegs + g) * kStatVals + tid] = s;
    }
}

// ------------------------------------------------------------------------------------------------
// Kernel 2: the fused tile kernel.  grid = (ceil(T / 32), B), block = 256, 2 CTAs / SM.
// ------------------------------------------------------------------------------------------------
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
    const ClipDesc d = p.desc[b];
    const float fill = p.skip_norm ? 0.f : __fdiv_rn(__fsub_rn(0.f, p.norm_mean), p.norm_div);
    float* out = p.out + static_cast<size_t>(b) * p.T * p.nmel;

    // rows of this tile that carry computed features: inside the clip and outside the time mask
    const int tile_hi = min(F0 + kFramesPerCta, d.m_eff);
    const int masked = max(0, min(tile_hi, d.t1) - max(F0, d.t0));
    const bool tile_live = (tile_hi - F0 - masked) > 0;

    const int fw0 = F0 + kFramesPerWarp * warp;
    unsigned live = 0;
#pragma unroll
    for (int q = 0; q < kFramesPerWarp; ++q) {
        const int fa = fw0 + q;
        if (fa < d.m_eff && !(fa >= d.t0 && fa < d.t1)) live |= 1u << q;
    }

    float acc[kMelSlots][kFramesPerWarp];
#pragma unroll
    for (int i = 0; i < kMelSlots; ++i)
#pragma unroll
        for (int q = 0; q < kFramesPerWarp; ++q) acc[i][q] = 0.f;

    if (tile_live) {
        // ---- constant tables -> shared memory; clip statistics ----
        {
            const float4* src = reinterpret_cast<const float4*>(p.tab);
            float4* dst = reinterpret_cast<float4*>(TAB);
            for (int i = tid; i < p.tab_floats / 4; i += kThreads) dst[i] = __ldg(src + i);
        }
        if (tid == 0) {
            const double* ps = p.partials + static_cast<size_t>(b) * kStatSegs * kStatVals;
            double s1 = 0.0, s2 = 0.0, s2c = 0.0;
            for (int g = 0; g < kStatSegs; ++g) {
                s1 += ps[g * kStatVals];
                s2 += ps[g * kStatVals + 1];
                s2c += ps[g * kStatVals + 2];
            }
            const float mean1 = static_cast<float>(s1 / d.n1);
            float mean2 = 0.f, mixmean = 0.f;
            if (d.n2 > 0) {
                mean2 = static_cast<float>(s2 / d.n2);
                const int nmin = min(d.n1, d.n2);
                // mean over n1 samples of l1*(w1-mean1) + l2*(w2'-mean2), w2' zero-padded / cut to n1
                const double r1 = s1 - static_cast<double>(d.n1) * mean1;
                const double r2 = s2c - static_cast<double>(nmin) * mean2;
                mixmean = static_cast<float>((static_cast<double>(d.l1) * r1 + static_cast<double>(d.l2) * r2) / d.n1);
            }
            XS[40] = mean1;
            XS[41] = mean2;
            XS[42] = mixmean;
        }
        __syncthreads();

        // ---- stage samples: x = mixed & mean-removed waveform, D = x[j] - p*x[j-1] ----
        const float mean1 = XS[40], mean2 = XS[41], mixmean = XS[42];
        const bool mix = d.n2 > 0;
        const int nmin = min(d.n1, d.n2);
        const float* w1 = p.wav + d.off1;
        const float* w2 = p.wav + d.off2;
        const int S0 = F0 * kFrameShift;
        const float pe = p.preemph, c1 = p.c1;
        auto xval = [&](float a, float c, int s) -> float {  // audioset_dataset.py:92 / :98-117, same op order
            if (s >= d.n1) return 0.f;
            float v = __fsub_rn(a, mean1);
            if (mix) {
                const float v2 = (s < nmin) ? __fsub_rn(c, mean2) : 0.f;
                v = __fsub_rn(__fadd_rn(__fmul_rn(d.l1, v), __fmul_rn(d.l2, v2)), mixmean);
            }
            return v;
        };
        constexpr int kGroups = kTileSamples / 4;  // 1340
#pragma unroll 1
        for (int it = 0; it < (kGroups + kThreads - 1) / kThreads; ++it) {
            const int g = tid + it * kThreads;
            const int u = 4 * g, s = S0 + u;
            const bool in = g < kGroups;
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (in) {
                const float4 a = load4(w1, s, d.n1, d.aligned & 1);
                float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
                if (mix) c = load4(w2, s, nmin, (d.aligned & 2) != 0);
                x.x = xval(a.x, c.x, s);
                x.y = xval(a.y, c.y, s + 1);
                x.z = xval(a.z, c.z, s + 2);
                x.w = xval(a.w, c.w, s + 3);
            }
            float prev = __shfl_up_sync(0xffffffffu, x.w, 1);
            if (lane == 0) {
                prev = 0.f;
                if (in && u > 0 && s - 1 < d.n1) {
                    const float a = __ldg(w1 + s - 1);
                    const float c = (mix && s - 1 < nmin) ? __ldg(w2 + s - 1) : 0.f;
                    prev = xval(a, c, s - 1);
                }
            }
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
        __syncthreads();

        if (live) {
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
                cmu = c1 * (sum * (1.0f / kFrameLen));
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
    }

    // ---- epilogue: log floor, SpecAugment fill, normalisation, coalesced streaming store ----
    const float eps = 1.1920928955078125e-07f;   // kaldi.py:22
    const float log_eps = -15.9423847198486328f;  // fp32 log(eps): what max(x, eps).log() gives on the floor
    const float rdiv = __frcp_rn(p.norm_div);
#pragma unroll
    for (int i = 0; i < kMelSlots; ++i) {
        const int j = lane + 32 * i;
        if (j >= p.nmel) continue;
        const bool fmasked = (j >= d.f0 && j < d.f1);
#pragma unroll
        for (int q = 0; q < kFramesPerWarp; ++q) {
            const int row = fw0 + q;
            if (row >= p.T) continue;
            float v = fill;
            if (((live >> q) & 1u) && !fmasked) {
                const float e = acc[i][q];
                float lg;
                if (p.fast_log) {
                    lg = (e > eps) ? __logf(e) : log_eps;
                } else {
                    lg = logf(fmaxf(e, eps));
                }
                if (p.skip_norm) {
                    v = lg;
                } else if (p.fast_log) {
                    // (lg - mean) / div as one Newton-corrected multiply by 1/div (correctly rounded but for
                    // rare last-bit cases; live cells only need 1e-3, the fill constant above is exact)
                    const float x = __fsub_rn(lg, p.norm_mean);
                    const float q0 = x * rdiv;
                    v = fmaf(fmaf(-q0, p.norm_div, x), rdiv, q0);
                } else {
                    v = __fdiv_rn(__fsub_rn(lg, p.norm_mean), p.norm_div);
                }
            }
            __stcs(out + static_cast<size_t>(row) * p.nmel + j, v);
        }
    }
}

}  // namespace psla

// ================================================================================================
// Host side: handle, tables, validation, launches (C ABI)
// ================================================================================================
using namespace psla;

struct psla_frontend {
    int device = 0;
    int sample_rate = 16000;
    int nmel = 128;
    float preemph = 0.97f;
    float low_freq = 20.f, high_freq = 0.f;
    std::vector<float> window;     // [400] analysis window (not halved)
    std::vector<float> mel_dense;  // [nmel][256]
    float* d_tab = nullptr;
    int tab_floats = 0;
    size_t smem_bytes = 0;
    int fast_log = 0;
    int64_t launches = 0;
    int last_cuda = 0;
    // pinned descriptor staging ring
    struct Slot {
        void* host = nullptr;
        size_t cap = 0;
        cudaEvent_t ev = nullptr;
        bool pending = false;
    } slots[4];
    int next_slot = 0;
    // forward_host device buffers
    float* d_wav = nullptr;
    size_t d_wav_cap = 0;
    float* d_out = nullptr;
    size_t d_out_cap = 0;
    void* d_ws = nullptr;
    size_t d_ws_cap = 0;
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

#define PSLA_CUDA(h, call)                       \
    do {                                         \
        cudaError_t e_ = (call);                 \
        if (e_ != cudaSuccess) {                 \
            (h)->last_cuda = static_cast<int>(e_); \
            return PSLA_E_CUDA;                  \
        }                                        \
    } while (0)

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int upload_tables(psla_frontend* h) {
    PackedMel pm;
    if (!pack_mel(h->mel_dense.data(), h->nmel, kNfft, &pm)) return PSLA_E_UNSUPPORTED;
    const int total = static_cast<int>(align_up(kTabMelW + pm.w.size(), 4));
    std::vector<float> buf(total, 0.f);
    for (int j = 0; j < kFrameLen; ++j) buf[kTabWin + j] = 0.5f * h->window[j];
    build_twiddles(buf.data() + kTabTw);
    std::memcpy(buf.data() + kTabMelStart, pm.start, sizeof(pm.start));
    std::memcpy(buf.data() + kTabMelCount, pm.count, sizeof(pm.count));
    std::memcpy(buf.data() + kTabSlotMax, pm.slot_max, sizeof(pm.slot_max));
    std::memcpy(buf.data() + kTabSlotOff, pm.slot_off, sizeof(pm.slot_off));
    std::memcpy(buf.data() + kTabMelW, pm.w.data(), pm.w.size() * sizeof(float));
    const size_t smem = static_cast<size_t>(kSmTab + total) * sizeof(float);
    if (smem > 113 * 1024) return PSLA_E_UNSUPPORTED;  // keep 2 CTAs / SM
    if (h->d_tab) {
        PSLA_CUDA(h, cudaDeviceSynchronize());
        cudaFree(h->d_tab);
        h->d_tab = nullptr;
    }
    PSLA_CUDA(h, cudaMalloc(&h->d_tab, total * sizeof(float)));
    PSLA_CUDA(h, cudaMemcpy(h->d_tab, buf.data(), total * sizeof(float), cudaMemcpyHostToDevice));
    h->tab_floats = total;
    h->smem_bytes = smem;
    PSLA_CUDA(h, cudaFuncSetAttribute(fbank_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    PSLA_CUDA(h, cudaFuncSetAttribute(fbank_tile_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    return PSLA_OK;
}

int64_t frames_of(int64_t n) { return n < kFrameLen ? -1 : 1 + (n - kFrameLen) / kFrameShift; }

}  // namespace

extern "C" {

const char* psla_version(void) { return "psla_b200 frontend 0.1 (sm_100a)"; }

const char* psla_strerror(int s) {
    switch (s) {
        case PSLA_OK: return "ok";
        case PSLA_E_INVALID: return "invalid argument";
        case PSLA_E_UNSUPPORTED: return "unsupported frontend geometry (need 16 kHz, 25 ms / 10 ms, <= 128 mel bins)";
        case PSLA_E_SHORT_CLIP: return "waveform shorter than one 25 ms frame (reference: AssertionError in kaldi.fbank)";
        case PSLA_E_MASK: return "mask offsets out of range";
        case PSLA_E_WORKSPACE: return "workspace too small or misaligned";
        case PSLA_E_CUDA: return "CUDA runtime error";
        case PSLA_E_NOMEM: return "out of memory";
        default: return "unknown status";
    }
}

int psla_last_cuda_error(const psla_frontend_t* h) { return h ? h->last_cuda : 0; }

int psla_frontend_create(psla_frontend_t** handle, int device, int sample_rate, int num_mel_bins, float frame_length_ms,
                         float frame_shift_ms, float preemph, float low_freq, float high_freq) {
    if (!handle) return PSLA_E_INVALID;
    *handle = nullptr;
    // int(sample_frequency * frame_shift * 0.001), kaldi.py:138-139
    const int shift = static_cast<int>(static_cast<double>(sample_rate) * static_cast<double>(frame_shift_ms) * 0.001);
    const int size = static_cast<int>(static_cast<double>(sample_rate) * static_cast<double>(frame_length_ms) * 0.001);
    if (shift != kFrameShift || size != kFrameLen) return PSLA_E_UNSUPPORTED;
    if (num_mel_bins < 4 || num_mel_bins > kMaxMel) return PSLA_E_UNSUPPORTED;  // kaldi.py:452 needs > 3
    if (!(preemph >= 0.f && preemph <= 1.f)) return PSLA_E_INVALID;              // kaldi.py:149
    const float nyq = 0.5f * sample_rate;
    const float hi = high_freq <= 0.f ? high_freq + nyq : high_freq;
    if (!(low_freq >= 0.f && low_freq < nyq && hi > 0.f && hi <= nyq && low_freq < hi)) return PSLA_E_INVALID;  // :460-462
    psla_frontend* h = new (std::nothrow) psla_frontend();
    if (!h) return PSLA_E_NOMEM;
    h->device = device;
    h->sample_rate = sample_rate;
    h->nmel = num_mel_bins;
    h->preemph = preemph;
    h->low_freq = low_freq;
    h->high_freq = high_freq;
    h->window.resize(kFrameLen);
    std::vector<float> half(kFrameLen);
    build_window_half(half.data(), kFrameLen);
    for (int j = 0; j < kFrameLen; ++j) h->window[j] = 2.f * half[j];
    h->mel_dense.resize(static_cast<size_t>(num_mel_bins) * (kNfft / 2));
    build_mel_dense(h->mel_dense.data(), num_mel_bins, kNfft, static_cast<float>(sample_rate), low_freq, high_freq);
    DeviceGuard guard(device);
    if (!guard.ok) {
        delete h;
        return PSLA_E_CUDA;
    }
    int st = upload_tables(h);
    if (st == PSLA_OK) {
        for (auto& s : h->slots)
            if (cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming) != cudaSuccess) st = PSLA_E_CUDA;
    }
    if (st != PSLA_OK) {
        psla_frontend_destroy(h);
        return st;
    }
    *handle = h;
    return PSLA_OK;
}

void psla_frontend_destroy(psla_frontend_t* h) {
    if (!h) return;
    DeviceGuard guard(h->device);
    cudaDeviceSynchronize();
    for (auto& s : h->slots) {
        if (s.host) cudaFreeHost(s.host);
        if (s.ev) cudaEventDestroy(s.ev);
    }
    if (h->d_tab) cudaFree(h->d_tab);
    if (h->d_wav) cudaFree(h->d_wav);
    if (h->d_out) cudaFree(h->d_out);
    if (h->d_ws) cudaFree(h->d_ws);
    delete h;
}

int psla_frontend_set_tables(psla_frontend_t* h, const float* window, const float* mel) {
    if (!h) return PSLA_E_INVALID;
    std::vector<float> w0 = h->window, m0 = h->mel_dense;
    if (window) std::memcpy(h->window.data(), window, sizeof(float) * kFrameLen);
    if (mel) std::memcpy(h->mel_dense.data(), mel, sizeof(float) * h->mel_dense.size());
    DeviceGuard guard(h->device);
    if (!guard.ok) return PSLA_E_CUDA;
    const int st = upload_tables(h);
    if (st != PSLA_OK) {  // keep the previous, valid tables
        h->window = w0;
        h->mel_dense = m0;
        upload_tables(h);
    }
    return st;
}

int psla_frontend_get_tables(const psla_frontend_t* h, float* window, float* mel) {
    if (!h) return PSLA_E_INVALID;
    if (window) std::memcpy(window, h->window.data(), sizeof(float) * kFrameLen);
    if (mel) std::memcpy(mel, h->mel_dense.data(), sizeof(float) * h->mel_dense.size());
    return PSLA_OK;
}

size_t psla_frontend_workspace_bytes(const psla_frontend_t* h, int n_clips) {
    if (!h || n_clips < 0) return 0;
    return align_up(static_cast<size_t>(n_clips) * sizeof(ClipDesc), 256) +
           static_cast<size_t>(n_clips) * kStatSegs * kStatVals * sizeof(double);
}

int psla_frontend_num_frames(const psla_frontend_t* h, int64_t n_samples) {
    (void)h;
    const int64_t m = frames_of(n_samples);
    return m > 0x7fffffff ? 0x7fffffff : static_cast<int>(m);
}

int64_t psla_frontend_launch_count(const psla_frontend_t* h) { return h ? h->launches : 0; }

int psla_frontend_set_fast_log(psla_frontend_t* h, int enable) {
    if (!h) return PSLA_E_INVALID;
    h->fast_log = enable ? 1 : 0;
    return PSLA_OK;
}

int psla_frontend_forward(psla_frontend_t* h, const float* wav, const int64_t* wav_offset, const int64_t* wav_length,
                          int n_wav, const int32_t* src1,
                          const int32_t* src2, const double* lambda, const int32_t* mask, int n_clips, int target_length,
                          float norm_mean, float norm_div, int skip_norm, float* out, void* workspace,
                          size_t workspace_bytes, void* stream_) {
    if (!h || !wav_offset || !wav_length || !src1 || n_wav < 0 || n_clips < 0 || target_length < 1) return PSLA_E_INVALID;
    if (n_clips == 0) return PSLA_OK;
    if (!wav || !out) return PSLA_E_INVALID;
    if ((reinterpret_cast<uintptr_t>(wav) & 15) || (reinterpret_cast<uintptr_t>(out) & 3)) return PSLA_E_INVALID;
    if (n_clips > 65535) return PSLA_E_INVALID;  // gridDim.y
    const size_t need = psla_frontend_workspace_bytes(h, n_clips);
    if (!workspace || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255)) return PSLA_E_WORKSPACE;
    if (!skip_norm && !(norm_div != 0.f)) return PSLA_E_INVALID;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);

    DeviceGuard guard(h->device);
    if (!guard.ok) return PSLA_E_CUDA;

    // ---- pinned staging slot ----
    psla_frontend::Slot& slot = h->slots[h->next_slot];
    const size_t desc_bytes = static_cast<size_t>(n_clips) * sizeof(ClipDesc);
    if (slot.pending) {
        PSLA_CUDA(h, cudaEventSynchronize(slot.ev));
        slot.pending = false;
    }
    if (slot.cap < desc_bytes) {
        if (slot.host) cudaFreeHost(slot.host);
        slot.host = nullptr;
        slot.cap = 0;
        const size_t cap = align_up(desc_bytes * 2, 4096);
        PSLA_CUDA(h, cudaHostAlloc(&slot.host, cap, cudaHostAllocDefault));
        slot.cap = cap;
    }
    ClipDesc* descs = static_cast<ClipDesc*>(slot.host);

    // ---- validate + build descriptors (errors surface synchronously, like the reference's exceptions) ----
    for (int i = 0; i < n_wav; ++i)
        if (wav_offset[i] < 0 || wav_length[i] < 0) return PSLA_E_INVALID;
    for (int b = 0; b < n_clips; ++b) {
        ClipDesc& d = descs[b];
        const int a = src1[b];
        if (a < 0 || a >= n_wav) return PSLA_E_INVALID;
        const int64_t n1 = wav_length[a];
        if (n1 < kFrameLen) return PSLA_E_SHORT_CLIP;  // kaldi.py:142
        if (n1 > 0x7fffffff - 8) return PSLA_E_INVALID;
        d.off1 = wav_offset[a];
        d.n1 = static_cast<int>(n1);
        d.off2 = 0;
        d.n2 = 0;
        d.l1 = 1.f;
        d.l2 = 0.f;
        const int c = src2 ? src2[b] : -1;
        if (c >= 0) {
            if (c >= n_wav || !lambda) return PSLA_E_INVALID;
            const int64_t n2 = wav_length[c];
            if (n2 < 1 || n2 > 0x7fffffff - 8) return PSLA_E_INVALID;
            d.off2 = wav_offset[c];
            d.n2 = static_cast<int>(n2);
            d.l1 = static_cast<float>(lambda[b]);        // audioset_dataset.py:116: python scalar -> fp32
            d.l2 = static_cast<float>(1.0 - lambda[b]);  // (1 - mix_lambda) evaluated in float64 first
        } else if (c != -1) {
            return PSLA_E_INVALID;
        }
        d.aligned = ((d.off1 & 3) == 0 ? 1 : 0) | ((d.off2 & 3) == 0 ? 2 : 0);
        d.f0 = d.f1 = d.t0 = d.t1 = 0;
        if (mask) {
            const int32_t* m = mask + 4 * b;
            if (m[0] < 0 || m[1] < 0 || m[2] < 0 || m[3] < 0) return PSLA_E_MASK;
            if (m[1] > 0 && static_cast<int64_t>(m[0]) + m[1] > h->nmel) return PSLA_E_MASK;
            if (m[3] > 0 && static_cast<int64_t>(m[2]) + m[3] > target_length) return PSLA_E_MASK;
            if (m[1] > 0) {
                d.f0 = m[0];
                d.f1 = m[0] + m[1];
            }
            if (m[3] > 0) {
                d.t0 = m[2];
                d.t1 = m[2] + m[3];
            }
        }
        const int64_t mfr = frames_of(n1);
        d.m_eff = static_cast<int>(mfr < target_length ? mfr : target_length);
    }

    // ---- enqueue ----
    ClipDesc* d_desc = static_cast<ClipDesc*>(workspace);
    double* d_part = reinterpret_cast<double*>(static_cast<char*>(workspace) + align_up(desc_bytes, 256));
    PSLA_CUDA(h, cudaMemcpyAsync(d_desc, descs, desc_bytes, cudaMemcpyHostToDevice, stream));
    PSLA_CUDA(h, cudaEventRecord(slot.ev, stream));
    slot.pending = true;
    h->next_slot = (h->next_slot + 1) % 4;

    clip_stats_kernel<<<dim3(kStatSegs, n_clips), 256, 0, stream>>>(wav, d_desc, d_part);
    PSLA_CUDA(h, cudaGetLastError());

    FbankParams p;
    p.wav = wav;
    p.desc = d_desc;
    p.partials = d_part;
    p.tab = h->d_tab;
    p.out = out;
    p.tab_floats = h->tab_floats;
    p.T = target_length;
    p.nmel = h->nmel;
    p.norm_mean = norm_mean;
    p.norm_div = norm_div;
    p.skip_norm = skip_norm ? 1 : 0;
    p.fast_log = h->fast_log;
    p.preemph = h->preemph;
    p.c1 = static_cast<float>(1.0 - static_cast<double>(h->preemph));
    const int tiles = (target_length + kFramesPerCta - 1) / kFramesPerCta;
    fbank_tile_kernel<<<dim3(tiles, n_clips), kThreads, h->smem_bytes, stream>>>(p);
    PSLA_CUDA(h, cudaGetLastError());
    h->launches += 2;
    return PSLA_OK;
}

float* psla_frontend_device_out(const psla_frontend_t* h) { return h ? h->d_out : nullptr; }

int psla_frontend_forward_host(psla_frontend_t* h, const float* wav_host, const int64_t* wav_offset,
                               const int64_t* wav_length, int n_wav,
                               const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask,
                               int n_clips, int target_length, float norm_mean, float norm_div, int skip_norm,
                               float* out_host, void* stream_) {
    if (!h || !wav_host || !wav_offset || !wav_length || n_wav < 1 || n_clips < 0 || target_length < 1) return PSLA_E_INVALID;
    if (n_clips == 0) return PSLA_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(h->device);
    if (!guard.ok) return PSLA_E_CUDA;
    int64_t total = 0;
    for (int i = 0; i < n_wav; ++i) {
        if (wav_offset[i] < 0 || wav_length[i] < 0) return PSLA_E_INVALID;
        if (wav_offset[i] + wav_length[i] > total) total = wav_offset[i] + wav_length[i];
    }
    if (total < 1) return PSLA_E_INVALID;
    const size_t wav_bytes = static_cast<size_t>(total) * sizeof(float);
    const size_t out_bytes = static_cast<size_t>(n_clips) * target_length * h->nmel * sizeof(float);
    const size_t ws_bytes = psla_frontend_workspace_bytes(h, n_clips);
    auto grow = [&](void** ptr, size_t* cap, size_t need) -> int {
        if (*cap >= need) return PSLA_OK;
        PSLA_CUDA(h, cudaDeviceSynchronize());  // in-flight work may still use the old buffer
        if (*ptr) cudaFree(*ptr);
        *ptr = nullptr;
        *cap = 0;
        const size_t cap_new = align_up(need + need / 4, 1 << 20);
        if (cudaMalloc(ptr, cap_new) != cudaSuccess) {
            cudaGetLastError();
            return PSLA_E_NOMEM;
        }
        *cap = cap_new;
        return PSLA_OK;
    };
    int st;
    if ((st = grow(reinterpret_cast<void**>(&h->d_wav), &h->d_wav_cap, wav_bytes)) != PSLA_OK) return st;
    if ((st = grow(reinterpret_cast<void**>(&h->d_out), &h->d_out_cap, out_bytes)) != PSLA_OK) return st;
    if ((st = grow(&h->d_ws, &h->d_ws_cap, ws_bytes)) != PSLA_OK) return st;
    PSLA_CUDA(h, cudaMemcpyAsync(h->d_wav, wav_host, wav_bytes, cudaMemcpyHostToDevice, stream));
    st = psla_frontend_forward(h, h->d_wav, wav_offset, wav_length, n_wav, src1, src2, lambda, mask, n_clips, target_length, norm_mean,
                               norm_div, skip_norm, h->d_out, h->d_ws, h->d_ws_cap, stream_);
    if (st != PSLA_OK) return st;
    if (out_host) PSLA_CUDA(h, cudaMemcpyAsync(out_host, h->d_out, out_bytes, cudaMemcpyDeviceToHost, stream));
    return PSLA_OK;
}

}  // extern "C"
