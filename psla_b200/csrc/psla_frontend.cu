// libpsla_frontend.so -- fused Kaldi-compatible log-mel frontend for sm_100a (B200).
//
// Replaces, for a batch of clips on one GPU, what the reference computes per item on CPU dataloader
// workers: AudiosetDataset._wav2fbank + the SpecAugment / normalisation tail of __getitem__
// (src/dataloaders/audioset_dataset.py:88-137, 177-192), whose arithmetic is
// torchaudio.compliance.kaldi.fbank (torchaudio/compliance/kaldi.py:514-645).
//
// Kernels per forward():
//   fbank_stream_kernel  persistent, four CTAs of 4 autonomous warps per SM: every warp walks along its share of the
//                        batch's frames 4 at a time -- stage 640 new (mixed, pre-emphasised) samples in its private
//                        shared-memory window, two 512-point complex FFTs (one per half-warp, two real frames each) in
//                        two-wide fp32 arithmetic, power, sparse mel, log, SpecAugment fill, normalisation, 128-byte row
//                        stores (see fbank_stream.cuh for the per-lane algorithm, stream_kernel.cuh for the kernel).
//   clip_stats_kernel    per-clip waveform sums (deterministic two-level reduction) for the whole-clip means of
//                        audioset_dataset.py:92,98-99,117.  Launched only for clips that need them: all clips in
//                        exact-clip-mean mode, otherwise only mixup clips whose partner is shorter.
//   noise_roll_kernel    the optional noise / roll augmentation (audioset_dataset.py:197-199), a separate pass (noise_roll.cuh).
// No cuFFT, no tensor cores (the mel projection is 504 non-zeros, not a dense contraction), no CPU path.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/psla_frontend.h"
#include "noise_roll.cuh"
#include "stream_kernel.cuh"
#include "generic_kernel.cuh"
#include "tables.h"

// ================================================================================================
// Host side: handle, tables, validation, launches (C ABI)
// ================================================================================================
using namespace psla;

constexpr int kMetaSlots = 4;
constexpr int kPartialSlots = 3;  // partial-sum buffers of the statistics kernel, rotated: the statistics pass of call k + 1 may run while the streaming kernel of call k still reads its own
constexpr int kCtasPerSm = 4;  // __launch_bounds__(kThreads, 4): 4 warps x 128 registers and ~53 KB of shared memory per CTA
constexpr size_t kSmemPerSm = 228 * 1024, kSmemPerCtaReserved = 1024, kSmemPerCtaMax = 227 * 1024;

// ---- scratch of the allocation-free entry point (psla_frontend_forward_static; no counterpart in the reference) ----
// A file-scope device array: every GPU that loads this library holds PSLA_STATIC_SCRATCH_BYTES of it until the process
// exits.  Laid out like a forward() workspace for PSLA_STATIC_MAX_CLIPS clips (descriptor slots, unused on this path, then
// the kPartialSlots partial-sum regions of clip_stats_kernel).  In HOST code `g_static_scratch` is only a stand-in for
// the symbol, not a device address: static_scratch_address() resolves it once per device with cudaGetSymbolAddress and
// only that pointer is ever passed to a kernel.
constexpr size_t kStaticScratchBytes = PSLA_STATIC_SCRATCH_BYTES;
constexpr size_t static_align(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr size_t kStaticScratchNeed =
    kMetaSlots * (static_align(size_t(PSLA_STATIC_MAX_CLIPS) * sizeof(ClipDesc), 256) + static_align((size_t(PSLA_STATIC_MAX_CLIPS) + 1) * sizeof(ClipCost), 256)) +
    kPartialSlots * static_align(size_t(PSLA_STATIC_MAX_CLIPS) * kStatSegs * kStatVals * sizeof(double), 256);
static_assert(kStaticScratchNeed <= kStaticScratchBytes, "PSLA_STATIC_SCRATCH_BYTES does not hold PSLA_STATIC_MAX_CLIPS clips");
static_assert(PSLA_STATIC_MAX_CLIPS <= kParamClips, "the allocation-free path passes descriptors as kernel parameters");
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

struct psla_frontend {
    int device = 0;
    int sample_rate = 16000;
    int nmel = 128;
    float preemph = 0.97f;
    float low_freq = 20.f, high_freq = 0.f;
    // frame geometry, kaldi.py:138-140: int(sr * 25 ms), int(sr * 10 ms), next power of two.  400 / 160 / 512 (16 kHz) runs
    // the streaming kernel; every other geometry (or generic_forced) the geometry-generic kernel of generic_kernel.cuh.
    int frame_len = kFrameLen, frame_shift = kFrameShift, nfft = kNfft;
    bool native_geometry = true;  // 400 / 160 / 512
    bool generic_forced = false;  // psla_frontend_set_generic_path
    bool generic() const { return !native_geometry || generic_forced; }
    float* d_gen = nullptr;       // tables of the generic kernel: window | tw | twn | mel_range | mel_w
    size_t gen_off_tw = 0, gen_off_twn = 0, gen_off_range = 0, gen_off_w = 0;
    std::vector<float> window;     // [frame_len] analysis window (not halved)
    std::vector<float> mel_dense;  // [nmel][nfft / 2]
    float* d_tab = nullptr;
    int tab_floats = 0;
    size_t smem_bytes = 0;
    int ctas_per_sm = kCtasPerSm;  // resident CTAs per SM the shared-memory footprint allows (4 with the 128-bin tables)
    int fast_log = 0;
    float lg2_eps = -23.0f;  // what the device's lg2.approx returns for the float epsilon (probed once in create())
    int exact_mean = 0;
    int layout = PSLA_OUT_TF_F32;
    bool mel_spec = false;
    int num_sms = 0;
    int timing = 0;
    long long* trace = nullptr;  // device buffer of the optional per-warp timeline
    CostModel cm = kDefaultCost;  // work-partition weights (psla_frontend_set_cost_model)
    struct EvTriple {
        cudaEvent_t e[3];
    };
    std::vector<EvTriple> timed;
    int64_t launches = 0;
    int last_cuda = 0;
    // Descriptor staging ring: call k fills pinned slot k % kMetaSlots, which is copied to device slot k % kMetaSlots of
    // the caller's workspace on `copy_stream` (so the copy is off the launch stream: kernels of consecutive calls run
    // back to back); `copied` orders the kernel after the copy, `done` (recorded after the kernel) guards slot reuse.
    struct Slot {
        void* host = nullptr;
        size_t cap = 0;
        cudaEvent_t copied = nullptr, done = nullptr;
        bool pending = false;
    } slots[kMetaSlots];
    int next_slot = 0;
    int next_partial = 0;
    cudaStream_t copy_stream = nullptr;
    int meta_on_copy_stream = 1;
    // host-side descriptor build buffers; batches of <= kParamClips clips can be passed as kernel parameters
    MetaBlock meta_host;
    std::vector<ClipDesc> big_descs;
    std::vector<ClipCost> big_costs;
    int param_meta = 1;
    // programmatic overlap of consecutive launches (psla_frontend_set_launch_overlap): 0 never, 1 between the launches of
    // ONE forward() call (default: the library itself enqueued the predecessor), 2 also across forward() calls (caller contract)
    int pdl = PSLA_OVERLAP_CALL;
    bool prev_valid = false;
    void* prev_stream = nullptr;
    // output ranges of the last kPrevOuts launches.  Every launch joins its predecessor before it exits (pdl_join in
    // stream_kernel.cuh) and the grid fills the resident-CTA capacity, so only ADJACENT launches can be in flight
    // together; three ranges are kept as margin for devices / carve-outs where more thread blocks are resident.
    static constexpr int kPrevOuts = 3;
    const char* prev_out_lo[kPrevOuts] = {nullptr, nullptr, nullptr};
    const char* prev_out_hi[kPrevOuts] = {nullptr, nullptr, nullptr};
    // forward_host device buffers
    float* d_wav = nullptr;
    size_t d_wav_cap = 0;
    float* d_out = nullptr;
    size_t d_out_cap = 0;
    void* d_ws = nullptr;
    size_t d_ws_cap = 0;
    // device address of g_static_scratch on `device` (resolved in create(); forward_static only)
    void* static_scratch = nullptr;
    int static_scratch_err = 0;
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

// Device address of g_static_scratch on the CURRENT device; resolved once per device (the symbol has one instance per
// device, and cudaGetSymbolAddress answers for the device that is current).  Returns the CUDA error code of the query.
int static_scratch_address(int device, void** ptr) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static void* cached[kMaxDevices] = {};
    *ptr = nullptr;
    std::lock_guard<std::mutex> lock(mu);
    if (device >= 0 && device < kMaxDevices && cached[device]) {
        *ptr = cached[device];
        return 0;
    }
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return static_cast<int>(e);
    }
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255)) return static_cast<int>(cudaErrorInvalidSymbol);
    if (device >= 0 && device < kMaxDevices) cached[device] = p;
    *ptr = p;
    return 0;
}

__global__ void probe_lg2_kernel(float x, float* out) { *out = lg2_ftz(x); }
int probe_lg2_eps(psla_frontend* h) {
    float* d = nullptr;
    PSLA_CUDA(h, cudaMalloc(&d, sizeof(float)));
    probe_lg2_kernel<<<1, 1>>>(1.1920928955078125e-07f, d);
    cudaError_t e = cudaMemcpy(&h->lg2_eps, d, sizeof(float), cudaMemcpyDeviceToHost);
    cudaFree(d);
    PSLA_CUDA(h, e);
    return PSLA_OK;
}

template <bool EXACT, bool MELSPEC, bool PMETA, class S, bool TRACE, int OUT = kOutTF>
cudaError_t prepare_one(size_t smem) {
    cudaError_t e = cudaFuncSetAttribute(fbank_stream_kernel<EXACT, MELSPEC, PMETA, S, TRACE, OUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(fbank_stream_kernel<EXACT, MELSPEC, PMETA, S, TRACE, OUT>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
}
#ifdef PSLA_SLIM
// Experimental builds (psla_b200/build.py build_variant): only the kernels the bench batch uses are instantiated --
// default clip-mean mode, specialised 128-bin mel path, fp32 samples, [B, T, F] output (both descriptor paths, with and
// without the timeline).  Everything else reports cudaErrorNotSupported.
template <class S>
cudaError_t launch_stream_kernel(const cudaLaunchConfig_t& cfg, bool exact, bool spec, bool pmeta, bool trace, int layout,
                                 const StreamParams& p, const MetaBlock& mb) {
    const MetaNone none = {0};
    if constexpr (!std::is_same_v<S, float>) return cudaErrorNotSupported;
    else {
        if (exact || !spec || layout != PSLA_OUT_TF_F32) return cudaErrorNotSupported;
        if (trace) {
            if (pmeta) return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, true, float, true>, p, mb);
            return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, false, float, true>, p, none);
        }
        if (pmeta) return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, true, float>, p, mb);
        return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, false, float>, p, none);
    }
}
cudaError_t prepare_all(size_t smem) {
    cudaError_t e = prepare_one<false, true, false, float, false>(smem);
    if (e == cudaSuccess) e = prepare_one<false, true, true, float, false>(smem);
    if (e == cudaSuccess) e = prepare_one<false, true, false, float, true>(smem);
    if (e == cudaSuccess) e = prepare_one<false, true, true, float, true>(smem);
    return e;
}
#else
template <bool EXACT, bool MELSPEC, bool TRACE>
cudaError_t prepare_kernel(size_t smem) {
    cudaError_t e = prepare_one<EXACT, MELSPEC, false, float, TRACE>(smem);
    if (e == cudaSuccess) e = prepare_one<EXACT, MELSPEC, true, float, TRACE>(smem);
    if (e == cudaSuccess) e = prepare_one<EXACT, MELSPEC, false, short, TRACE>(smem);
    if (e == cudaSuccess) e = prepare_one<EXACT, MELSPEC, true, short, TRACE>(smem);
    return e;
}

template <int OUT>
cudaError_t prepare_layout(size_t smem) {
    cudaError_t e = prepare_one<false, true, false, float, false, OUT>(smem);
    if (e == cudaSuccess) e = prepare_one<false, true, true, float, false, OUT>(smem);
    if (e == cudaSuccess) e = prepare_one<false, true, false, short, false, OUT>(smem);
    if (e == cudaSuccess) e = prepare_one<false, true, true, short, false, OUT>(smem);
    return e;
}

// launch of the variant selected at run time (the tracing variants exist for the default clip-mean mode only, the
// frequency-major layouts for the default mode without tracing and 128 mel bins)
template <class S>
cudaError_t launch_stream_kernel(const cudaLaunchConfig_t& cfg, bool exact, bool spec, bool pmeta, bool trace, int layout,
                                 const StreamParams& p, const MetaBlock& mb) {
    const MetaNone none = {0};
    if (layout == PSLA_OUT_FT_F32) {
        if (pmeta) return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, true, S, false, kOutFT>, p, mb);
        return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, false, S, false, kOutFT>, p, none);
    }
    if (layout == PSLA_OUT_FT_BF16) {
        if (pmeta) return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, true, S, false, kOutFTbf16>, p, mb);
        return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, false, S, false, kOutFTbf16>, p, none);
    }
    if (trace) {
        switch ((spec ? 2 : 0) | (pmeta ? 1 : 0)) {
            case 0: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, false, false, S, true>, p, none);
            case 1: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, false, true, S, true>, p, mb);
            case 2: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, false, S, true>, p, none);
            default: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, true, S, true>, p, mb);
        }
    }
    switch ((exact ? 4 : 0) | (spec ? 2 : 0) | (pmeta ? 1 : 0)) {
        case 0: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, false, false, S>, p, none);
        case 1: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, false, true, S>, p, mb);
        case 2: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, false, S>, p, none);
        case 3: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<false, true, true, S>, p, mb);
        case 4: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<true, false, false, S>, p, none);
        case 5: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<true, false, true, S>, p, mb);
        case 6: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<true, true, false, S>, p, none);
        default: return cudaLaunchKernelEx(&cfg, fbank_stream_kernel<true, true, true, S>, p, mb);
    }
}

cudaError_t prepare_all(size_t smem) {
    cudaError_t e = prepare_kernel<false, false, false>(smem);
    if (e == cudaSuccess) e = prepare_kernel<false, true, false>(smem);
    if (e == cudaSuccess) e = prepare_kernel<true, false, false>(smem);
    if (e == cudaSuccess) e = prepare_kernel<true, true, false>(smem);
    if (e == cudaSuccess) e = prepare_kernel<false, false, true>(smem);
    if (e == cudaSuccess) e = prepare_kernel<false, true, true>(smem);
    if (e == cudaSuccess) e = prepare_layout<kOutFT>(smem);
    if (e == cudaSuccess) e = prepare_layout<kOutFTbf16>(smem);
    return e;
}
#endif

// Tables of the geometry-generic kernel (generic_kernel.cuh), built for every handle (a few hundred KB at most): the 16 kHz
// geometry can be switched onto that kernel as well (psla_frontend_set_generic_path).
int upload_generic_tables(psla_frontend* h) {
    const int L = h->frame_len, N = h->nfft, M = N / 2, nmel = h->nmel;
    const size_t off_tw = align_up(static_cast<size_t>(L), 4);
    const size_t off_twn = off_tw + 2 * static_cast<size_t>(M);         // M complex
    const size_t off_range = off_twn + align_up(static_cast<size_t>(M) + 2, 4);  // M/2 + 1 complex
    const size_t off_w = off_range + align_up(2 * static_cast<size_t>(nmel), 4);
    const size_t total = off_w + static_cast<size_t>(nmel) * M;
    std::vector<float> buf(total, 0.f);
    std::memcpy(buf.data(), h->window.data(), sizeof(float) * L);
    const double two_pi = 6.283185307179586476925286766559;
    for (int j = 0; j < M; ++j) {
        buf[off_tw + 2 * j] = static_cast<float>(std::cos(two_pi * j / M));
        buf[off_tw + 2 * j + 1] = static_cast<float>(-std::sin(two_pi * j / M));
    }
    for (int k = 0; k <= M / 2; ++k) {
        buf[off_twn + 2 * k] = static_cast<float>(std::cos(two_pi * k / N));
        buf[off_twn + 2 * k + 1] = static_cast<float>(-std::sin(two_pi * k / N));
    }
    int* range = reinterpret_cast<int*>(buf.data() + off_range);
    for (int m = 0; m < nmel; ++m) {
        const float* row = h->mel_dense.data() + static_cast<size_t>(m) * M;
        int lo = 0, hi = 0;
        for (int k = 0; k < M; ++k)
            if (row[k] != 0.f) {
                if (hi == 0) lo = k;
                hi = k + 1;
            }
        range[2 * m] = lo;
        range[2 * m + 1] = hi;
    }
    std::memcpy(buf.data() + off_w, h->mel_dense.data(), sizeof(float) * static_cast<size_t>(nmel) * M);
    if (h->d_gen) {
        PSLA_CUDA(h, cudaDeviceSynchronize());
        cudaFree(h->d_gen);
        h->d_gen = nullptr;
    }
    PSLA_CUDA(h, cudaMalloc(&h->d_gen, total * sizeof(float)));
    PSLA_CUDA(h, cudaMemcpy(h->d_gen, buf.data(), total * sizeof(float), cudaMemcpyHostToDevice));
    h->gen_off_tw = off_tw;
    h->gen_off_twn = off_twn;
    h->gen_off_range = off_range;
    h->gen_off_w = off_w;
    const int smem = static_cast<int>(generic_smem_bytes(N));
    PSLA_CUDA(h, cudaFuncSetAttribute(fbank_generic_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    PSLA_CUDA(h, cudaFuncSetAttribute(fbank_generic_kernel<short>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    PSLA_CUDA(h, cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, h->device));
    return PSLA_OK;
}

int upload_tables(psla_frontend* h) {
    const int gst = upload_generic_tables(h);
    if (gst != PSLA_OK || !h->native_geometry) return gst;
    std::vector<float> buf;
    bool spec = false;
    if (!build_table_buffer(h->window.data(), h->mel_dense.data(), h->nmel, &buf, &spec)) return PSLA_E_UNSUPPORTED;
    const int total = static_cast<int>(buf.size());
    const size_t smem = (static_cast<size_t>(total) + static_cast<size_t>(kWarpsPerCta) * kWarpFloats) * sizeof(float);
    // four CTAs (tables + 4 warp windows each) fit one SM with the 128-bin tables; the larger tables of small bin counts
    // (wide triangles: more taps per lane) run with fewer resident CTAs -- the grid is sized to what is resident
    if (smem > kSmemPerCtaMax) return PSLA_E_UNSUPPORTED;
    const int fit = static_cast<int>(kSmemPerSm / (smem + kSmemPerCtaReserved));
    h->ctas_per_sm = fit < kCtasPerSm ? fit : kCtasPerSm;
    if (h->d_tab) {
        PSLA_CUDA(h, cudaDeviceSynchronize());
        cudaFree(h->d_tab);
        h->d_tab = nullptr;
    }
    PSLA_CUDA(h, cudaMalloc(&h->d_tab, total * sizeof(float)));
    PSLA_CUDA(h, cudaMemcpy(h->d_tab, buf.data(), total * sizeof(float), cudaMemcpyHostToDevice));
    h->tab_floats = total;
    h->smem_bytes = smem;
    h->mel_spec = spec;
    PSLA_CUDA(h, prepare_all(smem));
    PSLA_CUDA(h, cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, h->device));
    return PSLA_OK;
}

int64_t frames_of(const psla_frontend* h, int64_t n) {  // _get_strided, kaldi.py:63-67 (snip_edges)
    const int len = h ? h->frame_len : kFrameLen, shift = h ? h->frame_shift : kFrameShift;
    return n < len ? -1 : 1 + (n - len) / shift;
}

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
        case PSLA_E_STATIC_LIMIT:
            return "forward_static: more than PSLA_STATIC_MAX_CLIPS (128) clips in one call, or a handle mode that allocates per launch "
                   "(set_param_meta(0) / set_timing(1)); split the batch or use psla_frontend_forward with a workspace";
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
    int nfft = 1;
    while (nfft < size) nfft <<= 1;  // round_to_power_of_two, kaldi.py:140
    if (shift < 1 || size < 2 || nfft < kGenMinFft || nfft > kGenMaxFft) return PSLA_E_UNSUPPORTED;
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
    h->frame_len = size;
    h->frame_shift = shift;
    h->nfft = nfft;
    h->native_geometry = (size == kFrameLen && shift == kFrameShift && nfft == kNfft);
    h->window.resize(size);
    std::vector<float> half(size);
    build_window_half(half.data(), size);
    for (int j = 0; j < size; ++j) h->window[j] = 2.f * half[j];
    h->mel_dense.resize(static_cast<size_t>(num_mel_bins) * (nfft / 2));
    build_mel_dense(h->mel_dense.data(), num_mel_bins, nfft, static_cast<float>(sample_rate), low_freq, high_freq);
    DeviceGuard guard(device);
    if (!guard.ok) {
        delete h;
        return PSLA_E_CUDA;
    }
    int st = upload_tables(h);
    if (st == PSLA_OK) st = probe_lg2_eps(h);
    h->static_scratch_err = static_scratch_address(device, &h->static_scratch);  // reported by forward_static, nobody else needs it
    if (st == PSLA_OK) {
        for (auto& s : h->slots)
            if (cudaEventCreateWithFlags(&s.copied, cudaEventDisableTiming) != cudaSuccess ||
                cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming) != cudaSuccess)
                st = PSLA_E_CUDA;
        if (cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) != cudaSuccess) st = PSLA_E_CUDA;
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
    for (auto& t : h->timed)
        for (auto& e : t.e) cudaEventDestroy(e);
    for (auto& s : h->slots) {
        if (s.host) cudaFreeHost(s.host);
        if (s.copied) cudaEventDestroy(s.copied);
        if (s.done) cudaEventDestroy(s.done);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->d_tab) cudaFree(h->d_tab);
    if (h->d_gen) cudaFree(h->d_gen);
    if (h->d_wav) cudaFree(h->d_wav);
    if (h->d_out) cudaFree(h->d_out);
    if (h->d_ws) cudaFree(h->d_ws);
    delete h;
}

int psla_frontend_set_tables(psla_frontend_t* h, const float* window, const float* mel) {
    if (!h) return PSLA_E_INVALID;
    std::vector<float> w0 = h->window, m0 = h->mel_dense;
    if (window) std::memcpy(h->window.data(), window, sizeof(float) * h->window.size());
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
    if (window) std::memcpy(window, h->window.data(), sizeof(float) * h->window.size());
    if (mel) std::memcpy(mel, h->mel_dense.data(), sizeof(float) * h->mel_dense.size());
    return PSLA_OK;
}

size_t psla_frontend_workspace_bytes(const psla_frontend_t* h, int n_clips) {
    if (!h || n_clips < 0) return 0;
    const size_t meta = align_up(static_cast<size_t>(n_clips) * sizeof(ClipDesc), 256) +
                        align_up((static_cast<size_t>(n_clips) + 1) * sizeof(ClipCost), 256);
    return kMetaSlots * meta + kPartialSlots * align_up(static_cast<size_t>(n_clips) * kStatSegs * kStatVals * sizeof(double), 256);
}

int psla_frontend_noise_roll(psla_frontend_t* h, const float* in, float* out, const float* noise, const float* scale,
                             const int32_t* shift, int n_clips, int target_length, uint64_t seed, void* stream_) {
    if (!h || !in || !out || !scale || !shift || n_clips < 0 || target_length < 1 || in == out) return PSLA_E_INVALID;
    if (n_clips == 0) return PSLA_OK;
    if (h->layout != PSLA_OUT_TF_F32) return PSLA_E_UNSUPPORTED;  // rows are rolled: [B, T, F] only
    const size_t bytes = static_cast<size_t>(n_clips) * target_length * h->nmel * sizeof(float);
    const char *a = reinterpret_cast<const char*>(in), *b = reinterpret_cast<const char*>(out);
    if (a < b + bytes && b < a + bytes) return PSLA_E_INVALID;  // a roll cannot run in place
    for (int i = 0; i < n_clips; ++i)
        if (!(scale[i] == scale[i]) || shift[i] <= -target_length || shift[i] >= target_length) return PSLA_E_INVALID;
    DeviceGuard guard(h->device);
    if (!guard.ok) return PSLA_E_CUDA;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool vec = h->nmel == 128 && (reinterpret_cast<size_t>(in) % 16 == 0) && (reinterpret_cast<size_t>(out) % 16 == 0) &&
                     (!noise || reinterpret_cast<size_t>(noise) % 16 == 0);
    for (int c0 = 0; c0 < n_clips; c0 += kNoiseClips) {
        const int n = n_clips - c0 < kNoiseClips ? n_clips - c0 : kNoiseClips;
        NoiseMeta meta;
        for (int i = 0; i < n; ++i) {
            meta.scale[i] = scale[c0 + i];
            const int s = shift[c0 + i];
            meta.shift[i] = s < 0 ? s + target_length : s;  // torch.roll: element t moves to (t + shift) mod T
        }
        const size_t off = static_cast<size_t>(c0) * target_length * h->nmel;
        const long long rows = static_cast<long long>(n) * target_length;
        const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
        if (vec) noise_roll_kernel<true><<<grid, 256, 0, stream>>>(in + off, out + off, noise ? noise + off : nullptr, meta, n, target_length, h->nmel, seed, c0);
        else noise_roll_kernel<false><<<grid, 256, 0, stream>>>(in + off, out + off, noise ? noise + off : nullptr, meta, n, target_length, h->nmel, seed, c0);
        PSLA_CUDA(h, cudaGetLastError());
        h->launches += 1;
    }
    return PSLA_OK;
}

int psla_frontend_num_frames(const psla_frontend_t* h, int64_t n_samples) {
    const int64_t m = frames_of(h, n_samples);
    return m > 0x7fffffff ? 0x7fffffff : static_cast<int>(m);
}

int64_t psla_frontend_launch_count(const psla_frontend_t* h) { return h ? h->launches : 0; }

int psla_frontend_set_timing(psla_frontend_t* h, int enable) {
    if (!h) return PSLA_E_INVALID;
    h->timing = enable ? 1 : 0;
    return PSLA_OK;
}

int psla_frontend_get_timing(psla_frontend_t* h, double* stats_ms, double* tile_ms, int64_t* n_forward) {
    if (!h) return PSLA_E_INVALID;
    DeviceGuard guard(h->device);
    double a = 0.0, b = 0.0;
    for (auto& t : h->timed) {
        float m0 = 0.f, m1 = 0.f;
        PSLA_CUDA(h, cudaEventSynchronize(t.e[2]));
        PSLA_CUDA(h, cudaEventElapsedTime(&m0, t.e[0], t.e[1]));
        PSLA_CUDA(h, cudaEventElapsedTime(&m1, t.e[1], t.e[2]));
        a += m0;
        b += m1;
        for (auto& e : t.e) cudaEventDestroy(e);
    }
    if (stats_ms) *stats_ms = a;
    if (tile_ms) *tile_ms = b;
    if (n_forward) *n_forward = static_cast<int64_t>(h->timed.size());
    h->timed.clear();
    return PSLA_OK;
}

int psla_frontend_set_trace(psla_frontend_t* h, int64_t* device_buf, size_t capacity_records) {
    if (!h) return PSLA_E_INVALID;
    const size_t need = static_cast<size_t>(psla_frontend_trace_records(h));
    if (device_buf && capacity_records < need) return PSLA_E_WORKSPACE;
    static_assert(sizeof(long long) == sizeof(int64_t), "trace word");
    h->trace = reinterpret_cast<long long*>(device_buf);
    return PSLA_OK;
}

int psla_frontend_trace_records(const psla_frontend_t* h) {
    return h ? (h->num_sms > 0 ? h->num_sms : 148) * h->ctas_per_sm * kWarpsPerCta : 0;
}

int psla_frontend_trace_words(void) { return kTraceWords; }

int psla_frontend_set_meta_stream(psla_frontend_t* h, int enable) {
    if (!h) return PSLA_E_INVALID;
    h->meta_on_copy_stream = enable ? 1 : 0;
    return PSLA_OK;
}

int psla_frontend_set_launch_overlap(psla_frontend_t* h, int mode) {
    if (!h || mode < PSLA_OVERLAP_OFF || mode > PSLA_OVERLAP_STREAM) return PSLA_E_INVALID;
    h->pdl = mode;
    return PSLA_OK;
}

int psla_frontend_set_generic_path(psla_frontend_t* h, int enable) {
    if (!h) return PSLA_E_INVALID;
    h->generic_forced = enable != 0;
    return PSLA_OK;
}

int psla_frontend_get_geometry(const psla_frontend_t* h, int* frame_length, int* frame_shift, int* fft_size, int* generic_path) {
    if (!h) return PSLA_E_INVALID;
    if (frame_length) *frame_length = h->frame_len;
    if (frame_shift) *frame_shift = h->frame_shift;
    if (fft_size) *fft_size = h->nfft;
    if (generic_path) *generic_path = h->generic() ? 1 : 0;
    return PSLA_OK;
}

int psla_frontend_set_param_meta(psla_frontend_t* h, int enable) {
    if (!h) return PSLA_E_INVALID;
    h->param_meta = enable ? 1 : 0;
    return PSLA_OK;
}

int psla_frontend_set_cost_model(psla_frontend_t* h, int live, int dead, int cold) {
    // total cost of a batch must stay well inside int64 and a warp's budget inside int32
    if (!h || live < 1 || dead < 1 || cold < 0 || live > 4096 || dead > 4096 || cold > 4096) return PSLA_E_INVALID;
    h->cm.live = live;
    h->cm.dead = dead;
    h->cm.cold = cold;
    return PSLA_OK;
}

int psla_frontend_set_fast_log(psla_frontend_t* h, int enable) {
    if (!h) return PSLA_E_INVALID;
    h->fast_log = enable ? 1 : 0;
    return PSLA_OK;
}

int psla_frontend_set_exact_clip_mean(psla_frontend_t* h, int enable) {
    if (!h) return PSLA_E_INVALID;
    h->exact_mean = enable ? 1 : 0;
    return PSLA_OK;
}

int psla_frontend_set_output_layout(psla_frontend_t* h, int layout) {
    if (!h || layout < PSLA_OUT_TF_F32 || layout > PSLA_OUT_FT_BF16) return PSLA_E_INVALID;
    h->layout = layout;
    return PSLA_OK;
}

}  // extern "C"

namespace {
size_t out_elem_bytes(const psla_frontend* h) { return h->layout == PSLA_OUT_FT_BF16 ? 2 : 4; }

// Enqueue one launch (plus the statistics kernel where needed) for `n_clips` validated clips whose descriptors / costs are
// in `descs` / `costs` (h->meta_host's arrays when n_clips <= kParamClips).
int enqueue_batch(psla_frontend_t* h, const void* wav, bool pcm16, ClipDesc* descs, ClipCost* costs, int n_clips, long long total_cost,
                  bool any_stats, int target_length, float norm_mean, float norm_div, int skip_norm, float* out, void* workspace,
                  void* stream_, bool follows_own_launch) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const size_t desc_bytes = align_up(static_cast<size_t>(n_clips) * sizeof(ClipDesc), 256);
    const size_t cost_bytes = align_up((static_cast<size_t>(n_clips) + 1) * sizeof(ClipCost), 256);
    const size_t meta_bytes = desc_bytes + cost_bytes;
    const bool small = n_clips <= kParamClips;
    const int G = (target_length + kGroupFrames - 1) / kGroupFrames;
    // ---- enqueue ----
    // Small batches read their descriptors from the kernels' parameter space (the statistics kernel's too): nothing is
    // uploaded and, without a statistics pass, nothing sits between two launches on the stream.  Otherwise: pinned slot
    // -> device slot of the workspace on the copy stream, the kernels ordered after it by an event.
    const bool generic = h->generic();
    const bool param_meta = small && h->param_meta && !generic;  // the generic kernel reads its descriptors from device memory
    ClipDesc* d_desc = nullptr;
    ClipCost* d_cost = nullptr;
    const size_t part_bytes = align_up(static_cast<size_t>(n_clips) * kStatSegs * kStatVals * sizeof(double), 256);
    double* d_part = reinterpret_cast<double*>(static_cast<char*>(workspace) + kMetaSlots * meta_bytes + static_cast<size_t>(h->next_partial) * part_bytes);
    if (any_stats) h->next_partial = (h->next_partial + 1) % kPartialSlots;
    psla_frontend::Slot* slot = nullptr;
    if (!param_meta) {
        const int slot_index = h->next_slot;
        slot = &h->slots[slot_index];
        if (slot->pending) {
            PSLA_CUDA(h, cudaEventSynchronize(slot->done));  // the kernel that read this slot (pinned and device copy) is done
            slot->pending = false;
        }
        if (slot->cap < meta_bytes) {
            if (slot->host) cudaFreeHost(slot->host);
            slot->host = nullptr;
            slot->cap = 0;
            const size_t cap = align_up(meta_bytes * 2, 4096);
            PSLA_CUDA(h, cudaHostAlloc(&slot->host, cap, cudaHostAllocDefault));
            slot->cap = cap;
        }
        std::memcpy(slot->host, descs, static_cast<size_t>(n_clips) * sizeof(ClipDesc));
        std::memcpy(static_cast<char*>(slot->host) + desc_bytes, costs, (static_cast<size_t>(n_clips) + 1) * sizeof(ClipCost));
        char* d_meta = static_cast<char*>(workspace) + static_cast<size_t>(slot_index) * meta_bytes;
        d_desc = reinterpret_cast<ClipDesc*>(d_meta);
        d_cost = reinterpret_cast<ClipCost*>(d_meta + desc_bytes);
        if (h->meta_on_copy_stream) {
            PSLA_CUDA(h, cudaMemcpyAsync(d_meta, slot->host, meta_bytes, cudaMemcpyHostToDevice, h->copy_stream));
            PSLA_CUDA(h, cudaEventRecord(slot->copied, h->copy_stream));
            PSLA_CUDA(h, cudaStreamWaitEvent(stream, slot->copied, 0));
        } else {
            PSLA_CUDA(h, cudaMemcpyAsync(d_meta, slot->host, meta_bytes, cudaMemcpyHostToDevice, stream));
        }
        h->next_slot = (h->next_slot + 1) % kMetaSlots;
    }

    // ---- launch overlap (see psla_frontend_set_launch_overlap in the header) ----
    // Programmatic stream serialization lets a launch's thread blocks start on SM resources its PREDECESSOR ON THE STREAM
    // has already left -- if that predecessor signals early, which our kernels do and plain kernels / copies do not (behind
    // those the attribute changes nothing).  The streaming kernel does not wait before reading its inputs, so the attribute
    // is only attached when the predecessor is known to be one of our own launches:
    //   follows_own_launch     the previous chunk of this same forward() call (the library enqueued it a moment ago), or
    //   PSLA_OVERLAP_STREAM    the caller's promise that nothing that signals early itself sits between two calls.
    // Further conditions: descriptors in parameter space (no upload in between), output disjoint from those of the
    // launches that may still run, no tracing / timing.  A statistics pass is part of the chain: it may start behind the
    // previous call's streaming kernel (own partial-sum buffer), the streaming kernel of THIS call starts behind it and
    // waits for it only where a clip needs the sums (griddepcontrol.wait in enter_clip).
    const char* out_lo = reinterpret_cast<const char*>(out);
    const char* out_hi = out_lo + static_cast<size_t>(n_clips) * target_length * h->nmel * out_elem_bytes(h);
    bool disjoint = true;
    for (int i = 0; i < psla_frontend::kPrevOuts; ++i)
        disjoint = disjoint && (!h->prev_out_lo[i] || out_hi <= h->prev_out_lo[i] || out_lo >= h->prev_out_hi[i]);
    const bool own_pred = h->prev_valid && h->prev_stream == stream_ && (follows_own_launch || h->pdl == PSLA_OVERLAP_STREAM);
    const bool chain_ok = h->pdl != PSLA_OVERLAP_OFF && param_meta && !h->trace && !h->timing;
    const bool overlap_prev_call = chain_ok && own_pred && disjoint;  // the first kernel of this call behind the last one of the previous
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;

    psla_frontend::EvTriple ev{};
    if (h->timing) {
        for (auto& e : ev.e) PSLA_CUDA(h, cudaEventCreate(&e));
        PSLA_CUDA(h, cudaEventRecord(ev.e[0], stream));
    }
    if (any_stats) {
        cudaLaunchConfig_t sc = {};
        sc.gridDim = dim3(kStatSegs, n_clips);
        sc.blockDim = dim3(256);
        sc.dynamicSmemBytes = 0;
        sc.stream = stream;
        sc.attrs = attr;
        sc.numAttrs = overlap_prev_call ? 1 : 0;
        const MetaNone none = {0};
        const int stats_exact = (h->exact_mean || generic) ? 1 : 0;
        cudaError_t se;
        if (param_meta) {
            if (pcm16) se = cudaLaunchKernelEx(&sc, clip_stats_kernel<short, true>, static_cast<const short*>(wav), static_cast<const ClipDesc*>(nullptr), h->meta_host, d_part, stats_exact);
            else se = cudaLaunchKernelEx(&sc, clip_stats_kernel<float, true>, static_cast<const float*>(wav), static_cast<const ClipDesc*>(nullptr), h->meta_host, d_part, stats_exact);
        } else {
            if (pcm16) se = cudaLaunchKernelEx(&sc, clip_stats_kernel<short, false>, static_cast<const short*>(wav), static_cast<const ClipDesc*>(d_desc), none, d_part, stats_exact);
            else se = cudaLaunchKernelEx(&sc, clip_stats_kernel<float, false>, static_cast<const float*>(wav), static_cast<const ClipDesc*>(d_desc), none, d_part, stats_exact);
        }
        PSLA_CUDA(h, se);
        PSLA_CUDA(h, cudaGetLastError());
        h->launches += 1;
    }
    if (h->timing) PSLA_CUDA(h, cudaEventRecord(ev.e[1], stream));

    StreamParams p;
    p.wav = wav;
    p.desc = d_desc;
    p.cost = d_cost;
    p.partials = d_part;
    p.tab = h->d_tab;
    p.out = out;
    p.total_cost = total_cost;
    p.tab_floats = h->tab_floats;
    p.n_clips = n_clips;
    p.T = target_length;
    p.G = G;
    p.nmel = h->nmel;
    p.norm_mean = norm_mean;
    p.norm_div = norm_div;
    int fast_log = 0;
    // the value every masked / padded cell takes: (0 - mean) / std evaluated like torch does (fp32 sub, IEEE div);
    // likewise the value of a cell on the log floor, log(max(e, eps)) with e <= eps
    {
        const float log_eps = -15.9423847198486328f;  // fp32 log(1.1920929e-07)
        volatile float z = 0.0f - norm_mean;
        volatile float f = log_eps - norm_mean;
        p.fill = skip_norm ? 0.0f : z / norm_div;
        p.floor_val = skip_norm ? log_eps : f / norm_div;
        const double ln2 = 0.693147180559945309417;
        p.log_scale = skip_norm ? static_cast<float>(ln2) : static_cast<float>(ln2 / static_cast<double>(norm_div));
        p.log_bias = skip_norm ? 0.0f : static_cast<float>(-static_cast<double>(norm_mean) / static_cast<double>(norm_div));
        // The fast-log epilogue has no select for cells on the floor: nudge the bias (by a few ulps, ~1e-7) until
        // lg2(eps) * scale + bias (lg2(eps) as the device's lg2.approx returns it, probed in create()) is bit for bit the
        // reference's value.
        fast_log = h->fast_log;
        if (fast_log) {
            // candidates: the nominal bias moved in steps of half an ulp of the floor value (with skip_norm the nominal
            // bias is 0, whose own neighbours are denormals and would never get there)
            const float ulp = nextafterf(fabsf(p.floor_val), INFINITY) - fabsf(p.floor_val);
            const float nominal = p.log_bias;
            bool found = false;
            for (int k = 0; k <= 64 && !found; ++k) {
                for (int sgn = 1; sgn >= -1 && !found; sgn -= 2) {
                    const float cand = nominal + static_cast<float>(sgn * k) * 0.5f * ulp;
                    if (fmaf(h->lg2_eps, p.log_scale, cand) == p.floor_val) {
                        p.log_bias = cand;
                        found = true;
                    }
                }
            }
            if (!found) fast_log = 0;  // no such bias (never seen): the logf path evaluates the reference's expression verbatim
        }
    }
    p.skip_norm = skip_norm ? 1 : 0;
    p.fast_log = fast_log;
    p.preemph = h->preemph;
    p.c1 = static_cast<float>(1.0 - static_cast<double>(h->preemph));
    p.cm = h->cm;
    p.trace = h->trace;
    // behind our own statistics kernel the streaming kernel may always start early (it joins it where it needs the sums);
    // behind the previous call's streaming kernel under the conditions above
    const bool overlap_ok = any_stats ? chain_ok : overlap_prev_call;
    if (h->prev_stream != stream_)  // ranges written on another stream are ordered by the caller, not by us
        for (int i = 0; i < psla_frontend::kPrevOuts; ++i) h->prev_out_lo[i] = h->prev_out_hi[i] = nullptr;
    h->prev_valid = true;
    h->prev_stream = stream_;
    for (int i = psla_frontend::kPrevOuts - 1; i > 0; --i) {
        h->prev_out_lo[i] = h->prev_out_lo[i - 1];
        h->prev_out_hi[i] = h->prev_out_hi[i - 1];
    }
    h->prev_out_lo[0] = out_lo;
    h->prev_out_hi[0] = out_hi;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((h->num_sms > 0 ? h->num_sms : 148) * h->ctas_per_sm);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = h->smem_bytes;
    cfg.stream = stream;
    cfg.attrs = attr;
    cfg.numAttrs = overlap_ok ? 1 : 0;
    const bool trace = h->trace != nullptr;
    if (trace && (h->exact_mean || generic)) return PSLA_E_UNSUPPORTED;  // the timeline is compiled for the default clip-mean mode only
    cudaError_t le;
    if (generic) {
        // geometry-generic kernel (generic_kernel.cuh): a plain launch behind the statistics kernel, one CTA per 16 output rows
        GenericParams gp;
        gp.wav = wav;
        gp.desc = d_desc;
        gp.partials = d_part;
        gp.window = h->d_gen;
        gp.tw = reinterpret_cast<const float2*>(h->d_gen + h->gen_off_tw);
        gp.twn = reinterpret_cast<const float2*>(h->d_gen + h->gen_off_twn);
        gp.mel_range = reinterpret_cast<const int2*>(h->d_gen + h->gen_off_range);
        gp.mel_w = h->d_gen + h->gen_off_w;
        gp.out = out;
        gp.n_clips = n_clips;
        gp.T = target_length;
        gp.nmel = h->nmel;
        gp.L = h->frame_len;
        gp.S = h->frame_shift;
        gp.N = h->nfft;
        gp.log2M = 0;
        while ((2 << gp.log2M) < h->nfft) ++gp.log2M;
        gp.norm_mean = norm_mean;
        gp.norm_div = norm_div;
        gp.fill = p.fill;
        gp.floor_val = p.floor_val;
        gp.preemph = h->preemph;
        gp.skip_norm = skip_norm ? 1 : 0;
        const dim3 ggrid((target_length + kGenFramesPerCta - 1) / kGenFramesPerCta, n_clips);
        const size_t gsmem = generic_smem_bytes(h->nfft);
        if (pcm16) fbank_generic_kernel<short><<<ggrid, kGenWarps * 32, gsmem, stream>>>(gp);
        else fbank_generic_kernel<float><<<ggrid, kGenWarps * 32, gsmem, stream>>>(gp);
        le = cudaSuccess;
    } else {
        le = pcm16 ? launch_stream_kernel<short>(cfg, h->exact_mean, h->mel_spec, param_meta, trace, h->layout, p, h->meta_host)
                   : launch_stream_kernel<float>(cfg, h->exact_mean, h->mel_spec, param_meta, trace, h->layout, p, h->meta_host);
    }
    PSLA_CUDA(h, le);
    PSLA_CUDA(h, cudaGetLastError());
    if (slot) {
        PSLA_CUDA(h, cudaEventRecord(slot->done, stream));
        slot->pending = true;
    }
    if (h->timing) {
        PSLA_CUDA(h, cudaEventRecord(ev.e[2], stream));
        h->timed.push_back(ev);
    }
    h->launches += 1;
    return PSLA_OK;
}

int forward_any(psla_frontend_t* h, const void* wav, int64_t wav_samples, bool pcm16, const int64_t* wav_offset, const int64_t* wav_length, int n_wav,
                const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask, int n_clips, int target_length,
                float norm_mean, float norm_div, int skip_norm, float* out, void* workspace, size_t workspace_bytes, void* stream_) {
    if (!h || !wav_offset || !wav_length || !src1 || n_wav < 0 || n_clips < 0 || target_length < 1) return PSLA_E_INVALID;
    if (n_clips == 0) return PSLA_OK;
    if (!wav || !out) return PSLA_E_INVALID;
    if ((reinterpret_cast<uintptr_t>(wav) & 15) || (reinterpret_cast<uintptr_t>(out) & 3)) return PSLA_E_INVALID;
    if (n_clips > 65535) return PSLA_E_INVALID;  // gridDim.y of the statistics kernel
    const size_t need = psla_frontend_workspace_bytes(h, n_clips);
    if (!workspace || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255)) return PSLA_E_WORKSPACE;
    if (!skip_norm && !(norm_div != 0.f)) return PSLA_E_INVALID;
    if (h->layout != PSLA_OUT_TF_F32) {  // frequency-major: a group's 4 frames of a bin are one aligned vector store
        if (h->generic() || !h->mel_spec || h->nmel != kMaxMel || (target_length % kGroupFrames) != 0 || h->exact_mean || h->trace) return PSLA_E_UNSUPPORTED;
        if (reinterpret_cast<uintptr_t>(out) & 15) return PSLA_E_INVALID;
    }
    DeviceGuard guard(h->device);
    if (!guard.ok) return PSLA_E_CUDA;

    // ---- host-side descriptors (plain memory; they travel as kernel parameters or through a pinned slot below) ----
    const bool small = n_clips <= kParamClips;
    if (!small) {
        if (h->big_descs.size() < static_cast<size_t>(n_clips)) h->big_descs.resize(n_clips);
        if (h->big_costs.size() < static_cast<size_t>(n_clips) + 1) h->big_costs.resize(static_cast<size_t>(n_clips) + 1);
    }
    ClipDesc* descs = small ? h->meta_host.desc : h->big_descs.data();
    ClipCost* costs = small ? h->meta_host.cost : h->big_costs.data();
    const int G = (target_length + kGroupFrames - 1) / kGroupFrames;
    long long total_cost = 0;
    bool any_stats = false;

    // ---- validate + build descriptors (errors surface synchronously, like the reference's exceptions) ----
    // every waveform must lie inside the caller's buffer: a bad descriptor is an error here, not an out-of-bounds device read
    for (int i = 0; i < n_wav; ++i)
        if (wav_offset[i] < 0 || wav_length[i] < 0 || wav_length[i] > wav_samples || wav_offset[i] > wav_samples - wav_length[i]) return PSLA_E_INVALID;
    for (int b = 0; b < n_clips; ++b) {
        ClipDesc& d = descs[b];
        const int a = src1[b];
        if (a < 0 || a >= n_wav) return PSLA_E_INVALID;
        const int64_t n1 = wav_length[a];
        if (n1 < h->frame_len) return PSLA_E_SHORT_CLIP;  // kaldi.py:142
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
        d.flags = ((d.off1 & 3) == 0 ? kFlagAligned1 : 0) | ((d.off2 & 3) == 0 ? kFlagAligned2 : 0);
        // whole-clip means are needed in exact mode, and for the zero-padded tail of a shorter mixup partner
        if (h->exact_mean || h->generic() || (d.n2 > 0 && d.n2 < d.n1)) {
            d.flags |= kFlagStats;
            any_stats = true;
        }
        d.pad0 = d.pad1 = 0;
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
        const int64_t mfr = frames_of(h, n1);
        d.m_eff = static_cast<int>(mfr < target_length ? mfr : target_length);
        // dead groups (4 rows that are all padding / time-masked): [a_lo, a_hi) and [b_lo, G)
        ClipCost& cst = costs[b];
        int b_lo = (d.m_eff + kGroupFrames - 1) / kGroupFrames, a_lo = 0, a_hi = 0;
        if (d.t1 > d.t0) {
            const int lo = d.t0 < d.m_eff ? d.t0 : d.m_eff, hi = d.t1 < d.m_eff ? d.t1 : d.m_eff;
            if (hi >= d.m_eff) {  // the mask reaches the end of the computed rows: one dead run up to T
                b_lo = (lo + kGroupFrames - 1) / kGroupFrames;
            } else {
                a_lo = (lo + kGroupFrames - 1) / kGroupFrames;
                a_hi = hi / kGroupFrames;
                if (a_hi < a_lo) a_lo = a_hi = 0;
            }
        }
        if (b_lo > G) b_lo = G;
        if (a_hi > b_lo) a_hi = b_lo;
        if (a_lo > a_hi) a_lo = a_hi;
        cst.cum = total_cost;
        cst.a_lo = a_lo;
        cst.a_hi = a_hi;
        cst.b_lo = b_lo;
        cst.pad = 0;
        cst.pad2 = 0;
        total_cost += cost_before(h->cm, a_lo, a_hi, b_lo, G);
    }
    costs[n_clips].cum = total_cost;
    costs[n_clips].a_lo = costs[n_clips].a_hi = costs[n_clips].b_lo = costs[n_clips].pad = 0;
    costs[n_clips].pad2 = 0;

    // ---- enqueue ----
    // A batch above the parameter-space limit that needs no statistics pass goes out as several launches of <= kParamClips
    // clips each (balanced), every one with its descriptors as kernel parameters: no upload, and consecutive launches
    // overlap at their tails -- measured 1.98e6 instead of 1.73e6 clips/s at 256 eval clips.  (Not with the timeline or
    // the per-launch timing switched on: those look at one launch per call.)
    if (!small && !any_stats && h->param_meta && !h->trace && !h->timing) {
        const int n_chunks = (n_clips + kParamClips - 1) / kParamClips;
        const size_t clip_bytes = static_cast<size_t>(target_length) * h->nmel * out_elem_bytes(h);
        for (int c = 0; c < n_chunks; ++c) {
            const int c0 = static_cast<int>(static_cast<long long>(n_clips) * c / n_chunks);
            const int c1 = static_cast<int>(static_cast<long long>(n_clips) * (c + 1) / n_chunks);
            const int nc = c1 - c0;
            std::memcpy(h->meta_host.desc, descs + c0, static_cast<size_t>(nc) * sizeof(ClipDesc));
            for (int i = 0; i <= nc; ++i) {
                h->meta_host.cost[i] = costs[c0 + i];
                h->meta_host.cost[i].cum = costs[c0 + i].cum - costs[c0].cum;
            }
            h->meta_host.cost[nc].a_lo = h->meta_host.cost[nc].a_hi = h->meta_host.cost[nc].b_lo = 0;
            const int st = enqueue_batch(h, wav, pcm16, h->meta_host.desc, h->meta_host.cost, nc, costs[c1].cum - costs[c0].cum, false,
                                         target_length, norm_mean, norm_div, skip_norm,
                                         reinterpret_cast<float*>(reinterpret_cast<char*>(out) + static_cast<size_t>(c0) * clip_bytes), workspace,
                                         stream_, c > 0);
            if (st != PSLA_OK) return st;
        }
        return PSLA_OK;
    }
    return enqueue_batch(h, wav, pcm16, descs, costs, n_clips, total_cost, any_stats, target_length, norm_mean, norm_div, skip_norm, out,
                         workspace, stream_, false);
}

int forward_host_any(psla_frontend_t* h, const void* wav_host, int64_t wav_samples, bool pcm16, const int64_t* wav_offset, const int64_t* wav_length, int n_wav,
                     const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask, int n_clips, int target_length,
                     float norm_mean, float norm_div, int skip_norm, float* out_host, void* stream_) {
    if (!h || !wav_host || !wav_offset || !wav_length || n_wav < 1 || n_clips < 0 || target_length < 1) return PSLA_E_INVALID;
    if (n_clips == 0) return PSLA_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(h->device);
    if (!guard.ok) return PSLA_E_CUDA;
    int64_t total = 0;
    for (int i = 0; i < n_wav; ++i) {
        if (wav_offset[i] < 0 || wav_length[i] < 0 || wav_length[i] > wav_samples || wav_offset[i] > wav_samples - wav_length[i]) return PSLA_E_INVALID;
        if (wav_offset[i] + wav_length[i] > total) total = wav_offset[i] + wav_length[i];
    }
    if (total < 1) return PSLA_E_INVALID;
    const size_t wav_bytes = static_cast<size_t>(total) * (pcm16 ? sizeof(int16_t) : sizeof(float));
    const size_t out_bytes = static_cast<size_t>(n_clips) * target_length * h->nmel * out_elem_bytes(h);
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
    st = forward_any(h, h->d_wav, total, pcm16, wav_offset, wav_length, n_wav, src1, src2, lambda, mask, n_clips, target_length, norm_mean, norm_div,
                     skip_norm, h->d_out, h->d_ws, h->d_ws_cap, stream_);
    if (st != PSLA_OK) return st;
    if (out_host) PSLA_CUDA(h, cudaMemcpyAsync(out_host, h->d_out, out_bytes, cudaMemcpyDeviceToHost, stream));
    return PSLA_OK;
}
}  // namespace

extern "C" {

int psla_frontend_forward(psla_frontend_t* h, const float* wav, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length, int n_wav,
                          const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask, int n_clips,
                          int target_length, float norm_mean, float norm_div, int skip_norm, float* out, void* workspace,
                          size_t workspace_bytes, void* stream) {
    return forward_any(h, wav, wav_samples, false, wav_offset, wav_length, n_wav, src1, src2, lambda, mask, n_clips, target_length, norm_mean, norm_div,
                       skip_norm, out, workspace, workspace_bytes, stream);
}

int psla_frontend_forward_pcm16(psla_frontend_t* h, const int16_t* wav, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length, int n_wav,
                                const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask, int n_clips,
                                int target_length, float norm_mean, float norm_div, int skip_norm, float* out, void* workspace,
                                size_t workspace_bytes, void* stream) {
    return forward_any(h, wav, wav_samples, true, wav_offset, wav_length, n_wav, src1, src2, lambda, mask, n_clips, target_length, norm_mean, norm_div,
                       skip_norm, out, workspace, workspace_bytes, stream);
}

// No counterpart in the reference (see the header).  forward() over the library's static device scratch: with at most
// kParamClips clips and descriptors in parameter space, forward_any / enqueue_batch touch neither the host vectors nor a
// pinned slot, so the call consists of the validation loop and one or two cudaLaunchKernelEx.
int psla_frontend_forward_static(psla_frontend_t* h, const float* wav, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length,
                                 int n_wav, const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask, int n_clips,
                                 int target_length, float norm_mean, float norm_div, int skip_norm, float* out, void* stream) {
    if (!h) return PSLA_E_INVALID;
    if (n_clips > PSLA_STATIC_MAX_CLIPS) return PSLA_E_STATIC_LIMIT;
    if (!h->param_meta || h->timing || h->generic()) return PSLA_E_STATIC_LIMIT;  // a pinned descriptor slot / timing events would be created
    if (h->static_scratch_err != 0 || !h->static_scratch) {
        h->last_cuda = h->static_scratch_err ? h->static_scratch_err : static_cast<int>(cudaErrorInvalidSymbol);
        return PSLA_E_CUDA;
    }
    if (n_clips > 0 && psla_frontend_workspace_bytes(h, n_clips) > kStaticScratchBytes) return PSLA_E_STATIC_LIMIT;
    return forward_any(h, wav, wav_samples, false, wav_offset, wav_length, n_wav, src1, src2, lambda, mask, n_clips, target_length, norm_mean, norm_div,
                       skip_norm, out, h->static_scratch, kStaticScratchBytes, stream);
}

float* psla_frontend_device_out(const psla_frontend_t* h) { return h ? h->d_out : nullptr; }

int psla_frontend_forward_host(psla_frontend_t* h, const float* wav_host, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length, int n_wav,
                               const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask, int n_clips,
                               int target_length, float norm_mean, float norm_div, int skip_norm, float* out_host, void* stream) {
    return forward_host_any(h, wav_host, wav_samples, false, wav_offset, wav_length, n_wav, src1, src2, lambda, mask, n_clips, target_length, norm_mean,
                            norm_div, skip_norm, out_host, stream);
}

int psla_frontend_forward_host_pcm16(psla_frontend_t* h, const int16_t* wav_host, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length,
                                     int n_wav, const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask,
                                     int n_clips, int target_length, float norm_mean, float norm_div, int skip_norm, float* out_host,
                                     void* stream) {
    return forward_host_any(h, wav_host, wav_samples, true, wav_offset, wav_length, n_wav, src1, src2, lambda, mask, n_clips, target_length, norm_mean,
                            norm_div, skip_norm, out_host, stream);
}

}  // extern "C"
