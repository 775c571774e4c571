// Device code of libpsla_frontend.so: the per-clip statistics kernel and the warp-autonomous streaming fbank kernel.
// (Host side / C ABI: psla_frontend.cu.  Per-lane FFT / mel phases: fbank_stream.cuh.)
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "fbank_stream.cuh"

namespace psla {

// output layouts (psla_frontend_set_output_layout): values of PSLA_OUT_* in include/psla_frontend.h
constexpr int kOutTF = 0, kOutFT = 1, kOutFTbf16 = 2;

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
    int flags;             // bit 0: off1 % 4 == 0, bit 1: off2 % 4 == 0 (128-bit loads allowed), bit 2: clip statistics needed
    int pad0, pad1;
};
static_assert(sizeof(ClipDesc) == 64, "ClipDesc layout");
constexpr int kFlagAligned1 = 1, kFlagAligned2 = 2, kFlagStats = 4;

// Work partition: a group (4 output rows) costs `live` if any of its rows carries computed features (+ `cold` if it is
// the first group of a run of live groups: its whole 880-sample window is staged without a prefetch), `dead` if all four
// rows are padding / time-masked (constant rows).  Dead groups of a clip are [a_lo, a_hi) and [b_lo, G).
struct CostModel {
    int live, dead, cold;
};
struct alignas(16) ClipCost {  // 32 bytes; entry n_clips holds the total in `cum`
    long long cum;  // cost of all clips before this one
    int a_lo, a_hi, b_lo, pad;
    long long pad2;
};
static_assert(sizeof(ClipCost) == 32, "ClipCost layout");
// measured on B200 (scripts/gpu_trace.py): a live group ~10.3k cycles, a cold start ~+8k, a dead group ~0.3k in a run
constexpr CostModel kDefaultCost = {64, 2, 48};

constexpr int kStatSegs = 32;  // partial sums per clip: one CTA each (8 left the pass latency-bound: 17 us for the 6 clips of an FSD50K batch of 24 that need it)
constexpr int kStatVals = 3;  // sum(w1), sum(w2), sum(w2[0:min(n1,n2)])

constexpr int kWarpsPerCta = 4;   // four CTAs per SM
constexpr int kThreads = kWarpsPerCta * 32;

struct StreamParams {
    const void* wav;  // fp32 samples, or int16 PCM (value / 32768, what a decoder's float conversion yields) -- see Pcm<S>
    const ClipDesc* desc;
    const ClipCost* cost;
    const double* partials;
    const float* tab;
    float* out;
    long long total_cost;
    int tab_floats;
    int n_clips;
    int T;  // target_length
    int G;  // groups per clip = ceil(T / 4)
    int nmel;
    float norm_mean, norm_div, fill;  // fill = (0 - mean) / div in IEEE fp32 (0 if skip_norm)
    float floor_val;                  // (log(eps) - mean) / div in IEEE fp32: the value of a cell on the log floor
    float log_scale, log_bias;        // fast log: value = lg2(e) * log_scale + log_bias
    int skip_norm;
    int fast_log;
    float preemph, c1;
    CostModel cm;
    long long* trace;  // optional [gridDim.x * kWarpsPerCta][kTraceWords] per-warp timeline (psla_frontend_set_trace), else nullptr
};
// words per warp: [0] globaltimer at warp start (ns)  [1] at warp end  [2] clock64 cycles before the first group | total << 32
// [3] smid | live groups << 16 | dead groups << 32 | cold starts << 48  [4 + i] i-th loop step of the warp (i < kTraceSteps):
// cycles since warp start | kind << 40 (1 live, 2 cold-started live, 3 run of dead groups; 4.. phase markers inside a step:
// 4 clip entered, 5 cold window staged, 6 pass 1 done, 7 pass 2 + power done, 8 mel done, 9 epilogue done, 10 cold loads issued, 11 prefetch issued) | groups << 44
constexpr int kTraceSteps = 124;
constexpr int kTraceWords = 4 + kTraceSteps;

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
__device__ __forceinline__ ClipCost load_cost(const ClipCost* __restrict__ p) {
    union {
        ClipCost c;
        int4 v[2];
    } u;
    const int4* src = reinterpret_cast<const int4*>(p);
    u.v[0] = __ldg(src);
    u.v[1] = __ldg(src + 1);
    return u.c;
}

// Where the kernel reads the clip descriptors and the cost table from: device memory (any batch size; uploaded per call),
// or, for batches of at most kParamClips clips, the kernel's own parameter space (no upload, nothing between two launches
// on the stream, which is what lets consecutive launches overlap; see psla_frontend.cu).
constexpr int kParamClips = 128;
struct MetaBlock {
    ClipCost cost[kParamClips + 1];
    ClipDesc desc[kParamClips];
};
struct MetaNone {
    int unused;
};
struct GlobalMeta {
    const ClipDesc* desc_;
    const ClipCost* cost_;
    static constexpr bool kWarpSearch = true;  // lanes may read different entries at full speed
    __device__ __forceinline__ ClipDesc desc(int b) const { return load_desc(desc_ + b); }
    __device__ __forceinline__ ClipCost cost(int b) const { return load_cost(cost_ + b); }
    __device__ __forceinline__ long long cum(int b) const { return __ldg(&cost_[b].cum); }
};
struct ParamMeta {
    const MetaBlock* blk;
    static constexpr bool kWarpSearch = false;  // constant-bank reads want a warp-uniform index
    __device__ __forceinline__ ClipDesc desc(int b) const { return blk->desc[b]; }
    __device__ __forceinline__ ClipCost cost(int b) const { return blk->cost[b]; }
    __device__ __forceinline__ long long cum(int b) const { return blk->cost[b].cum; }
};

// Sample formats.  Pcm<S>::reg_t holds 4 consecutive samples as loaded (fp32: 16 bytes, int16: 8 bytes); the conversion to
// fp32 happens when the samples are consumed, so that a prefetch is nothing but loads.  int16 PCM converts as
// value * 2^-15, exactly what a decoder's float conversion yields (torchaudio.load normalises int16 by 32768).
template <class S>
struct Pcm;
template <>
struct Pcm<float> {
    using reg_t = float4;
    static __device__ __forceinline__ reg_t zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
    static __device__ __forceinline__ reg_t load_vec(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
    static __device__ __forceinline__ float load1(const float* p) { return __ldg(p); }
    static __device__ __forceinline__ reg_t pack(float x, float y, float z, float w) { return make_float4(x, y, z, w); }
    static __device__ __forceinline__ float4 to_float4(const reg_t& r) { return r; }
};
template <>
struct Pcm<short> {
    using reg_t = uint2;
    static constexpr float kScale = 1.0f / 32768.0f;
    static __device__ __forceinline__ reg_t zero() { return make_uint2(0u, 0u); }
    static __device__ __forceinline__ reg_t load_vec(const short* p) { return __ldg(reinterpret_cast<const uint2*>(p)); }
    static __device__ __forceinline__ float load1(const short* p) { return static_cast<float>(__ldg(p)) * kScale; }
    static __device__ __forceinline__ reg_t pack(float x, float y, float z, float w) {  // values are exact int16 / 32768
        const unsigned a = static_cast<unsigned short>(static_cast<short>(x * 32768.0f)), b = static_cast<unsigned short>(static_cast<short>(y * 32768.0f));
        const unsigned c = static_cast<unsigned short>(static_cast<short>(z * 32768.0f)), d = static_cast<unsigned short>(static_cast<short>(w * 32768.0f));
        return make_uint2(a | (b << 16), c | (d << 16));
    }
    static __device__ __forceinline__ float4 to_float4(const reg_t& r) {
        float4 v;
        v.x = static_cast<float>(static_cast<short>(r.x & 0xffffu)) * kScale;
        v.y = static_cast<float>(static_cast<int>(r.x) >> 16) * kScale;
        v.z = static_cast<float>(static_cast<short>(r.y & 0xffffu)) * kScale;
        v.w = static_cast<float>(static_cast<int>(r.y) >> 16) * kScale;
        return v;
    }
};

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Programmatic dependent launch (PDL), the two halves every launch of the streaming kernel executes:
//   pdl_release()  at kernel entry: the NEXT kernel on the stream, if the host launched it with programmatic stream
//                  serialization, may start its thread blocks as soon as SM resources free up;
//   pdl_join()     on every exit path, after the thread's last global store: wait until the grid(s) this launch was
//                  allowed to start ahead of have completed and flushed.  A launch therefore never COMPLETES before its
//                  predecessor -- stream order stays transitive for whatever is enqueued next (a model step, a copy, an
//                  event) -- and because waiting thread blocks keep their SM slots, at most two consecutive launches are
//                  ever in flight together.  When the launch was not made with the attribute both are no-ops.
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_join() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Level of a waveform from FIVE samples spread over it (first, quartiles, last): their median.  The default clip-mean mode
// subtracts this instead of the clip mean (see "Waveform conditioning" below).  One sample would do for the algebra -- any
// constant is removed again by the per-frame DC removal -- but not for fp32: if that sample is an outlier (a click, a
// clipped sample, a loud onset before a quiet body) every staged value sits at the outlier's magnitude and the quiet
// content loses 1 ulp of IT per sample -- measured: 1 % of the cells of a quiet 16-bit clip with a full-scale click at
// sample 0 off by up to 1.2e-2 (profiles/r02a/parity.json).  The median of five keeps |x - level| within ~2x of |x - mean|
// for any clip with up to two outliers among the five, at the cost of five scattered loads per cold start.
__device__ __forceinline__ float median3(float x, float y, float z) { return fmaxf(fminf(x, y), fminf(fmaxf(x, y), z)); }
template <class S>
__device__ __forceinline__ float robust_level(const S* __restrict__ w, int n) {
    const int q1 = n >> 2, q2 = n >> 1;
    const float a = Pcm<S>::load1(w), b = Pcm<S>::load1(w + q1), c = Pcm<S>::load1(w + q2), d = Pcm<S>::load1(w + q1 + q2),
                e = Pcm<S>::load1(w + n - 1);
    const float f = fmaxf(fminf(a, b), fminf(c, d));  // drops the lowest of the first four
    const float g = fminf(fmaxf(a, b), fmaxf(c, d));  // drops the highest
    return median3(e, f, g);
}

// 4 consecutive samples [i, i+4) of a waveform of n samples, zeros past the end
template <class S>
__device__ __forceinline__ typename Pcm<S>::reg_t load4(const S* __restrict__ w, int i, int n, bool aligned) {
    if (aligned && i + 3 < n) return Pcm<S>::load_vec(w + i);
    const float x = (i < n) ? Pcm<S>::load1(w + i) : 0.f;
    const float y = (i + 1 < n) ? Pcm<S>::load1(w + i + 1) : 0.f;
    const float z = (i + 2 < n) ? Pcm<S>::load1(w + i + 2) : 0.f;
    const float v = (i + 3 < n) ? Pcm<S>::load1(w + i + 3) : 0.f;
    return Pcm<S>::pack(x, y, z, v);
}

// sum of w[lo, hi) by one 256-thread CTA's thread `tid` (strided groups of 4 samples, 4 loads in flight)
template <class S>
__device__ __forceinline__ float strided_sum(const S* __restrict__ w, int lo, int hi, int tid, bool aligned) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int i = lo + 4 * tid;
    if (aligned) {
        for (; i + 3 * 1024 + 3 < hi; i += 4 * 1024) {
            const float4 a = Pcm<S>::to_float4(Pcm<S>::load_vec(w + i));
            const float4 b = Pcm<S>::to_float4(Pcm<S>::load_vec(w + i + 1024));
            const float4 c = Pcm<S>::to_float4(Pcm<S>::load_vec(w + i + 2048));
            const float4 d = Pcm<S>::to_float4(Pcm<S>::load_vec(w + i + 3072));
            s0 += (a.x + a.y) + (a.z + a.w);
            s1 += (b.x + b.y) + (b.z + b.w);
            s2 += (c.x + c.y) + (c.z + c.w);
            s3 += (d.x + d.y) + (d.z + d.w);
        }
    }
    for (; i < hi; i += 1024) {
        const float4 v = Pcm<S>::to_float4(load4<S>(w, i, hi, aligned));
        s0 += (v.x + v.y) + (v.z + v.w);
    }
    return (s0 + s1) + (s2 + s3);
}

// ------------------------------------------------------------------------------------------------
// Kernel 1 (only launched when some clip needs it): per-clip sums for the whole-clip means of
// audioset_dataset.py:92 / :98-99 / :117.  grid = (kStatSegs, B), block = 256.  Deterministic: fixed partition, fixed
// reduction tree; the streaming kernel adds the kStatSegs partials in a fixed order.
// ------------------------------------------------------------------------------------------------
// PMETA: the descriptors come in the kernel's parameter space (batches of <= kParamClips clips: nothing is uploaded).
template <class S, bool PMETA>
__global__ void __launch_bounds__(256) clip_stats_kernel(const S* __restrict__ wav, const ClipDesc* __restrict__ desc,
                                                         const __grid_constant__ std::conditional_t<PMETA, MetaBlock, MetaNone> meta,
                                                         double* __restrict__ partials, int exact) {
    const int b = blockIdx.y, g = blockIdx.x, tid = threadIdx.x;
    // part of the launch chain (psla_frontend.cu): the streaming kernel behind it may start at once and joins it where a
    // clip needs these sums; this kernel itself joins ITS predecessor before it exits (stream order stays transitive)
    pdl_release();
    ClipDesc d;
    if constexpr (PMETA) d = meta.desc[b];
    else d = load_desc(desc + b);
    if (!(d.flags & kFlagStats)) {
        pdl_join();
        return;
    }
    // exact == 0 (default clip-mean mode): only the partner's mean is used (enter_clip), the primary waveform is not read
    float s1 = 0.f, s2 = 0.f, s2c = 0.f;
    if (exact) {
        const int seg = ((d.n1 + 4 * kStatSegs - 1) / (4 * kStatSegs)) * 4;
        const int lo = min(d.n1, g * seg), hi = min(d.n1, lo + seg);
        s1 = strided_sum(wav + d.off1, lo, hi, tid, d.flags & kFlagAligned1);
    }
    if (d.n2 > 0) {
        const int nmin = min(d.n1, d.n2);
        const int seg = ((d.n2 + 4 * kStatSegs - 1) / (4 * kStatSegs)) * 4;
        const int lo = min(d.n2, g * seg), hi = min(d.n2, lo + seg);
        const bool al = (d.flags & kFlagAligned2) != 0;
        s2 = strided_sum(wav + d.off2, lo, hi, tid, al);
        if (exact) {
            if (nmin >= hi) s2c = s2;                                               // segment entirely kept
            else if (nmin > lo) s2c = strided_sum(wav + d.off2, lo, nmin, tid, al);  // the one straddling segment
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
#pragma unroll
        for (int w = 0; w < 8; ++w) s += red[w][tid];
        partials[(static_cast<size_t>(b) * kStatSegs + g) * kStatVals + tid] = s;
    }
    pdl_join();
}

// ------------------------------------------------------------------------------------------------
// Per-clip state of a warp.  It lives in the warp's shared-memory window, not in registers: the FFT phases need
// almost the whole 128-register budget, and anything loop-carried would be spilled to local memory around them.
// Every use is a volatile shared-memory read (so the compiler cannot cache it in a register across the FFT).
// ------------------------------------------------------------------------------------------------
// Words that are used together share a 16-byte quad, so one LDS.128 (cs_q) brings them.
enum ClipStateWord {
    CS_N1 = 0, CS_NMIN, CS_L1, CS_L2,                                    // quad 0: conditioning
    CS_SUB1, CS_MEAN2, CS_MIXMEAN, CS_FLAGS,                             // quad 1: conditioning
    CS_OFF1 /* 2 words */, CS_OFF1_HI, CS_OFF2 /* 2 words */, CS_OFF2_HI,  // quad 2: waveform addresses
    CS_F0, CS_F1, CS_T0, CS_T1,                                          // quad 3: masks
    CS_MEFF, CS_ALO, CS_AHI, CS_BLO,                                     // quad 4: live rows / dead group ranges
    // loop state of the warp (same reason: nothing may stay live in registers across the FFT phases)
    CS_B, CS_G, CS_BUDGET, CS_NEXT /* kNext* bits for the group being processed */,  // quad 5
    CS_CARRY, CS_TR_LIVE, CS_TR_DEAD, CS_TR_COLD,                        // quad 6
    CS_TR_T0 /* 2 words */, CS_TR_T0_HI, CS_TR_C0 /* 2 words */, CS_TR_C0_HI, CS_TR_STEP, CS_END
};
constexpr int kNextLive = 1, kNextFull = 2;
static_assert(CS_END <= kCsWords && CS_SUB1 % 4 == 0 && CS_OFF1 % 4 == 0 && CS_F0 % 4 == 0 && CS_MEFF % 4 == 0 && CS_B % 4 == 0 &&
                  CS_CARRY % 4 == 0 && CS_TR_T0 % 4 == 0 && kWsCS % 4 == 0 && kWarpFloats % 4 == 0,
              "clip-state layout");
constexpr int kCsMix = 8;  // CS_FLAGS: kFlagAligned1 | kFlagAligned2 | kCsMix | (number of leading groups with a full window) << kCsFullShift
constexpr int kCsFullShift = 4;

__device__ __forceinline__ int cs_i(const int* cs, int w) { return *reinterpret_cast<const volatile int*>(cs + w); }
__device__ __forceinline__ float cs_f(const int* cs, int w) { return __int_as_float(cs_i(cs, w)); }
__device__ __forceinline__ long long cs_ll(const int* cs, int w) { return *reinterpret_cast<const volatile long long*>(cs + w); }
// one quad (w % 4 == 0); volatile like the scalar reads
__device__ __forceinline__ int4 cs_q(const int* cs, int w) {
    int4 v;
    asm volatile("ld.volatile.shared.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(static_cast<unsigned>(__cvta_generic_to_shared(cs + w))));
    return v;
}
__device__ __forceinline__ void cs_set_q(int* cs, int w, int a, int b, int c, int d) {
    asm volatile("st.volatile.shared.v4.s32 [%0], {%1, %2, %3, %4};" ::"r"(static_cast<unsigned>(__cvta_generic_to_shared(cs + w))), "r"(a),
                 "r"(b), "r"(c), "r"(d)
                 : "memory");
}

// ------------------------------------------------------------------------------------------------
// Waveform conditioning of one clip (audioset_dataset.py:91-117).
//   EXACT   x = (l1*(a - mean1) + l2*(c' - mean2)) - mixmean in the reference's fp32 op order, c' the partner zero-padded
//           AFTER its mean removal (:104-106) or cut (:109); the three means come from clip_stats_kernel.
//   fused   every constant the reference subtracts from the whole clip is removed again by the per-frame DC removal
//           (kaldi.py:183-186), so x + const gives the same features (up to fp32 rounding).  The constant used here is
//           a five-sample median of each waveform (robust_level) instead of its mean -- no pass over the clip is
//           needed, and the values stay small when the recording has a DC offset:   x' = l1*a + l2*c - (l1*A + l2*C).
//           The one place a mean survives is the zero-padded tail of a SHORTER partner: there c' - mean2 = 0, i.e.
//           c is replaced by mean2 (clip statistics are computed for those clips only).
// FULL: the staged range lies inside both waveforms and both are 16-byte aligned -- no bounds checks, no selects.
// ------------------------------------------------------------------------------------------------
template <class S>
struct MixState {
    const S* w1;
    const S* w2;
    float l1, l2;
    float sub1;     // EXACT: mean1;  fused: K = l1*level(a) (+ l2*level(c))
    float mean2;    // EXACT: mean2;  fused: mean2 where the partner is shorter (tail fill)
    float mixmean;  // EXACT only
    int n1, nmin;
    int gfull1;  // groups [0, gfull1) of the clip can be staged without bounds checks (computed once per clip, enter_clip)
    bool mix, al1, al2;
};

template <class S>
__device__ __forceinline__ MixState<S> load_mix(const int* cs, const void* wav_) {
    const S* wav = static_cast<const S*>(wav_);
    MixState<S> m;
    const int4 q0 = cs_q(cs, CS_N1), q1 = cs_q(cs, CS_SUB1), q2 = cs_q(cs, CS_OFF1);
    const int fl = q1.w;
    (void)wav;  // quad 2 holds the waveforms' ADDRESSES (enter_clip adds the buffer base once per clip, not twice per step)
    m.w1 = reinterpret_cast<const S*>((static_cast<unsigned long long>(static_cast<unsigned>(q2.y)) << 32) | static_cast<unsigned>(q2.x));
    m.w2 = reinterpret_cast<const S*>((static_cast<unsigned long long>(static_cast<unsigned>(q2.w)) << 32) | static_cast<unsigned>(q2.z));
    m.n1 = q0.x;
    m.nmin = q0.y;
    m.l1 = __int_as_float(q0.z);
    m.l2 = __int_as_float(q0.w);
    m.sub1 = __int_as_float(q1.x);
    m.mean2 = __int_as_float(q1.y);
    m.mixmean = __int_as_float(q1.z);
    m.mix = (fl & kCsMix) != 0;
    m.al1 = (fl & kFlagAligned1) != 0;
    m.al2 = (fl & kFlagAligned2) != 0;
    m.gfull1 = fl >> kCsFullShift;
    return m;
}

// the window of group g, [640 g, 640 g + 880), can be staged without bounds checks: it lies inside both waveforms and both
// are 16-byte aligned
template <class S>
__device__ __forceinline__ bool window_full(const MixState<S>& m, int g) {
    return g < m.gfull1;
}
__device__ __forceinline__ int full_groups(int n1, int nmin, bool mix, int flags) {
    const bool ok = (flags & kFlagAligned1) && (!mix || (flags & kFlagAligned2));
    const int lim = mix ? nmin : n1;
    return (ok && lim >= kGroupSamples) ? (lim - kGroupSamples) / kStepSamples + 1 : 0;
}

template <bool FULL, class S>
__device__ __forceinline__ float mix_sample_exact(const MixState<S>& m, float a, float c, int s) {
    if (!FULL && s >= m.n1) return 0.f;
    const float v1 = __fsub_rn(a, m.sub1);
    if (!m.mix) return v1;
    const float v2 = (FULL || s < m.nmin) ? __fsub_rn(c, m.mean2) : 0.f;
    return __fsub_rn(__fadd_rn(__fmul_rn(m.l1, v1), __fmul_rn(m.l2, v2)), m.mixmean);
}
// 4 consecutive samples starting at s.  The fused form is one expression for mixed and plain clips: a plain clip carries
// l1 = 1, l2 = 0 (enter_clip), and 1*a + (0*c - K) == a - K bit for bit -- two two-wide FMAs per sample pair, no selects.
template <bool EXACT, bool FULL, class S>
__device__ __forceinline__ float4 mix4(const MixState<S>& m, const float4& a, float4 c, int s) {
    float4 x;
    if (EXACT) {
        x.x = mix_sample_exact<FULL>(m, a.x, c.x, s);
        x.y = mix_sample_exact<FULL>(m, a.y, c.y, s + 1);
        x.z = mix_sample_exact<FULL>(m, a.z, c.z, s + 2);
        x.w = mix_sample_exact<FULL>(m, a.w, c.w, s + 3);
        return x;
    }
    if (!FULL) {  // past the partner's end its (removed) mean takes the place of the zero padding
        c.x = (s < m.nmin) ? c.x : m.mean2;
        c.y = (s + 1 < m.nmin) ? c.y : m.mean2;
        c.z = (s + 2 < m.nmin) ? c.z : m.mean2;
        c.w = (s + 3 < m.nmin) ? c.w : m.mean2;
    }
    const cf nk = bc(-m.sub1);
    const cf lo = cfma(bc(m.l1), mk(a.x, a.y), cfma(bc(m.l2), mk(c.x, c.y), nk));
    const cf hi = cfma(bc(m.l1), mk(a.z, a.w), cfma(bc(m.l2), mk(c.z, c.w), nk));
    x = make_float4(lo.x, lo.y, hi.x, hi.y);
    if (!FULL) {
        x.x = (s < m.n1) ? x.x : 0.f;
        x.y = (s + 1 < m.n1) ? x.y : 0.f;
        x.z = (s + 2 < m.n1) ? x.z : 0.f;
        x.w = (s + 3 < m.n1) ? x.w : 0.f;
    }
    return x;
}

// ------------------------------------------------------------------------------------------------
// Warp-private staging of window samples [U0, U1) of the group window ([0, 240): the part a group shares with its
// predecessor, only staged on a cold start; [240, 880): the 640 new samples of every step).  issue() only loads (so the
// latency can hide behind the mel / epilogue of the previous group), commit() conditions the samples and writes
//   WS[wpos(u, par)] = x[u] - p*x[u-1],  BS[par][u/16] = sum of 16 samples,  XS[par][u/160] = (1-p)*x[u]  (u % 160 == 0),
// where the part the NEXT group shares (u >= 640) lands in the other parity's arrays at u - 640: samples in OV[par ^ 1]
// only (this group's pass 1 reads them there), block sums in both arrays (this group's last frame needs them too).
// Nothing is ever moved between steps.
// `carry` is x of the last sample staged so far (pre-emphasis neighbour of the next staged sample).
// ------------------------------------------------------------------------------------------------
template <int U0, int U1>
struct StageGeom {
    static_assert(U0 % 16 == 0 && U1 % 16 == 0 && U0 < U1, "whole 16-sample blocks");
    static constexpr int kFloat4 = (U1 - U0) / 4;         // 60 or 160
    static constexpr int kIters = (kFloat4 + 31) / 32;    // 2 or 5
    static constexpr int kLastLane = (kFloat4 - 1) % 32;  // lane holding the range's last float4
};

template <int U0, int U1, bool FULL, class S, int NIT>
__device__ __forceinline__ void stage_issue(const MixState<S>& m, int S0, int lane, typename Pcm<S>::reg_t (&a)[NIT],
                                            typename Pcm<S>::reg_t (&c)[NIT]) {
    using Geo = StageGeom<U0, U1>;
    static_assert(NIT >= Geo::kIters, "register arrays too small");
    if (FULL && Geo::kFloat4 % 32 == 0) {  // whole iterations, no bounds: one base pointer per waveform, immediate offsets
        const S* p1 = m.w1 + (S0 + U0 + 4 * lane);
#pragma unroll
        for (int it = 0; it < Geo::kIters; ++it) a[it] = Pcm<S>::load_vec(p1 + 128 * it);
        if (m.mix) {
            const S* p2 = m.w2 + (S0 + U0 + 4 * lane);
#pragma unroll
            for (int it = 0; it < Geo::kIters; ++it) c[it] = Pcm<S>::load_vec(p2 + 128 * it);
        } else {
#pragma unroll
            for (int it = 0; it < Geo::kIters; ++it) c[it] = Pcm<S>::zero();
        }
        return;
    }
#pragma unroll
    for (int it = 0; it < Geo::kIters; ++it) {
        const int L = lane + 32 * it;
        const int s = S0 + U0 + 4 * L;
        a[it] = Pcm<S>::zero();
        c[it] = a[it];
        if ((it + 1) * 32 <= Geo::kFloat4 || L < Geo::kFloat4) {
            if (FULL) {
                a[it] = Pcm<S>::load_vec(m.w1 + s);
                if (m.mix) c[it] = Pcm<S>::load_vec(m.w2 + s);
            } else {
                a[it] = load4<S>(m.w1, s, m.n1, m.al1);
                if (m.mix) c[it] = load4<S>(m.w2, s, m.nmin, m.al2);
            }
        }
    }
}

template <bool EXACT, int U0, int U1, bool FULL, class S, int NIT>
__device__ __forceinline__ void stage_commit(const MixState<S>& m, int S0, int lane, float pe, float c1,
                                             const typename Pcm<S>::reg_t (&a)[NIT], const typename Pcm<S>::reg_t (&c)[NIT],
                                             float* __restrict__ WS, int par, float& carry) {
    using Geo = StageGeom<U0, U1>;
    float* BSp = WS + kWsBS + par * kBsFloats;
    float* BSo = WS + kWsBS + (par ^ 1) * kBsFloats;
    float* XSp = WS + kWsXS + par * kXsFloats;
    float* XSo = WS + kWsXS + (par ^ 1) * kXsFloats;
    // piece bases with the piece's first window sample folded in: sample u of the piece sits at base[u]
    float* p_lo = WS + kWsOV + par * kOvFloats;                // u < 240
    float* p_mid = WS + kWsDM - kOverlapSamples;               // 240 <= u < 640 (+ 16 from u = 320 on)
    float* p_hi = WS + kWsOV + (par ^ 1) * kOvFloats - kStepSamples;  // u >= 640
    float last_w = carry;
#pragma unroll
    for (int it = 0; it < Geo::kIters; ++it) {
        const int L = lane + 32 * it;
        const int u = U0 + 4 * L;
        const int s = S0 + u;
        const bool in = ((it + 1) * 32 <= Geo::kFloat4) || L < Geo::kFloat4;
        const float4 fa = Pcm<S>::to_float4(a[it]), fc = Pcm<S>::to_float4(c[it]);
        const float4 x = mix4<EXACT, FULL>(m, fa, fc, s);
        // the pre-emphasis neighbour x[u - 1]: lane l - 1's last sample of THIS iteration, for lane 0 lane 31's last sample
        // of the PREVIOUS one (the incoming carry at first) -- one rotating shuffle: lane 31 sends the older value
        const float prev = __shfl_sync(0xffffffffu, lane == 31 ? last_w : x.w, (lane + 31) & 31);
        last_w = x.w;
        float s4 = (x.x + x.y) + (x.z + x.w);
        s4 += __shfl_xor_sync(0xffffffffu, s4, 1);
        s4 += __shfl_xor_sync(0xffffffffu, s4, 2);
        if (in) {
            float4 dd;
            dd.x = fmaf(-pe, prev, x.x);
            dd.y = fmaf(-pe, x.x, x.y);
            dd.z = fmaf(-pe, x.y, x.z);
            dd.w = fmaf(-pe, x.z, x.w);
            // the iteration's first / last float4 decide at compile time which pieces it can touch
            constexpr int kU0 = U0;
            const int u_first = kU0 + 128 * it, u_last = u_first + 124;
            float* dst;
            if (u_last < kOverlapSamples) dst = p_lo + u;
            else if (u_first >= kStepSamples) dst = p_hi + u;
            else if (u_first >= 320 && u_last < kStepSamples) dst = p_mid + 16 + u;
            else if (u_first >= kOverlapSamples && u_last < 320) dst = p_mid + u;
            else if (u_first >= kOverlapSamples && u_last < kStepSamples) dst = p_mid + u + (u >= 320 ? 16 : 0);  // straddles the padding
            else if (u_first >= 320 && u_first < kStepSamples) dst = (u < kStepSamples) ? p_mid + 16 + u : p_hi + u;  // straddles 640
            else dst = (u < kOverlapSamples) ? p_lo + u : (u < kStepSamples ? p_mid + u + (u >= 320 ? 16 : 0) : p_hi + u);
            *reinterpret_cast<float4*>(dst) = dd;
            if ((lane & 3) == 0) {
                const int blk = U0 / 16 + (L >> 2);
                BSp[blk] = s4;
                if (u_last >= kStepSamples && blk >= kStepSamples / 16) BSo[blk - kStepSamples / 16] = s4;
            }
            // u % 160 == 0  <=>  (U0/4 + L) % 40 == 0: at most one lane per iteration, known at compile time
            const int q = (U0 / 4 + 32 * it + 39) / 40;  // first multiple of 40 at or after this iteration's first float4
            if (U0 / 4 + L == 40 * q) {
                if (q < 4) XSp[xs_pos(q)] = c1 * x.x;
                else XSo[xs_pos(q - 4)] = c1 * x.x;
            }
        }
    }
    carry = __shfl_sync(0xffffffffu, last_w, Geo::kLastLane);
}

// ------------------------------------------------------------------------------------------------
// Work partition helpers (see ClipCost / CostModel).  psla_frontend.cu evaluates the same cost_before() on the host.
// ------------------------------------------------------------------------------------------------
// the group opens a run of live groups of its clip (group 0, or the group after the dead range [a_lo, a_hi))
__host__ __device__ __forceinline__ bool run_start(int a_lo, int a_hi, int g) { return g == 0 || (g == a_hi && a_lo < a_hi); }
__host__ __device__ __forceinline__ bool is_dead(int a_lo, int a_hi, int b_lo, int g) { return (g >= a_lo && g < a_hi) || g >= b_lo; }
// cost of groups [0, g) of the clip
__host__ __device__ __forceinline__ int cost_before(const CostModel& cm, int a_lo, int a_hi, int b_lo, int g) {
    const int ga = g < a_lo ? a_lo : (g > a_hi ? a_hi : g);
    const int n_dead = (ga - a_lo) + ((g > b_lo ? g : b_lo) - b_lo);
    const bool has_a = a_lo < a_hi;
    const int starts = ((g > 0 && (has_a ? a_lo : b_lo) > 0) ? 1 : 0) + ((has_a && a_hi < b_lo && g > a_hi) ? 1 : 0);
    return cm.live * (g - n_dead) + cm.dead * n_dead + cm.cold * starts;
}

// First group (b, g) in clip-major order whose preceding cost is >= target; returns that preceding cost.
// Warp-cooperative (all lanes return the same).  From device memory the clip is found by a 128-ary search over the
// cumulative costs, one round trip to L2 per level instead of one per bit; from parameter space by a plain binary search.
template <class MS>
__device__ __forceinline__ long long find_start(const MS& ms, const CostModel& cm, int n_clips, int G, long long target, int lane,
                                                int& b_out, int& g_out) {
    int lo = 0, hi = n_clips;  // largest b in [lo, hi] with cum[b] <= target (cum[0] == 0)
    if (MS::kWarpSearch) {
        while (lo < hi) {
            const int step = (hi - lo + 127) >> 7;
            long long v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = ms.cum(min(hi, lo + (4 * lane + j + 1) * step));
            int cnt = 0;  // candidates lo + k*step, k = 1..128 (clipped to hi) are monotone: the first cnt of them pass
#pragma unroll
            for (int j = 0; j < 4; ++j) cnt += __popc(__ballot_sync(0xffffffffu, v[j] <= target));
            const int nlo = min(hi, lo + cnt * step);
            if (cnt < 128) hi = min(hi, lo + (cnt + 1) * step) - 1;
            lo = nlo;
        }
    } else {
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (ms.cum(mid) <= target) lo = mid;
            else hi = mid - 1;
        }
    }
    b_out = lo;
    g_out = 0;
    const ClipCost c = ms.cost(lo);
    if (lo >= n_clips) return c.cum;
    const int rem = static_cast<int>(target - c.cum);
    if (rem <= 0) return c.cum;
    int glo = 1, ghi = G;  // smallest g with cost_before(g) >= rem  (cost_before(G) > rem)
    while (glo < ghi) {
        const int mid = (glo + ghi) >> 1;
        if (cost_before(cm, c.a_lo, c.a_hi, c.b_lo, mid) >= rem) ghi = mid;
        else glo = mid + 1;
    }
    if (glo >= G) {
        b_out = lo + 1;
        return ms.cum(lo + 1);
    }
    g_out = glo;
    return c.cum + cost_before(cm, c.a_lo, c.a_hi, c.b_lo, glo);
}

// `count` floats starting at `ptr` := v, by the whole warp (16-byte stores once aligned)
__device__ __forceinline__ void warp_fill(float* __restrict__ ptr, long long count, float v, int lane) {
    const int head = static_cast<int>((16 - (reinterpret_cast<size_t>(ptr) & 15)) & 15) >> 2;
    if (lane < head && lane < count) __stcs(ptr + lane, v);
    if (count <= head) return;
    ptr += head;
    count -= head;
    const long long n4 = count >> 2;
    const float4 v4 = make_float4(v, v, v, v);
    for (long long i = lane; i < n4; i += 32) __stcs(reinterpret_cast<float4*>(ptr) + i, v4);
    const int tail = static_cast<int>(count & 3);
    if (lane < tail) __stcs(ptr + 4 * n4 + lane, v);
}

// load a clip's descriptor, derive its conditioning constants, park everything in the warp's clip-state words
template <bool EXACT, class S, class MS>
__device__ __forceinline__ void enter_clip(const StreamParams& p, const MS& ms, int b, int lane, int* cs) {
    const ClipDesc d = ms.desc(b);
    const ClipCost cc = ms.cost(b);
    const bool mix = d.n2 > 0;
    const int nmin = min(d.n1, d.n2);
    float sub1 = 0.f, mean2 = 0.f, mixmean = 0.f;
    if (d.flags & kFlagStats) {  // whole-clip means from the statistics kernel's partial sums (fixed tree)
        // the statistics kernel is this launch's predecessor on the stream; when this launch was allowed to start ahead
        // of its completion, THIS is where it has to be complete and its sums visible (a no-op otherwise)
        pdl_join();
        const double* ps = p.partials + static_cast<size_t>(b) * kStatSegs * kStatVals;
        double s1 = 0.0, s2 = 0.0, s2c = 0.0;
        if (lane < kStatSegs) {
            s1 = __ldcg(ps + lane * kStatVals);  // from L2: written by a kernel that may have run concurrently with this one
            s2 = __ldcg(ps + lane * kStatVals + 1);
            s2c = __ldcg(ps + lane * kStatVals + 2);
        }
#pragma unroll
        for (int o = kStatSegs / 2; o > 0; o >>= 1) {
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            s2c += __shfl_xor_sync(0xffffffffu, s2c, o);
        }
        s1 = __shfl_sync(0xffffffffu, s1, 0);
        s2 = __shfl_sync(0xffffffffu, s2, 0);
        s2c = __shfl_sync(0xffffffffu, s2c, 0);
        const float mean1 = static_cast<float>(s1 / d.n1);
        if (mix) mean2 = static_cast<float>(s2 / d.n2);
        if (EXACT) {
            sub1 = mean1;
            if (mix) {  // mean over n1 samples of l1*(w1-mean1) + l2*(w2'-mean2), w2' zero-padded / cut to n1
                const double r1 = s1 - static_cast<double>(d.n1) * mean1;
                const double r2 = s2c - static_cast<double>(nmin) * mean2;
                mixmean = static_cast<float>((static_cast<double>(d.l1) * r1 + static_cast<double>(d.l2) * r2) / d.n1);
            }
        }
    }
    // (fused mode: CS_SUB1 is filled at every cold start of the warp inside the clip: robust_level, five loads)
    __syncwarp();
    if (lane == 0) {
        cs_set_q(cs, CS_N1, d.n1, nmin, __float_as_int(mix ? d.l1 : 1.0f), __float_as_int(mix ? d.l2 : 0.0f));
        cs_set_q(cs, CS_SUB1, __float_as_int(sub1), __float_as_int(mean2), __float_as_int(mixmean),
                 (d.flags & (kFlagAligned1 | kFlagAligned2)) | (mix ? kCsMix : 0) | (full_groups(d.n1, nmin, mix, d.flags) << kCsFullShift));
        const unsigned long long a1 = reinterpret_cast<unsigned long long>(static_cast<const S*>(p.wav) + d.off1);
        const unsigned long long a2 = reinterpret_cast<unsigned long long>(static_cast<const S*>(p.wav) + d.off2);
        cs_set_q(cs, CS_OFF1, static_cast<int>(a1), static_cast<int>(a1 >> 32), static_cast<int>(a2), static_cast<int>(a2 >> 32));
        cs_set_q(cs, CS_F0, d.f0, d.f1, d.t0, d.t1);
        cs_set_q(cs, CS_MEFF, d.m_eff, cc.a_lo, cc.a_hi, cc.b_lo);
    }
    __syncwarp();
}

// log floor, SpecAugment fill, normalisation, coalesced streaming stores of the group's 4 rows.
//   FASTLOG   (ln(e) - mean) / div  =  lg2(e) * (ln2 / div) - mean / div : one MUFU.LG2 and one FMA per cell (abs error
//             ~1e-6 in log-mel units); cells on the float-epsilon floor take the host-computed IEEE constant, so they
//             stay bit-identical to the reference.
//   !FASTLOG  logf + IEEE subtraction / division, the reference's expression verbatim.
//   NMEL      compile-time row length (every (slot, lane) then owns a column and row offsets are immediates); 0: p.nmel
__device__ __forceinline__ float lg2_ftz(float x) {  // one MUFU.LG2: the argument is >= eps, never subnormal
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <bool FASTLOG, int NMEL>
__device__ __forceinline__ void epilogue(const StreamParams& p, const int* cs, const int* mel_bin, int lane, int f0, float* __restrict__ out,
                                         const float (&acc)[kMelSlots][kGroupFrames]) {
    const float eps = 1.1920928955078125e-07f;  // kaldi.py:22
    const float fill = p.fill;
    const int nmel = NMEL ? NMEL : p.nmel;
    const int4 mq = cs_q(cs, CS_F0);
    const int mf0 = mq.x, mf1 = mq.y, t0 = mq.z, t1 = mq.w, m_eff = cs_i(cs, CS_MEFF);
    float* col[kMelSlots];  // the lane's four columns at row f0
    bool keep[kMelSlots];   // column survives the frequency mask
    bool own[kMelSlots];
    // FASTLOG: value = lg2(max(e, eps)) * la + lb with the frequency mask folded into the constants of the column
    // (masked: 0 * lg + fill == fill exactly).  The host picks log_bias such that lg2(eps) * la + lb is exactly the
    // reference's value of a cell on the float-epsilon floor, so floored cells need no select either.
    float la[kMelSlots], lb[kMelSlots];
#pragma unroll
    for (int i = 0; i < kMelSlots; ++i) {
        const int mb = mel_bin[32 * i + lane];
        own[i] = NMEL != 0 || mb >= 0;
        keep[i] = !(mb >= mf0 && mb < mf1);
        col[i] = out + static_cast<size_t>(f0) * nmel + max(mb, 0);
        la[i] = keep[i] ? p.log_scale : 0.f;
        lb[i] = keep[i] ? p.log_bias : fill;
    }
    const int rows = min(kGroupFrames, p.T - f0);
    if (FASTLOG && NMEL != 0 && f0 + kGroupFrames <= min(m_eff, p.T) && (f0 + kGroupFrames <= t0 || f0 >= t1)) {
        // all four rows carry features (the common case): no row selects, no row predicates
#pragma unroll
        for (int q = 0; q < kGroupFrames; ++q) {
#pragma unroll
            for (int i = 0; i < kMelSlots; ++i) __stcs(col[i] + q * nmel, fmaf(lg2_ftz(fmaxf(acc[i][q], eps)), la[i], lb[i]));
        }
        return;
    }
    if (FASTLOG) {
        // straight-line: 16 independent cells (their MUFU / FMA latencies overlap), rows that are padding or time-masked by
        // a select, rows past target_length by predicated stores -- no branch fences the scheduler
#pragma unroll
        for (int q = 0; q < kGroupFrames; ++q) {
            const int row = f0 + q;
            const bool rlive = row < m_eff && !(row >= t0 && row < t1);  // warp-uniform
#pragma unroll
            for (int i = 0; i < kMelSlots; ++i) {
                float v = fmaf(lg2_ftz(fmaxf(acc[i][q], eps)), la[i], lb[i]);
                v = rlive ? v : fill;
                if (own[i] && q < rows) __stcs(col[i] + q * nmel, v);
            }
        }
        return;
    }
#pragma unroll
    for (int q = 0; q < kGroupFrames; ++q) {
        if (q < rows) {  // warp-uniform
            const int row = f0 + q;
            const bool rlive = row < m_eff && !(row >= t0 && row < t1);  // warp-uniform
            if (!rlive) {
#pragma unroll
                for (int i = 0; i < kMelSlots; ++i)
                    if (own[i]) __stcs(col[i] + q * nmel, fill);
            } else {
#pragma unroll
                for (int i = 0; i < kMelSlots; ++i) {
                    const float lg = logf(fmaxf(acc[i][q], eps));
                    float v = p.skip_norm ? lg : __fdiv_rn(__fsub_rn(lg, p.norm_mean), p.norm_div);
                    v = keep[i] ? v : fill;
                    if (own[i]) __stcs(col[i] + q * nmel, v);
                }
            }
        }
    }
}

// ---- frequency-major output (the layout the model's first convolution consumes: Models.py:106-107 turns [B, T, F] into
// [B, 1, F, T] with unsqueeze + transpose) -- OUT = 1: fp32 [B, F, T], OUT = 2: bf16 [B, F, T].  target_length is a multiple
// of 4 here (checked on the host), so a group's 4 frames of one mel bin are ONE aligned 16-byte (8-byte) store and no
// row is cut: a lane writes its 4 bins with 4 vector stores instead of 16 scalar ones.  Same arithmetic, same values.
template <bool BF16>
__device__ __forceinline__ void store_frames4(void* __restrict__ clip_out, size_t elem, float v0, float v1, float v2, float v3) {
    if (BF16) {
        const __nv_bfloat162 lo = __float22bfloat162_rn(make_float2(v0, v1)), hi = __float22bfloat162_rn(make_float2(v2, v3));
        uint2 u;
        u.x = *reinterpret_cast<const unsigned*>(&lo);
        u.y = *reinterpret_cast<const unsigned*>(&hi);
        __stcs(reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(clip_out) + elem), u);
    } else {
        __stcs(reinterpret_cast<float4*>(static_cast<float*>(clip_out) + elem), make_float4(v0, v1, v2, v3));
    }
}
// rows [row0, row1) of every mel bin of one clip := v  (row0, row1 multiples of 4)
template <bool BF16>
__device__ __forceinline__ void warp_fill_ft(void* __restrict__ clip_out, int T, int nmel, int row0, int row1, float v, int lane) {
    const int n4 = (row1 - row0) >> 2;
    const int total = nmel * n4;
    for (int i = lane; i < total; i += 32) {
        const int m = i / n4, j = i - m * n4;
        store_frames4<BF16>(clip_out, static_cast<size_t>(m) * T + row0 + 4 * j, v, v, v, v);
    }
}
template <bool FASTLOG, int NMEL, bool BF16>
__device__ __forceinline__ void epilogue_ft(const StreamParams& p, const int* cs, const int* mel_bin, int lane, int f0, void* __restrict__ clip_out,
                                            const float (&acc)[kMelSlots][kGroupFrames]) {
    const float eps = 1.1920928955078125e-07f;  // kaldi.py:22
    const float fill = p.fill;
    const int4 mq = cs_q(cs, CS_F0);
    const int mf0 = mq.x, mf1 = mq.y, t0 = mq.z, t1 = mq.w, m_eff = cs_i(cs, CS_MEFF);
    bool rlive[kGroupFrames];  // warp-uniform
#pragma unroll
    for (int q = 0; q < kGroupFrames; ++q) rlive[q] = f0 + q < m_eff && !(f0 + q >= t0 && f0 + q < t1);
#pragma unroll
    for (int i = 0; i < kMelSlots; ++i) {
        const int mb = mel_bin[32 * i + lane];
        const bool own = NMEL != 0 || mb >= 0;
        const bool keep = !(mb >= mf0 && mb < mf1);
        const float la = keep ? p.log_scale : 0.f, lb = keep ? p.log_bias : fill;
        float v[kGroupFrames];
#pragma unroll
        for (int q = 0; q < kGroupFrames; ++q) {
            if (FASTLOG) {
                v[q] = fmaf(lg2_ftz(fmaxf(acc[i][q], eps)), la, lb);
            } else {
                const float lg = logf(fmaxf(acc[i][q], eps));
                v[q] = p.skip_norm ? lg : __fdiv_rn(__fsub_rn(lg, p.norm_mean), p.norm_div);
                v[q] = keep ? v[q] : fill;
            }
            v[q] = rlive[q] ? v[q] : fill;
        }
        if (own) store_frames4<BF16>(clip_out, static_cast<size_t>(max(mb, 0)) * p.T + f0, v[0], v[1], v[2], v[3]);
    }
}

__device__ __forceinline__ void cs_set(int* cs, int w, int v, int lane) {
    if (lane == 0) *reinterpret_cast<volatile int*>(cs + w) = v;
}

// ------------------------------------------------------------------------------------------------
// Optional per-warp timeline (StreamParams::trace): where the time of a launch goes, warp by warp.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ long long global_ns() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void trace_begin(int* cs, int lane) {
    if (lane == 0) {
        *reinterpret_cast<volatile long long*>(cs + CS_TR_T0) = global_ns();
        *reinterpret_cast<volatile long long*>(cs + CS_TR_C0) = clock64();
        cs[CS_TR_LIVE] = 0;
        cs[CS_TR_DEAD] = 0;
        cs[CS_TR_COLD] = 0;
        cs[CS_TR_STEP] = 0;
    }
    __syncwarp();
}
__device__ __forceinline__ void trace_mark(int* cs, int lane) {  // end of the prologue: C0 := cycles spent so far
    if (lane == 0) {
        const long long c0 = cs_ll(cs, CS_TR_C0);
        // low word: prologue cycles; the start clock stays recoverable as (now - prologue)
        *reinterpret_cast<volatile long long*>(cs + CS_TR_C0) = (c0 & ((1ll << 40) - 1)) | ((clock64() - c0) << 40);
    }
    __syncwarp();
}
__device__ __forceinline__ void trace_count(const StreamParams& p, int* cs, int lane, int word, int n) {
    if (lane == 0) {
        const int warp = threadIdx.x >> 5;
        const int step = cs_i(cs, CS_TR_STEP);
        *reinterpret_cast<volatile int*>(cs + CS_TR_STEP) = step + 1;
        if (step < kTraceSteps) {
            const long long c0 = cs_ll(cs, CS_TR_C0) & ((1ll << 40) - 1);
            const long long kind = word == CS_TR_LIVE ? 1 : (word == CS_TR_COLD ? 2 : 3);
            p.trace[(static_cast<size_t>(blockIdx.x) * kWarpsPerCta + warp) * kTraceWords + 4 + step] =
                ((clock64() - c0) & ((1ll << 40) - 1)) | (kind << 40) | (static_cast<long long>(n) << 44);
        }
        *reinterpret_cast<volatile int*>(cs + word) = cs_i(cs, word) + n;
    }
}
__device__ __forceinline__ void trace_phase(const StreamParams& p, int* cs, int lane, int kind) {
    __syncwarp();
    if (lane == 0) {
        const int step = cs_i(cs, CS_TR_STEP);
        if (step < kTraceSteps) {
            *reinterpret_cast<volatile int*>(cs + CS_TR_STEP) = step + 1;
            const long long c0 = cs_ll(cs, CS_TR_C0) & ((1ll << 40) - 1);
            p.trace[(static_cast<size_t>(blockIdx.x) * kWarpsPerCta + (threadIdx.x >> 5)) * kTraceWords + 4 + step] =
                ((clock64() - c0) & ((1ll << 40) - 1)) | (static_cast<long long>(kind) << 40);
        }
    }
}
__device__ __forceinline__ void trace_end(const StreamParams& p, const int* cs, int lane) {
    __syncwarp();
    if (lane == 0) {
        const int warp = threadIdx.x >> 5;
        const long long packed = cs_ll(cs, CS_TR_C0);
        const long long c0 = packed & ((1ll << 40) - 1), pro = static_cast<unsigned long long>(packed) >> 40;
        const long long total = (clock64() - c0) & ((1ll << 40) - 1);
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        long long* t = p.trace + (static_cast<size_t>(blockIdx.x) * kWarpsPerCta + warp) * kTraceWords;
        t[0] = cs_ll(cs, CS_TR_T0);
        t[1] = global_ns();
        t[2] = pro | (total << 32);
        t[3] = static_cast<long long>(smid) | (static_cast<long long>(cs_i(cs, CS_TR_LIVE) + cs_i(cs, CS_TR_COLD)) << 16) |
               (static_cast<long long>(cs_i(cs, CS_TR_DEAD)) << 32) | (static_cast<long long>(cs_i(cs, CS_TR_COLD)) << 48);
    }
}

constexpr int kStageIters = StageGeom<kOverlapSamples, kGroupSamples>::kIters;  // float4 per lane and waveform of one step

// one live group whose window is staged: frame means, pass 1, pass 2, power, (prefetch), mel, epilogue.
// Everything it needs beyond the kernel parameters comes from the warp's shared-memory state (CS_*).  The prefetch
// leaves the next group's 640 new samples in ra / rc; the caller commits them at the top of its next step.
template <bool MELSPEC, class S, bool TRACE, int OUT>
__device__ __forceinline__ void process_group(const StreamParams& p, const float* TAB, float* WS, int lane, const int b, const int gn,
                                              typename Pcm<S>::reg_t (&ra)[kStageIters], typename Pcm<S>::reg_t (&rc)[kStageIters]) {
    const int par = (gn >> 2) & 1;
    const float* BS = WS + kWsBS + par * kBsFloats;
    int* cs = reinterpret_cast<int*>(WS + kWsCS);
    float* EX = WS + kWsEX;
    const int h = lane >> 4, n2 = lane & 15;

    // per-frame means (kaldi.py:183-186) from the 16-sample block sums: lane 8 f + j adds 3 (4) blocks, xor tree
    float cmu_x, cmu_y;
    {
        const float* bs = BS + 10 * (lane >> 3) + (lane & 7);
        float a = (bs[0] + bs[8]) + bs[16];
        if ((lane & 7) == 0) a += bs[24];
        a += __shfl_xor_sync(0xffffffffu, a, 1);
        a += __shfl_xor_sync(0xffffffffu, a, 2);
        a += __shfl_xor_sync(0xffffffffu, a, 4);
        const float cmu = p.c1 * (a * (1.0f / kFrameLen));
        cmu_x = __shfl_sync(0xffffffffu, cmu, 16 * h);
        cmu_y = __shfl_sync(0xffffffffu, cmu, 16 * h + 8);
    }
    {
        cf z[32];
        pair_pass1(n2, h, par, WS, TAB + kTabWin, cmu_x, cmu_y, z);
        __syncwarp();  // EX aliases the middle piece of the staged samples: every lane has its samples in registers
        pair_pass1_store(n2, TAB + kTabTwz, EX + h * kExSecondPair, z);
    }
    __syncwarp();
    if (TRACE && p.trace) trace_phase(p, cs, lane, 6);
    {
        // pass 2 re-maps the lanes: pair h2 = lane & 1, row t = lane >> 1 (see kExSecondPair)
        const int h2 = lane & 1, t = lane >> 1;
        cf za[16], zb[16];
        pair_pass2_fft(t, EX + h2 * kExSecondPair, TAB + kTabTw, za, zb);
        __syncwarp();  // P aliases EX: all reads done before any write
        pair_power_store(t, h2, za, zb, EX);
    }
    __syncwarp();
    if (TRACE && p.trace) trace_phase(p, cs, lane, 7);
    {
        if (gn & kNextLive) {  // prefetch the next group's 640 new samples; they are consumed after the epilogue
            const MixState<S> m = load_mix<S>(cs, p.wav);
            const int S1 = ((gn >> 2) + 1) * kStepSamples;
            if (gn & kNextFull) stage_issue<kOverlapSamples, kGroupSamples, true>(m, S1, lane, ra, rc);
            else stage_issue<kOverlapSamples, kGroupSamples, false>(m, S1, lane, ra, rc);
            if (TRACE && p.trace) trace_phase(p, cs, lane, 11);
        }
    }
    float acc[kMelSlots][kGroupFrames];
    {
        MelView mt;
        mt.start = reinterpret_cast<const int*>(TAB + kTabMelStart);
        mt.bin = reinterpret_cast<const int*>(TAB + kTabMelBin);
        mt.info = reinterpret_cast<const int*>(TAB + kTabMelInfo);
        mt.w = TAB + kTabMelW;
        warp_mel<MELSPEC>(lane, EX, mt, acc);
    }
    if (TRACE && p.trace) trace_phase(p, cs, lane, 8);
    {
        const int f0 = (gn >> 2) * kGroupFrames;
        const int* mel_bin = reinterpret_cast<const int*>(TAB + kTabMelBin);
        // MELSPEC is only selected for 128 mel bins: every (slot, lane) then owns a column
        if constexpr (OUT == kOutTF) {
            float* out = p.out + static_cast<size_t>(b) * p.T * p.nmel;
            if (p.fast_log) epilogue<true, MELSPEC ? kMaxMel : 0>(p, cs, mel_bin, lane, f0, out, acc);
            else epilogue<false, MELSPEC ? kMaxMel : 0>(p, cs, mel_bin, lane, f0, out, acc);
        } else {
            constexpr bool BF16 = OUT == kOutFTbf16;
            void* out = reinterpret_cast<char*>(p.out) + static_cast<size_t>(b) * p.T * p.nmel * (BF16 ? 2 : 4);
            if (p.fast_log) epilogue_ft<true, MELSPEC ? kMaxMel : 0, BF16>(p, cs, mel_bin, lane, f0, out, acc);
            else epilogue_ft<false, MELSPEC ? kMaxMel : 0, BF16>(p, cs, mel_bin, lane, f0, out, acc);
        }
    }
    if (TRACE && p.trace) trace_phase(p, cs, lane, 9);
}

// ------------------------------------------------------------------------------------------------
// Kernel 2: the streaming kernel.  grid = four CTAs per SM, block = 4 warps, no CTA barrier after the table load.
// Warp r of W owns the groups whose preceding cost lies in [r*C/W, (r+1)*C/W)  (clip-major order).
// ------------------------------------------------------------------------------------------------
// TRACE: the per-warp timeline (StreamParams::trace) is compiled in; the production variants carry none of its checks
// OUT: output layout (kOutTF: fp32 [B, T, F], the reference's; kOutFT / kOutFTbf16: [B, F, T] for the model's first convolution)
template <bool EXACT, bool MELSPEC, bool PMETA, class S, bool TRACE = false, int OUT = kOutTF>
__global__ void __launch_bounds__(kThreads, 4)
    fbank_stream_kernel(const __grid_constant__ StreamParams p, const __grid_constant__ std::conditional_t<PMETA, MetaBlock, MetaNone> meta) {
    using MS = std::conditional_t<PMETA, ParamMeta, GlobalMeta>;
    MS ms;
    if constexpr (PMETA) ms.blk = &meta;
    else {
        ms.desc_ = p.desc;
        ms.cost_ = p.cost;
    }
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* TAB = smem;
    float* WS = smem + p.tab_floats + warp * kWarpFloats;
    int* cs = reinterpret_cast<int*>(WS + kWsCS);
    // The next launch on the stream may begin as soon as SMs free up (only effective when the host launched it with
    // programmatic stream serialization, which it does only when that launch cannot conflict with this one).
    pdl_release();
    // ---- prologue: the table copy and the partition search overlap (both are round trips to L2) ----
    // (Measured in r02 and left as it is -- the other CTAs of the SM hide it: the whole table copied with cp.async while
    // the search runs, 48.25 vs 48.09 us per launch; the two 64-bit divisions below replaced by a host-computed
    // quotient / remainder, 48.3 us.)
    const int n4 = p.tab_floats / 4;
    const float4* tsrc = reinterpret_cast<const float4*>(p.tab);
    float4 tv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid < n4) tv = __ldg(tsrc + tid);
    if (TRACE && p.trace) trace_begin(cs, lane);
    // this warp's share of the work
    const long long W = static_cast<long long>(gridDim.x) * kWarpsPerCta;
    const long long r = static_cast<long long>(warp) * gridDim.x + blockIdx.x;  // neighbouring ranges on different SMs
    const long long c_lo = p.total_cost * r / W, c_hi = p.total_cost * (r + 1) / W;
    const bool has_work = c_lo < c_hi;
    if (has_work) {
        int b, g;
        const long long start_cost = find_start(ms, p.cm, p.n_clips, p.G, c_lo, lane, b, g);
        // groups are taken while the cost consumed is < budget
        if (lane == 0) cs_set_q(cs, CS_B, b, g, static_cast<int>(c_hi - start_cost), 0);
        // pull the first group's window towards the SM while the tables are copied and the CTA barrier is passed: its
        // loads (issued after the barrier) are the one DRAM latency of the launch nothing else hides.  No registers held.
        if (b < p.n_clips) {
            const ClipDesc d = ms.desc(b);
            constexpr int kLine = 128 / static_cast<int>(sizeof(S));
            const int s0 = g * kStepSamples + kLine * lane;
            if (kLine * lane < kGroupSamples + kLine) {
                if (s0 < d.n1) prefetch_l2(static_cast<const S*>(p.wav) + d.off1 + s0);
                if (s0 < min(d.n1, d.n2)) prefetch_l2(static_cast<const S*>(p.wav) + d.off2 + s0);
            }
        }
    }
    {
        float4* dst = reinterpret_cast<float4*>(TAB);
        if (tid < n4) dst[tid] = tv;
        // the rest in batches of 8 loads in flight per thread (one round trip to L2 per batch, not per float4)
        for (int i0 = tid + kThreads; i0 < n4; i0 += 8 * kThreads) {
            float4 t8[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (i0 + j * kThreads < n4) t8[j] = __ldg(tsrc + i0 + j * kThreads);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (i0 + j * kThreads < n4) dst[i0 + j * kThreads] = t8[j];
        }
    }
    __syncthreads();
    if (!has_work) {
        if (TRACE && p.trace) trace_end(p, cs, lane);
        pdl_join();
        return;
    }
    int loaded_b = -1;
    bool staged = false;  // the current group's overlap samples are in the window and its new samples in ra / rc
    typename Pcm<S>::reg_t ra[kStageIters], rc[kStageIters];
    if (TRACE && p.trace) trace_mark(cs, lane);

    // loop state: clip, group and remaining budget stay in registers (three words live across the FFT phases); everything
    // per clip lives in the warp's shared-memory words (CS_*)
    int b, g, budget;
    {
        const int4 ls = cs_q(cs, CS_B);
        b = ls.x;
        g = ls.y;
        budget = ls.z;
    }
    while (true) {
        if (b >= p.n_clips || budget <= 0) break;
        if (b != loaded_b) {
            enter_clip<EXACT, S>(p, ms, b, lane, cs);
            loaded_b = b;
            staged = false;
            if (TRACE && p.trace) trace_phase(p, cs, lane, 4);
        }
        const int4 dq = cs_q(cs, CS_MEFF);  // m_eff, a_lo, a_hi, b_lo
        if (is_dead(dq.y, dq.z, dq.w, g)) {
            // a run of constant rows: take as many of its groups as the budget covers, one contiguous fill
            const int run_end = (g >= dq.w) ? p.G : dq.z;
            const int n = min(run_end - g, (budget + p.cm.dead - 1) / p.cm.dead);
            if (TRACE && p.trace) trace_count(p, cs, lane, CS_TR_DEAD, n);
            const int row0 = kGroupFrames * g, row1 = min(p.T, kGroupFrames * (g + n));
            if constexpr (OUT == kOutTF) {
                warp_fill(p.out + (static_cast<size_t>(b) * p.T + row0) * p.nmel, static_cast<long long>(row1 - row0) * p.nmel, p.fill, lane);
            } else {
                constexpr bool BF16 = OUT == kOutFTbf16;
                warp_fill_ft<BF16>(reinterpret_cast<char*>(p.out) + static_cast<size_t>(b) * p.T * p.nmel * (BF16 ? 2 : 4), p.T, p.nmel, row0, row1,
                                   p.fill, lane);
            }
            const bool wrap = g + n == p.G;
            budget -= n * p.cm.dead;
            b = wrap ? b + 1 : b;
            g = wrap ? 0 : g + n;
            staged = false;
            continue;
        }
        if (TRACE && p.trace) trace_count(p, cs, lane, staged ? CS_TR_LIVE : CS_TR_COLD, 1);
        bool next_live;
        int gn;
        {
            MixState<S> m = load_mix<S>(cs, p.wav);
            const int S0 = g * kStepSamples;
            const bool full = window_full(m, g);
            const int par = g & 1;
            float carry;
            if (!staged) {  // cold start of a run of groups: the whole window is loaded here, nothing was prefetched
                constexpr int NH = StageGeom<0, kOverlapSamples>::kIters;
                typename Pcm<S>::reg_t ha[NH], hc[NH];
                float a0 = 0.f, c0 = 0.f;
                if (!EXACT) {  // the constant removed from the clip (see mix_sample): a robust level of each waveform
                    // (costs +0.6 us per 100-clip launch over reading the first sample; measured and rejected in r02:
                    // prefetching the ten samples into L2 in the prologue, 50.7 vs 49.9 us, and loading them there behind the
                    // table copy, 50.4 us)
                    a0 = robust_level<S>(m.w1, m.n1);
                    if (m.mix) c0 = robust_level<S>(m.w2, m.nmin);
                }
                if (full) {
                    stage_issue<0, kOverlapSamples, true>(m, S0, lane, ha, hc);
                    stage_issue<kOverlapSamples, kGroupSamples, true>(m, S0, lane, ra, rc);
                } else {
                    stage_issue<0, kOverlapSamples, false>(m, S0, lane, ha, hc);
                    stage_issue<kOverlapSamples, kGroupSamples, false>(m, S0, lane, ra, rc);
                }
                if (TRACE && p.trace) trace_phase(p, cs, lane, 10);
                if (!EXACT) {
                    m.sub1 = m.mix ? fmaf(m.l1, a0, m.l2 * c0) : a0;
                    cs_set(cs, CS_SUB1, __float_as_int(m.sub1), lane);
                }
                carry = 0.f;
                if (full) stage_commit<EXACT, 0, kOverlapSamples, true>(m, S0, lane, p.preemph, p.c1, ha, hc, WS, par, carry);
                else stage_commit<EXACT, 0, kOverlapSamples, false>(m, S0, lane, p.preemph, p.c1, ha, hc, WS, par, carry);
            } else {
                carry = cs_f(cs, CS_CARRY);
            }
            if (full) stage_commit<EXACT, kOverlapSamples, kGroupSamples, true>(m, S0, lane, p.preemph, p.c1, ra, rc, WS, par, carry);
            else stage_commit<EXACT, kOverlapSamples, kGroupSamples, false>(m, S0, lane, p.preemph, p.c1, ra, rc, WS, par, carry);
            if (TRACE && p.trace) trace_phase(p, cs, lane, 5);
            budget -= p.cm.live + (run_start(dq.y, dq.z, g) ? p.cm.cold : 0);
            // (g is live: it lies in [0, a_lo) or in [a_hi, b_lo); the run it belongs to ends at a_lo or at b_lo <= G)
            next_live = g + 1 < (g < dq.y ? dq.y : dq.w) && budget > 0;
            const bool full_next = next_live && window_full(m, g + 1);
            gn = (g << 2) | (next_live ? kNextLive : 0) | (full_next ? kNextFull : 0);
            if (lane == 0) cs[CS_CARRY] = __float_as_int(carry);
        }
        __syncwarp();
        process_group<MELSPEC, S, TRACE, OUT>(p, TAB, WS, lane, b, gn, ra, rc);
        staged = next_live;
        {  // advance
            const bool wrap = g + 1 == p.G;
            b = wrap ? b + 1 : b;
            g = wrap ? 0 : g + 1;
        }
        __syncwarp();
    }
    if (TRACE && p.trace) trace_end(p, cs, lane);
    pdl_join();
}

}  // namespace psla
