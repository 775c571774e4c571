// Per-lane phases of the warp-autonomous streaming fbank kernel (sm_100a), also compilable by g++ for the CPU
// lane-stepping harness in tests/emu/ (see fft_core.cuh).
//
// Geometry (fixed: 16 kHz, 25 ms / 10 ms => 400-sample frames every 160 samples, 512-point FFT;
// torchaudio/compliance/kaldi.py:138-140 as called at src/dataloaders/audioset_dataset.py:119-120).
//
// Work unit of a WARP: a GROUP of 4 consecutive frames of one clip = 880 staged samples.  A warp walks along a
// clip group by group; the 240 samples two consecutive groups share stay where they are in its private shared-memory
// window (two overlap buffers used alternately), only 640 new samples are staged per step.  No CTA-level barrier exists after the table load.
//
// Inside a group each HALF-WARP turns a frame pair (x, y) into ONE 512-point complex FFT of z = x + i*y
// (32 complex values = 64 registers per lane):
//   pass 1   lane n2 (0..15) holds z[16*n1 + n2], n1 = 0..31 (only n1 < 25 is non-zero: 400 = 25*16).  The y frame
//            starts 160 samples = 10 n1-steps after the x frame, so ONE run of 35 staged values feeds both.
//            -> 32-point DIT FFT over n1 in registers (FMA-only butterflies, first stage pruned for the 7 zero inputs) -> shared memory
//            EX[k1][n2], untwiddled (row 0 pre-multiplied by W_32^-n2, see below).  EX aliases the middle of the staged samples (kWsEX).
//   pass 2   lane t reads rows ka = t and kb = 32 - t (lane 0: rows 0 and 16; pass 1 stored them side by side), multiplies row ka by
//            TW[t][n2] = W_512^(t*n2) and row kb by conj(TW[t][n2]): because W_512^((32-t)*n2) = W_16^n2 * conj(TW[t][n2])
//            and a modulation by W_16^n2 only rotates the 16-point FFT's output index by one, both rows share one
//            table row.  Two 16-point DIT FFTs over n2 in registers.  Z[k] and Z[512-k] are then in the SAME lane, so
//            the two real spectra separate without any exchange (the 1/2 is folded into the window table); the power
//            spectra go to shared memory as P[bin][frame 0..3].
//            Lane 0's table row is W_32^n2 (what row 16 needs after conj + rotation); pass 1 compensates row 0.
//   mel      the whole warp: lane l owns 4 mel bins (one per slot); per tap ONE 128-bit load brings the bin's power of
//            all 4 frames; sparse triangle weights as fp32 FMAs.
#pragma once

#include "fft_core.cuh"

namespace psla {

// 8- and 16-byte shared-memory accesses (plain scalar code on the host harness)
PSLA_HD void st2(float* p, float a, float b) {
#if defined(__CUDA_ARCH__)
    *reinterpret_cast<float2*>(p) = make_float2(a, b);
#else
    p[0] = a;
    p[1] = b;
#endif
}
PSLA_HD void st4(float* p, float a, float b, float c, float d) {
#if defined(__CUDA_ARCH__)
    *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
#else
    p[0] = a;
    p[1] = b;
    p[2] = c;
    p[3] = d;
#endif
}
PSLA_HD void ld2(const float* p, float& a, float& b) {
#if defined(__CUDA_ARCH__)
    const float2 v = *reinterpret_cast<const float2*>(p);
    a = v.x;
    b = v.y;
#else
    a = p[0];
    b = p[1];
#endif
}
PSLA_HD void ld4(const float* p, float& a, float& b, float& c, float& d) {
#if defined(__CUDA_ARCH__)
    const float4 v = *reinterpret_cast<const float4*>(p);
    a = v.x;
    b = v.y;
    c = v.z;
    d = v.w;
#else
    a = p[0];
    b = p[1];
    c = p[2];
    d = p[3];
#endif
}

constexpr int kFrameLen = 400;
constexpr int kFrameShift = 160;
constexpr int kNfft = 512;
constexpr int kNumBins = kNfft / 2 + 1;  // 257
constexpr int kMaxMel = 128;

constexpr int kGroupFrames = 4;                                                  // frames per warp step
constexpr int kGroupSamples = (kGroupFrames - 1) * kFrameShift + kFrameLen;      // 880
constexpr int kStepSamples = kGroupFrames * kFrameShift;                         // 640 new samples per step
constexpr int kOverlapSamples = kGroupSamples - kStepSamples;                    // 240

// Staged-sample layout of a warp's window.  A group's 880 samples live in three pieces, so that the 240 samples two
// consecutive groups share never have to be moved:
//   OV[par]      window samples [0, 240)    -- written by the PREVIOUS group's staging (or a cold start)
//   DM           window samples [240, 640)  -- with 16 floats of padding at sample 320
//   OV[par ^ 1]  window samples [640, 880)  -- the next group's [0, 240)
// par = group index & 1.  Bank rule (pass 1 reads, per instruction, sample u in half-warp 0 and u + 320 in half-warp 1):
// both OV buffers start on bank 0, DM on bank 16, and the padding at 320 keeps the two half-warps on disjoint bank halves.
constexpr int kOvFloats = 256;  // >= kOverlapSamples, multiple of 32
// physical float offset inside the warp window of window sample u (see kWsOV / kWsDM below)
PSLA_HDC int wpos(int u, int par) {
    return u < 240 ? par * kOvFloats + u : (u < 640 ? 2 * kOvFloats + 16 + (u - 240) + (u >= 320 ? 16 : 0) : (par ^ 1) * kOvFloats + (u - 640));
}
constexpr int kBsFloats = 56;  // 55 sums of 16 samples (one such array per parity: sums 40..54 are also the next group's 0..14)
constexpr int kXsFloats = 64;  // (1-p) * x at u = 0, 160, 320, 480 (per parity; u = 640, 800 are the next group's 0, 160), at xs_pos(u / 160)
// Frames 0, 1 sit on bank 0 and frames 2, 3 on bank 16: the banks no other lane of the half-warp reads its staged sample
// from in the same pass-1 load (lane 0 reads XS where lanes 1..15 read consecutive staged samples).
PSLA_HDC int xs_pos(int f) { return (f & 1) * 32 + (f >> 1) * 16; }
// Exchange buffer of one frame pair, organised by the ROW PAIRS pass 2 consumes: EX[t][n2] = (Z1[ka][n2], Z1[kb][n2]) with
// ka = t, kb = 32 - t (t = 0: rows 0 and 16) -- 16 bytes per (t, n2), so pass 1 stores and pass 2 loads are all 128-bit.
// Row stride 68 floats: the 8 lanes of a quarter-warp (4 consecutive t x 2 pairs) read 8 distinct 16-byte bank groups.
// (The compiler spends ~45 MOVs per group on bringing the four registers of each 128-bit store together.  Storing the
// two rows in separate arrays with 64-bit stores -- no MOVs, same wavefronts, conflict-free -- was measured in r02:
// 50.1 instead of 48.2 us per launch; twice the store instructions through the memory-instruction queue cost more.)
constexpr int kExRowFloats = 68;
constexpr int kExPairFloats = 16 * kExRowFloats;  // 1088 floats = 4352 B
constexpr int kTwRowFloats = 36;                  // row stride of the inter-pass twiddle table TW[t][n2]
// Pass 2 interleaves the two pairs over the lanes (lane = 2 t + h, so that a lane pair stores the 4 frames of a bin as
// one contiguous 16 bytes); the second pair's buffer starts 16 floats later so that (t, 0) and (t, 1) hit different banks.
constexpr int kExSecondPair = kExPairFloats + 16;
constexpr int kExFloats = kExSecondPair + kExPairFloats;  // 2192
// Power spectra of the warp's 4 frames, P[bin][frame], alias the exchange buffers once pass 2 has read them.
constexpr int kPFloats = 4 * (kNfft / 2);  // bins 0..255 (bin 256 never carries mel weight: kaldi.py:627 pads a zero column)
static_assert(kPFloats <= kExFloats, "P must fit in the warp's exchange buffers");

// per-warp shared-memory window (floats).  The exchange buffers ALIAS the middle piece DM of the staged samples: once
// every lane of the warp holds its pass-1 inputs in registers DM is dead until the next step commits its new samples --
// by then the power spectra (which alias EX) have been consumed by the mel stage.  The two overlap buffers sit below
// EX and survive.  11.2 KB per warp: four CTAs of 4 warps (each with its own copy of the tables) fit one SM, so a
// following launch takes over SM resources in small steps.
constexpr int kWsOV = 0;                      // OV[0], OV[1]
constexpr int kWsEX = 2 * kOvFloats;          // 512: multiple of 32
constexpr int kWsDM = kWsEX + 16;             // bank 16 (see above); inside EX
constexpr int kDmFloats = 400 + 16;
constexpr int kWsBS = kWsEX + kExFloats;      // BS[0], BS[1]
constexpr int kWsXS = kWsBS + 2 * kBsFloats;  // XS[0], XS[1]
constexpr int kWsCS = kWsXS + 2 * kXsFloats;  // per-clip state of the warp (kept out of the register file, see stream_kernel.cuh)
constexpr int kCsWords = 36;
constexpr int kWarpFloats = kWsCS + kCsWords;  // 2868 floats = 11472 B
static_assert(kOvFloats >= kOverlapSamples && kWsEX % 32 == 0 && kWsDM % 32 == 16 && kDmFloats + 16 <= kExFloats, "EX aliases DM only");
static_assert(wpos(0, 0) == kWsOV && wpos(240, 0) == kWsDM && wpos(240, 1) == kWsDM && wpos(640, 0) == kOvFloats && wpos(640, 1) == 0, "window pieces");
static_assert((kWsEX * 4) % 16 == 0 && (kWsCS * 4) % 16 == 0 && (kWarpFloats * 4) % 16 == 0, "16-byte alignment of vector regions");

// Constant tables (floats), copied once per CTA into shared memory.
constexpr int kWinRow = 28;                            // WINT[n2][n1], n1 < 25, zero padded: 0.5 * window[16*n1 + n2]
constexpr int kTabWin = 0;                             // [16][28]
constexpr int kTabTw = kTabWin + 16 * kWinRow;         // [16][36]: TW[t][n2] (re, im), row stride 36 floats
constexpr int kTabTwz = kTabTw + 16 * kTwRowFloats;    // [16] (re, im): W_32^(-n2), applied to row 0 in pass 1
constexpr int kTabMelStart = kTabTwz + 32;             // [4][32] ints: first P bin of (slot, lane)
constexpr int kTabMelBin = kTabMelStart + 128;         // [4][32] ints: mel bin owned by (slot, lane), -1 = none
constexpr int kTabMelInfo = kTabMelBin + 128;          // [8] ints: slot_max[4], slot_off[4] (rows of W before the slot)
constexpr int kTabMelW = kTabMelInfo + 8;              // [sum slot_max][32] weights
static_assert(kTabTw % 4 == 0 && kTabTwz % 4 == 0 && kTabMelW % 4 == 0, "table regions are float4 aligned");

constexpr int kMelSlots = 4;
constexpr int kMelMaxCount = 128;  // per-row tap bound of the generic path (the table must also fit in shared memory)
// tap counts the specialised (fully unrolled) mel path is compiled for: 16 kHz / 128 bins needs (2, 3, 6, 10)
constexpr int kMelSpec0 = 2, kMelSpec1 = 3, kMelSpec2 = 6, kMelSpec3 = 10;
// (A lane-major weight table read with six 128-bit loads instead of 21 scalar ones was measured in r02: 49.7 instead of
// 48.0 us per launch -- 24 weight registers at once in front of the accumulation chains.)

// ---------------------------------------------------------------------------------------------
// Pass 1 of half-warp h, lane n2: frames 2h (x) and 2h+1 (y) of the group.
//   WS    the warp window: staged d[u] = x[u] - p*x[u-1] at wpos(u, par),  XS[par][f] = (1-p) * x[160 f]
//   cmu_* per-frame means pre-multiplied by (1-p): per-frame DC removal (kaldi.py:183-186) and pre-emphasis
//         (:193-198) commute into  d[j] - (1-p)*mu  (first sample: (1-p)*(x0 - mu), replicate padding)
// ---------------------------------------------------------------------------------------------
// pair_pass1_load fills z[0..24] (the windowed, mean-removed samples); pair_pass1 = load + the pruned 32-point FFT.
// (Calling the load with a literal parity from two branches -- so that the piece offsets fold into immediates -- was
// measured in r02: 58.0 instead of 49.9 us per launch; the merge of the two branches breaks the register pairing of the
// two-wide arithmetic that follows.  The parity stays a run-time value.)
PSLA_HD void pair_pass1_load(int n2, int h, int par, const float* __restrict__ WS, const float* __restrict__ WINT, float cmu_x, float cmu_y,
                             cf (&z)[32]) {
    {
        // element n1 of the lane's run is window sample u = 320 h + 16 n1 + n2; x frame: n1 = 0..24, y frame: n1 = 10..34.
        // The run crosses the window pieces at fixed n1 (15: u = 240, 20: u = 320 / 640), so three base pointers with
        // the piece offsets folded in make every load an immediate offset: v[n1] = p(n1)[16 n1].
        const float* ovp = WS + kWsOV + par * kOvFloats;
        const float* ovo = WS + kWsOV + (par ^ 1) * kOvFloats;
        const float* dm = WS + kWsDM;
        const float* pa = (h == 0 ? ovp : dm + 96) + n2;        // n1 < 15   h = 0: u < 240   h = 1: DM (u >= 320: + 16 padding)
        const float* pb = (h == 0 ? dm - 240 : dm + 96) + n2;   // n1 < 20   h = 0: 240 <= u < 320
        const float* pc = (h == 0 ? dm - 224 : ovo - 320) + n2;  // n1 >= 20  h = 0: u >= 320 (+ 16)   h = 1: u >= 640
        const float* XS = WS + kWsXS + par * kXsFloats;
        // replicate-padded first samples: lane 0 reads (1-p)*x0 from XS instead of d[0]; an address select, so that
        // lane 0 does not get a code path of its own
        const float* x0p = (n2 == 0) ? XS + xs_pos(2 * h) : pa;
        const float* y0p = (n2 == 0) ? XS + xs_pos(2 * h + 1) : pa + 160;
        float w[kWinRow];
        static_for<0, kWinRow / 4>([&](auto q_) {
            constexpr int q = decltype(q_)::value;
            ld4(WINT + kWinRow * n2 + 4 * q, w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        });
        float v[35];
        static_for<0, 35>([&](auto n1_) {
            constexpr int n1 = decltype(n1_)::value;
            v[n1] = (n1 == 0) ? x0p[0] : (n1 < 15 ? pa : (n1 < 20 ? pb : pc))[16 * n1];
        });
        const float vy0 = y0p[0];
        static_for<0, 25>([&](auto n1_) {
            constexpr int n1 = decltype(n1_)::value;
            z[n1].x = (v[n1] - cmu_x) * w[n1];
            z[n1].y = ((n1 == 0 ? vy0 : v[n1 + 10]) - cmu_y) * w[n1];
        });
    }
}
PSLA_HD void pair_pass1(int n2, int h, int par, const float* __restrict__ WS, const float* __restrict__ WINT, float cmu_x, float cmu_y,
                        cf (&z)[32]) {
    pair_pass1_load(n2, h, par, WS, WINT, cmu_x, cmu_y, z);
    fft_dit<32, 5, 25>(z);  // inputs n1 >= 25 are zero (400 = 25 * 16)
}
// The stores of pass 1, separate because EX aliases DM: every lane of the warp must have read its samples before any
// lane stores (__syncwarp on the device, loop order on the host).
PSLA_HD void pair_pass1_store(int n2, const float* __restrict__ TWZ, float* __restrict__ EX, const cf (&z)[32]) {
    float zr, zi;
    ld2(TWZ + 2 * n2, zr, zi);
    static_for<0, 16>([&](auto t_) {
        constexpr int t = decltype(t_)::value;
        constexpr int ka = t, kb = (t != 0) ? 32 - t : 16;
        cf a = z[brev<5>(ka)];
        const cf b = z[brev<5>(kb)];
        if constexpr (t == 0) a = cmul(a, zr, zi);  // row 0 is read by lane 0 of pass 2, whose table row is W_32^n2: compensate
        st4(EX + t * kExRowFloats + 4 * n2, a.x, a.y, b.x, b.y);
    });
}

// ---------------------------------------------------------------------------------------------
// Pass 2a of lane t: rows ka / kb times TW[t] / conj(TW[t]), two 16-point DIT FFTs.
//   a[brev4(k2)]            = Z[ka + 32*k2]
//   b[brev4((k2 + 1) % 16)] = Z[kb + 32*k2]      (the W_16^n2 modulation rotates the output index)
// ---------------------------------------------------------------------------------------------
PSLA_HD void pair_pass2_fft(int t, const float* __restrict__ EX, const float* __restrict__ TW, cf (&a)[16], cf (&b)[16]) {
    const float* row = EX + t * kExRowFloats;  // (Z1[t][n], Z1[32 - t][n]) for n = 0..15  (t = 0: rows 0 and 16)
    const float* tw = TW + t * kTwRowFloats;
    static_for<0, 16>([&](auto n_) {
        constexpr int n = decltype(n_)::value;
        ld4(row + 4 * n, a[n].x, a[n].y, b[n].x, b[n].y);
    });
    static_for<0, 8>([&](auto q_) {
        constexpr int n = 2 * decltype(q_)::value;
        float w0r, w0i, w1r, w1i;
        ld4(tw + 2 * n, w0r, w0i, w1r, w1i);
        if constexpr (n != 0) {  // TW[t][0] == 1
            a[n] = cmul(a[n], w0r, w0i);
            b[n] = cmul_conj(b[n], w0r, w0i);
        }
        a[n + 1] = cmul(a[n + 1], w1r, w1i);
        b[n + 1] = cmul_conj(b[n + 1], w1r, w1i);
    });
    fft_dit<16, 4>(a);
    fft_dit<16, 4>(b);
}

// ---------------------------------------------------------------------------------------------
// Pass 2b: separation of the two real spectra + power, stored as P[bin][2h], P[bin][2h + 1].
//   X[k] = Z[k] + conj(Z[512-k]),  Y[k] = -i * (Z[k] - conj(Z[512-k]))   (1/2 folded into the window)
// Lanes 1..15, slot k2:  Z[t + 32*k2] = A[k2] pairs with Z[(32-t) + 32*(15-k2)] = B[(16-k2) % 16]  (A, B: natural-order
// views of a, b).  Slots k2 < 8 give bin t + 32*k2; slots k2 >= 8 give the mirror bin (32-t) + 32*(15-k2) (|X[512-k]| =
// |X[k]| for real input).  Lane 0 holds rows 0 and 16, whose partners sit in the SAME row: row 0 pairs A[k2] with
// A[(16-k2) % 16] (bins 32*k2), row 16 pairs B[(k2+1) % 16] with B[16-k2] (bins 16 + 32*(15-k2)).  Both differ from
// the generic pairing in ONE operand with the same index, so lane 0 costs 16 complex selects, no code path of its own.
// Separate from 2a because P aliases the exchange buffers: every lane of the warp must have finished its pass-2 reads
// before any lane stores (__syncwarp on the device, loop order on the host).
// ---------------------------------------------------------------------------------------------
PSLA_HD void pair_power_store(int t, int h, const cf (&a)[16], const cf (&b)[16], float* __restrict__ P) {
    const bool t0 = (t == 0);
    float* plo = P + 4 * t + 2 * h;                          // bin t + 32*k2
    float* phi = P + 4 * (t0 ? 16 : 32 - t) + 2 * h;         // bin hibase + 32*(15-k2)
    static_for<0, 16>([&](auto k2_) {
        constexpr int k2 = decltype(k2_)::value;
        constexpr int ia = brev<4>(k2);              // A[k2]
        constexpr int ib = brev<4>((16 - k2) % 16);  // B[(16-k2) % 16]
        cf u, v;
        if constexpr (k2 < 8) {
            u = a[ia];
            constexpr int ia2 = brev<4>((16 - k2) % 16);  // lane 0: A[(16-k2) % 16]
            v.x = t0 ? a[ia2].x : b[ib].x;
            v.y = t0 ? a[ia2].y : b[ib].y;
        } else {
            constexpr int ib2 = brev<4>((k2 + 1) % 16);  // lane 0: B[(k2+1) % 16]
            u.x = t0 ? b[ib2].x : a[ia].x;
            u.y = t0 ? b[ib2].y : a[ia].y;
            v = b[ib];
        }
        const cf x = cadd(u, conj_(v));  // X = u + conj(v)
        const cf y = csub(u, conj_(v));  // i*Y = u - conj(v)
        const float px = fmaf(x.x, x.x, x.y * x.y);
        const float py = fmaf(y.x, y.x, y.y * y.y);
        if constexpr (k2 < 8) st2(plo + 128 * k2, px, py);
        else st2(phi + 128 * (15 - k2), px, py);
    });
}

// ---------------------------------------------------------------------------------------------
// Mel projection for the 4 frames of a warp: acc[i][f] = sum_r W[off_i + r][lane] * P[start_i + r][f]
// R0..R3 >= 0: compile-time tap counts (fully unrolled, weights zero padded); R0 < 0: run-time counts from `info`.
// ---------------------------------------------------------------------------------------------
struct MelView {
    const int* start;  // [4][32]
    const int* bin;    // [4][32]
    const int* info;   // slot_max[4], slot_off[4]
    const float* w;    // [rows][32]
};

template <int R, int OFF>
PSLA_HD void mel_slot_static(int lane, const float* __restrict__ P, const float* __restrict__ w, int start, float (&acc)[kGroupFrames]) {
    cf a01 = mk(0.f, 0.f), a23 = mk(0.f, 0.f);  // frames (0, 1) and (2, 3): one two-wide FMA each per tap
    const float* p = P + 4 * start;
    static_for<0, R>([&](auto r_) {
        constexpr int r = decltype(r_)::value;
        const float wv = w[32 * (OFF + r) + lane];
        cf p01, p23;
        ld4(p + 4 * r, p01.x, p01.y, p23.x, p23.y);
        a01 = cfma(p01, bc(wv), a01);
        a23 = cfma(p23, bc(wv), a23);
    });
    acc[0] = a01.x;
    acc[1] = a01.y;
    acc[2] = a23.x;
    acc[3] = a23.y;
}

template <bool SPEC>
PSLA_HD void warp_mel(int lane, const float* __restrict__ P, const MelView& mt, float (&acc)[kMelSlots][kGroupFrames]) {
    if constexpr (SPEC) {
        mel_slot_static<kMelSpec0, 0>(lane, P, mt.w, mt.start[lane], acc[0]);
        mel_slot_static<kMelSpec1, kMelSpec0>(lane, P, mt.w, mt.start[32 + lane], acc[1]);
        mel_slot_static<kMelSpec2, kMelSpec0 + kMelSpec1>(lane, P, mt.w, mt.start[64 + lane], acc[2]);
        mel_slot_static<kMelSpec3, kMelSpec0 + kMelSpec1 + kMelSpec2>(lane, P, mt.w, mt.start[96 + lane], acc[3]);
    } else {
        for (int i = 0; i < kMelSlots; ++i) {
            const float* p = P + 4 * mt.start[32 * i + lane];
            const float* w = mt.w + 32 * mt.info[4 + i] + lane;
            const int rmax = mt.info[i];
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            for (int r = 0; r < rmax; ++r) {
                const float wv = w[32 * r];
                float p0, p1, p2, p3;
                ld4(p + 4 * r, p0, p1, p2, p3);
                a0 = fmaf(wv, p0, a0);
                a1 = fmaf(wv, p1, a1);
                a2 = fmaf(wv, p2, a2);
                a3 = fmaf(wv, p3, a3);
            }
            acc[i][0] = a0;
            acc[i][1] = a1;
            acc[i][2] = a2;
            acc[i][3] = a3;
        }
    }
}

}  // namespace psla
