// Per-lane phases of the fused Kaldi-fbank tile kernel (sm_100a), also compilable by g++ for the
// CPU lane-stepping harness in tests/emu/ (see fft_core.cuh).
//
// Geometry (fixed: 16 kHz, 25 ms / 10 ms  =>  400-sample frames every 160 samples, 512-point FFT;
// torchaudio/compliance/kaldi.py:138-140 as called at src/dataloaders/audioset_dataset.py:119-120).
//
// Work unit: a FRAME PAIR (two consecutive frames x, y of one clip) is transformed as ONE 512-point
// complex FFT of z = x + i*y by a HALF-WARP (16 lanes, 32 complex values = 64 registers per lane):
//
//   pass 1   lane n2 (0..15) holds z[16*n1 + n2], n1 = 0..31 (only n1 < 25 is non-zero: 400 = 25*16)
//            -> 32-point DIF FFT over n1 in registers -> times W_512^(n2*k1) -> shared memory EX[k1][n2]
//   pass 2   lane t reads rows k1 = t and k1 = 32 - t (lane 0: rows 0 and 16), two 16-point DIF FFTs over
//            n2 in registers -> Z[k1 + 32*k2].  With this row assignment Z[k] and Z[512-k] are always in
//            the SAME lane, so the two real spectra separate without any exchange:
//              X[k] = Z[k] + conj(Z[512-k]),  Y[k] = -i * (Z[k] - conj(Z[512-k]))      (the 1/2 is folded
//            into the window table), and the power spectra |X|^2, |Y|^2 go to shared memory P[frame][bin].
//   mel      the whole warp (4 frames): lane l owns mel bins l, l+32, l+64, l+96; sparse triangle weights
//            as fp32 FMAs, log with the float-epsilon floor, SpecAugment fill, (v - mean) / div, store.
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

constexpr int kWarpsPerCta = 8;
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kFramesPerWarp = 4;
constexpr int kFramesPerCta = kWarpsPerCta * kFramesPerWarp;                    // 32
constexpr int kTileSamples = (kFramesPerCta - 1) * kFrameShift + kFrameLen;     // 5360
constexpr int kBlk16PerTile = kTileSamples / 16;                                // 335

// Staged-sample layout: 16 floats of padding after every 320 samples so that the two half-warps of a
// warp (whose frame pairs start 320 samples apart, i.e. on the same bank) read disjoint bank halves.
PSLA_HDC int dpad(int u) { return u + 16 * (u / 320); }
constexpr int kStageFloats = ((kTileSamples - 1) + 16 * ((kTileSamples - 1) / 320) + 1 + 31) / 32 * 32;  // 5632

// Exchange buffer of one frame pair: 32 rows (k1) x 16 complex (n2), row stride 18 complex (=36 floats)
// so that pass-2 row reads (LDS.128, one row per lane) are bank-conflict free.
constexpr int kExRowFloats = 36;
constexpr int kExPairFloats = 32 * kExRowFloats;  // 1152 floats = 4608 B
// Power-spectrum rows (alias the warp's two exchange buffers once pass 2 has consumed them):
// row stride 264 == 8 (mod 16) so the two half-warps' stores land on disjoint bank halves.
constexpr int kPRowFloats = 264;
static_assert(4 * kPRowFloats <= 2 * kExPairFloats, "P rows must fit in the warp's exchange buffers");

// Sparse mel table in shared memory.  Lane l, slot i owns mel bin j = l + 32*i:
//   start[j], count[j] : first FFT bin with non-zero weight and number of consecutive non-zero weights
//   w[i][r][l]         : weight of FFT bin start[j]+r for mel bin j  (r < slot_max[i]; zero padded)
constexpr int kMelSlots = 4;
constexpr int kMelMaxCount = 32;  // per-slot r-loop bound limit (16 kHz / 128 bins needs 10)

struct MelTableView {
    const short* start;      // [128]
    const short* count;      // [128]
    const int* slot_max;     // [4]
    const int* slot_off;     // [4]  offset (in floats) of slot i inside w
    const float* w;          // packed [sum_i slot_max[i]][32]
};

// ---------------------------------------------------------------------------------------------
// Pass 1: load the two frames of this half-warp's pair, per-frame DC removal + pre-emphasis (applied
// as d[j] - (1-p)*mu with d[j] = x[j] - p*x[j-1] staged once per sample) + Hann window (x0.5),
// 32-point FFT, inter-pass twiddle, store to the exchange buffer.
//   D       staged d[] in dpad() layout,  XS[lf] = (1-p) * x[first sample of local frame lf]
//   lfx     local index (within the CTA tile) of the x frame (even); the y frame is lfx + 1
//   mu_*    per-frame means pre-multiplied by (1-p)
// ---------------------------------------------------------------------------------------------
PSLA_HD void pair_pass1(int n2, int lfx, const float* __restrict__ D, const float* __restrict__ XS,
                        const float* __restrict__ WIN, const float* __restrict__ TW, float* __restrict__ EX,
                        float cmu_x, float cmu_y) {
    float re[32], im[32];
    const int half = lfx >> 1;                         // (160*lfx) / 320
    const float* dx = D + 160 * lfx + 16 * half + n2;  // x frame, element 16*n1 + n2 at dx[16*n1 + 16*(n1>=20)]
    const float* dy = dx + 160;                        // y frame,                     dy[16*n1 + 16*(n1>=10)]
    // replicate-padded first sample (kaldi.py:193-198): lane 0 reads (1-p)*x0 from XS instead of d[0];
    // done as an address select so that lane 0 does not get a code path of its own
    const float* d0x = (n2 == 0) ? XS + lfx : dx;
    const float* d0y = (n2 == 0) ? XS + lfx + 1 : dy;
    static_for<0, 25>([&](auto n1_) {
        constexpr int n1 = decltype(n1_)::value;
        const float vx = (n1 == 0) ? d0x[0] : dx[16 * n1 + (n1 >= 20 ? 16 : 0)];
        const float vy = (n1 == 0) ? d0y[0] : dy[16 * n1 + (n1 >= 10 ? 16 : 0)];
        const float w = WIN[16 * n1 + n2];
        re[n1] = (vx - cmu_x) * w;
        im[n1] = (vy - cmu_y) * w;
    });
    static_for<25, 32>([&](auto n1_) {
        re[decltype(n1_)::value] = 0.f;
        im[decltype(n1_)::value] = 0.f;
    });
    fft_dif<32>(re, im);
    static_for<0, 32>([&](auto r_) {
        constexpr int r = decltype(r_)::value;
        constexpr int k1 = brev<5>(r);
        float ar = re[r], ai = im[r];
        if constexpr (k1 != 0) {
            float wr, wi;
            ld2(TW + 2 * (k1 * 16 + n2), wr, wi);
            const float t = ar * wr - ai * wi;
            ai = ar * wi + ai * wr;
            ar = t;
        }
        st2(EX + k1 * kExRowFloats + 2 * n2, ar, ai);
    });
}

// One 16-point FFT over n2 of exchange row k1:  out (re, im)[r] = Z[k1 + 32 * brev4(r)]
PSLA_HD void row_fft16(const float* __restrict__ EX, int k1, float (&re)[16], float (&im)[16]) {
    const float* row = EX + k1 * kExRowFloats;
    static_for<0, 8>([&](auto q_) {
        constexpr int n = 2 * decltype(q_)::value;
        ld4(row + 2 * n, re[n], im[n], re[n + 1], im[n + 1]);
    });
    fft_dif<16>(re, im);
}

// ---------------------------------------------------------------------------------------------
// Pass 2a: the two 16-point row FFTs of lane t (rows t and 32-t; lane 0: rows 0 and 16).
// Pass 2b: separation of the two real spectra + power, Px[bin] / Py[bin] for this lane's bins.
// They are separate functions because P aliases the exchange buffer: every lane of the warp must have
// finished 2a (its reads) before any lane starts 2b (__syncwarp on the device, loop order on the host).
// ---------------------------------------------------------------------------------------------
PSLA_HD void pair_pass2_fft(int t, const float* __restrict__ EX, float (&ar)[16], float (&ai)[16], float (&br)[16],
                            float (&bi)[16]) {
    const int ka = (t != 0) ? t : 0;
    const int kb = (t != 0) ? 32 - t : 16;
    row_fft16(EX, ka, ar, ai);  // Z[ka + 32*k2]
    row_fft16(EX, kb, br, bi);  // Z[kb + 32*k2]
}

PSLA_HD void pair_power_store(int t, const float (&ar)[16], const float (&ai)[16], const float (&br)[16],
                              const float (&bi)[16], float* __restrict__ Px, float* __restrict__ Py) {
    // Lanes 1..15: a[r] = Z[t + 32*k2], b[r] = Z[(32-t) + 32*k2] with k2 = brev4(r), and
    //   512 - (t + 32*k2) = (32-t) + 32*(15-k2)   =>   the conjugate partner of a[r] is b[15 - r].
    // Lane 0 holds rows 0 and 16, whose partners sit in the SAME row (row 0: k2 <-> (16-k2)%16, row 16:
    // k2 <-> 15-k2).  Instead of a divergent code path, lane 0 permutes its registers with selects so that the
    // generic (a'[r], b'[15-r]) pairing produces bins 0,32,..,224 (slots k2 < 8) and 16,48,..,240 (slots k2 >= 8);
    // bin 256 is the one extra value only lane 0 computes.
    const bool t0 = (t == 0);
    float pr[16], pi[16], qr[16], qi[16];
    static_for<0, 16>([&](auto r_) {
        constexpr int r = decltype(r_)::value;
        constexpr int k2 = brev<4>(r);
        if constexpr (k2 < 8) {
            pr[r] = ar[r];
            pi[r] = ai[r];
            constexpr int s = brev<4>(8 + k2);
            qr[r] = t0 ? br[s] : br[r];
            qi[r] = t0 ? bi[s] : bi[r];
        } else {
            constexpr int sa = brev<4>(k2 - 8), sb = brev<4>((k2 + 1) % 16);
            pr[r] = t0 ? br[sa] : ar[r];
            pi[r] = t0 ? bi[sa] : ai[r];
            qr[r] = t0 ? ar[sb] : br[r];
            qi[r] = t0 ? ai[sb] : bi[r];
        }
    });
    static_for<0, 16>([&](auto r_) {
        constexpr int r = decltype(r_)::value;
        constexpr int k2 = brev<4>(r);
        constexpr int rp = 15 - r;  // brev4(15 - k2)
        const float xr = pr[r] + qr[rp], xi = pi[r] - qi[rp];
        const float yr = pi[r] + qi[rp], yi = qr[rp] - pr[r];
        int bin;
        if constexpr (k2 < 8) bin = t + 32 * k2;
        else bin = t0 ? (16 + 32 * (k2 - 8)) : ((32 - t) + 32 * (15 - k2));
        Px[bin] = xr * xr + xi * xi;
        Py[bin] = yr * yr + yi * yi;
    });
    if (t0) {  // Nyquist bin: Z[256] pairs with itself
        constexpr int r8 = brev<4>(8);
        const float xr = ar[r8] + ar[r8], yr = ai[r8] + ai[r8];
        Px[256] = xr * xr;
        Py[256] = yr * yr;
    }
}

// ---------------------------------------------------------------------------------------------
// Mel projection for the 4 frames of a warp: acc[i][f] = sum_r w[i][r][lane] * P[f][start + r]
// ---------------------------------------------------------------------------------------------
PSLA_HD void warp_mel(int lane, const float* __restrict__ P, const MelTableView& mt, float (&acc)[kMelSlots][kFramesPerWarp]) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < kMelSlots; ++i) {
        const int j = lane + 32 * i;
        const int s = mt.start[j];
        const float* w = mt.w + mt.slot_off[i] + lane;
        const int rmax = mt.slot_max[i];
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        // weights are zero padded up to rmax, so lanes with fewer taps just add 0 * P[some valid bin]
        for (int r = 0; r < rmax; ++r) {
            const float wv = w[32 * r];
            const int k = (s + r < kNumBins - 1) ? s + r : kNumBins - 1;
            const float* p = P + k;
            a0 = fmaf(wv, p[0], a0);
            a1 = fmaf(wv, p[kPRowFloats], a1);
            a2 = fmaf(wv, p[2 * kPRowFloats], a2);
            a3 = fmaf(wv, p[3 * kPRowFloats], a3);
        }
        acc[i][0] = a0;
        acc[i][1] = a1;
        acc[i][2] = a2;
        acc[i][3] = a3;
    }
}

}  // namespace psla
