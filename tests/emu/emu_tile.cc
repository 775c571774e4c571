// CPU lane-stepping harness for the sm_100a tile kernel (TEST ONLY, built by tests/test_emu.py with g++).
//
// It runs the very same per-lane functions the CUDA kernel calls (psla_b200/csrc/fbank_tile.cuh),
// one simulated lane after another with the __syncwarp points turned into loop boundaries, on a
// staged tile built by plain scalar code.  It checks index maps / twiddles / the two-pass FFT /
// the separation / the sparse mel layout without a GPU.  It is not a fallback: nothing in
// psla_b200 can load it.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../psla_b200/csrc/fbank_tile.cuh"
#include "../../psla_b200/csrc/tables.h"

using namespace psla;

extern "C" {

// x: mean-removed (and mixed) waveform, n samples.  out: [m][num_bins] raw log-mel (floor applied).
// mel_dense: optional [num_bins][256] override (nullptr = native table).  window_half likewise.
int emu_fbank(const float* x, int n, int num_bins, const float* mel_dense, const float* window_half, float* out) {
    if (n < kFrameLen) return -1;
    const int m = 1 + (n - kFrameLen) / kFrameShift;
    std::vector<float> win(kFrameLen), tw(32 * 16 * 2), dense(num_bins * 256);
    if (window_half) std::memcpy(win.data(), window_half, sizeof(float) * kFrameLen);
    else build_window_half(win.data(), kFrameLen);
    if (mel_dense) std::memcpy(dense.data(), mel_dense, sizeof(float) * num_bins * 256);
    else build_mel_dense(dense.data(), num_bins, kNfft, 16000.f, 20.f, 0.f);
    build_twiddles(tw.data());
    PackedMel pm;
    if (!pack_mel(dense.data(), num_bins, kNfft, &pm)) return -2;
    MelTableView mt{pm.start, pm.count, pm.slot_max, pm.slot_off, pm.w.data()};

    const float p = 0.97f;
    const float c1 = static_cast<float>(1.0 - static_cast<double>(p));
    const float eps = 1.1920928955078125e-07f;
    std::vector<float> D(kStageFloats), XS(kFramesPerCta + 2), BS(kBlk16PerTile + 1);
    std::vector<float> EXw(2 * kExPairFloats);

    for (int F0 = 0; F0 < m; F0 += kFramesPerCta) {
        const int S0 = F0 * kFrameShift;
        // ---- staging (scalar restatement of the kernel's cooperative load) ----
        std::fill(D.begin(), D.end(), 0.f);
        for (int u = 0; u < kTileSamples; ++u) {
            const int s = S0 + u;
            const float xc = s < n ? x[s] : 0.f;
            const float xp = (u > 0 && s - 1 < n) ? x[s - 1] : 0.f;
            D[dpad(u)] = fmaf(-p, xp, xc);
            if (u % kFrameShift == 0 && u / kFrameShift < kFramesPerCta) XS[u / kFrameShift] = c1 * xc;
        }
        for (int q = 0; q < kBlk16PerTile; ++q) {
            // same association as the kernel: 4 x (sum of 4) combined by two xor-shuffle steps
            float s4[4];
            for (int g = 0; g < 4; ++g) {
                const int s = S0 + 16 * q + 4 * g;
                float v[4];
                for (int e = 0; e < 4; ++e) v[e] = (s + e < n) ? x[s + e] : 0.f;
                s4[g] = (v[0] + v[1]) + (v[2] + v[3]);
            }
            BS[q] = (s4[0] + s4[1]) + (s4[2] + s4[3]);
        }
        for (int w = 0; w < kWarpsPerCta; ++w) {
            float cmu[4];
            for (int q = 0; q < 4; ++q) {
                const int lf = 4 * w + q;
                float sum = 0.f;
                for (int k = 0; k < 25; ++k) sum += BS[10 * lf + k];
                cmu[q] = c1 * (sum * (1.0f / kFrameLen));
            }
            float* P = EXw.data();
            float regs[32][4][16];
            for (int lane = 0; lane < 32; ++lane) {  // pass 1
                const int h = lane >> 4, n2 = lane & 15;
                pair_pass1(n2, 4 * w + 2 * h, D.data(), XS.data(), win.data(), tw.data(), EXw.data() + h * kExPairFloats,
                           cmu[2 * h], cmu[2 * h + 1]);
            }
            for (int lane = 0; lane < 32; ++lane) {  // pass 2a (reads)   -- __syncwarp --
                const int h = lane >> 4, t = lane & 15;
                pair_pass2_fft(t, EXw.data() + h * kExPairFloats, regs[lane][0], regs[lane][1], regs[lane][2], regs[lane][3]);
            }
            for (int lane = 0; lane < 32; ++lane) {  // pass 2b (P aliases EX) -- __syncwarp --
                const int h = lane >> 4, t = lane & 15;
                pair_power_store(t, regs[lane][0], regs[lane][1], regs[lane][2], regs[lane][3], P + (2 * h) * kPRowFloats,
                                 P + (2 * h + 1) * kPRowFloats);
            }
            for (int lane = 0; lane < 32; ++lane) {  // mel                    -- __syncwarp --
                float acc[kMelSlots][kFramesPerWarp];
                warp_mel(lane, P, mt, acc);
                for (int i = 0; i < kMelSlots; ++i)
                    for (int f = 0; f < kFramesPerWarp; ++f) {
                        const int row = F0 + 4 * w + f, j = lane + 32 * i;
                        if (row < m && j < num_bins) out[row * num_bins + j] = logf(fmaxf(acc[i][f], eps));
                    }
            }
        }
    }
    return m;
}

// expose the native tables for table-parity tests
void emu_tables(int num_bins, float* window_half, float* mel_dense, float* tw) {
    build_window_half(window_half, kFrameLen);
    build_mel_dense(mel_dense, num_bins, kNfft, 16000.f, 20.f, 0.f);
    build_twiddles(tw);
}

// raw 512-point complex FFT through pass1/pass2 machinery is covered by emu_fbank; this one checks
// the register FFT templates alone against a direct DFT (N = 16 and 32).
void emu_fft(int n, const float* in_re, const float* in_im, float* out_re, float* out_im) {
    if (n == 32) {
        float re[32], im[32];
        for (int i = 0; i < 32; ++i) { re[i] = in_re[i]; im[i] = in_im[i]; }
        fft_dif<32>(re, im);
        for (int r = 0; r < 32; ++r) { out_re[brev<5>(r)] = re[r]; out_im[brev<5>(r)] = im[r]; }
    } else {
        float re[16], im[16];
        for (int i = 0; i < 16; ++i) { re[i] = in_re[i]; im[i] = in_im[i]; }
        fft_dif<16>(re, im);
        for (int r = 0; r < 16; ++r) { out_re[brev<4>(r)] = re[r]; out_im[brev<4>(r)] = im[r]; }
    }
}
}
