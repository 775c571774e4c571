// CPU lane-stepping harness for the sm_100a streaming kernel (TEST ONLY, built by tests/test_emu.py with g++).
//
// It runs the very same per-lane functions the CUDA kernel calls (psla_b200/csrc/fbank_stream.cuh), one simulated
// lane after another with the __syncwarp points turned into loop boundaries, on a staged group window built by plain
// scalar code.  It checks index maps / twiddles / the two-pass FFT / the separation / the sparse mel layout / the
// table buffer without a GPU.  It is not a fallback: nothing in psla_b200 can load it.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../psla_b200/csrc/fbank_stream.cuh"
#include "../../psla_b200/csrc/tables.h"

using namespace psla;

extern "C" {

// Lane order of one phase.  seed == 0: ascending.  Otherwise a seeded shuffle, different for every phase: between two
// __syncwarp points the hardware gives no order between the lanes of a warp, so the result must not depend on it -- any
// lane that read what another lane writes inside the same phase (a shared-memory race: the exchange buffer aliases the
// staged samples, the power spectra alias the exchange buffer) would show up as a difference between two seeds.
static void lane_order(unsigned seed, unsigned phase, int (&ord)[32]) {
    for (int i = 0; i < 32; ++i) ord[i] = i;
    if (seed == 0) return;
    unsigned long long s = 0x9E3779B97F4A7C15ull * (seed + 1) + 0xD1B54A32D192ED03ull * (phase + 1);
    for (int i = 31; i > 0; --i) {
        s = s * 6364136223846793005ull + 1442695040888963407ull;
        const int j = static_cast<int>((s >> 33) % static_cast<unsigned>(i + 1));
        const int t = ord[i];
        ord[i] = ord[j];
        ord[j] = t;
    }
}

// x: mean-removed (and mixed) waveform, n samples.  out: [m][num_bins] raw log-mel (floor applied).
// mel_dense: optional [num_bins][256] override (nullptr = native table); window: optional [400] analysis window.
// force_generic_mel != 0 runs the run-time-loop mel path even when the unrolled one applies.
int emu_fbank_ordered(const float* x, int n, int num_bins, const float* mel_dense, const float* window, int force_generic_mel,
                      unsigned order_seed, float* out);
int emu_fbank(const float* x, int n, int num_bins, const float* mel_dense, const float* window, int force_generic_mel,
              float* out) {
    return emu_fbank_ordered(x, n, num_bins, mel_dense, window, force_generic_mel, 0u, out);
}
int emu_fbank_ordered(const float* x, int n, int num_bins, const float* mel_dense, const float* window, int force_generic_mel,
                      unsigned order_seed, float* out) {
    if (n < kFrameLen) return -1;
    const int m = 1 + (n - kFrameLen) / kFrameShift;
    std::vector<float> win(kFrameLen), dense(num_bins * 256);
    if (window) std::memcpy(win.data(), window, sizeof(float) * kFrameLen);
    else {
        build_window_half(win.data(), kFrameLen);
        for (auto& v : win) v *= 2.f;
    }
    if (mel_dense) std::memcpy(dense.data(), mel_dense, sizeof(float) * num_bins * 256);
    else build_mel_dense(dense.data(), num_bins, kNfft, 16000.f, 20.f, 0.f);
    std::vector<float> tab;
    bool spec = false;
    if (!build_table_buffer(win.data(), dense.data(), num_bins, &tab, &spec)) return -2;
    if (force_generic_mel) spec = false;
    const float* TAB = tab.data();
    MelView mt{reinterpret_cast<const int*>(TAB + kTabMelStart), reinterpret_cast<const int*>(TAB + kTabMelBin),
               reinterpret_cast<const int*>(TAB + kTabMelInfo), TAB + kTabMelW};

    const float p = 0.97f;
    const float c1 = static_cast<float>(1.0 - static_cast<double>(p));
    const float eps = 1.1920928955078125e-07f;
    std::vector<float> W(kWarpFloats, 0.f);
    float* EX = W.data() + kWsEX;

    for (int f0 = 0; f0 < m; f0 += kGroupFrames) {
        const int S0 = f0 * kFrameShift;
        const int par = (f0 / kGroupFrames) & 1;  // the kernel alternates its two overlap buffers from group to group
        float* BS = W.data() + kWsBS + par * kBsFloats;
        float* XS = W.data() + kWsXS + par * kXsFloats;
        // ---- staging (scalar restatement of the kernel's warp-private staging) ----
        for (int u = 0; u < kGroupSamples; ++u) {
            const int s = S0 + u;
            const float xc = s < n ? x[s] : 0.f;
            const float xp = (u > 0 && s - 1 < n) ? x[s - 1] : 0.f;
            W[wpos(u, par)] = fmaf(-p, xp, xc);
            if (u % kFrameShift == 0 && u < kStepSamples) XS[xs_pos(u / kFrameShift)] = c1 * xc;
        }
        for (int q = 0; q < kGroupSamples / 16; ++q) {
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
        // ---- per-frame means: lane 8 f + j adds blocks 10 f + j (+8, +16, and +24 for j == 0), xor tree over j ----
        float cmu[4];
        for (int f = 0; f < 4; ++f) {
            float part[8];
            for (int j = 0; j < 8; ++j) {
                float a = (BS[10 * f + j] + BS[10 * f + j + 8]) + BS[10 * f + j + 16];
                if (j == 0) a += BS[10 * f + 24];
                part[j] = a;
            }
            for (int o = 1; o < 8; o <<= 1) {
                float nxt[8];
                for (int j = 0; j < 8; ++j) nxt[j] = part[j] + part[j ^ o];
                std::memcpy(part, nxt, sizeof(part));
            }
            cmu[f] = c1 * (part[0] * (1.0f / kFrameLen));
        }
        cf regs[32][2][16];
        static cf z1[32][32];
        int ord[32];
        const unsigned ph = 8u * static_cast<unsigned>(f0 / kGroupFrames);
        lane_order(order_seed, ph + 0, ord);
        for (int li = 0, lane = ord[0]; li < 32; lane = ord[++li < 32 ? li : 0]) {  // pass 1 (reads the staged samples)
            const int h = lane >> 4, n2 = lane & 15;
            pair_pass1(n2, h, par, W.data(), TAB + kTabWin, cmu[2 * h], cmu[2 * h + 1], z1[lane]);
        }
        lane_order(order_seed, ph + 1, ord);
        for (int li = 0, lane = ord[0]; li < 32; lane = ord[++li < 32 ? li : 0]) {  // pass 1 stores (EX aliases the middle of the window)   -- __syncwarp --
            const int h = lane >> 4, n2 = lane & 15;
            pair_pass1_store(n2, TAB + kTabTwz, EX + h * kExSecondPair, z1[lane]);
        }
        lane_order(order_seed, ph + 2, ord);
        for (int li = 0, lane = ord[0]; li < 32; lane = ord[++li < 32 ? li : 0]) {  // pass 2a (reads)   -- __syncwarp --
            const int h = lane & 1, t = lane >> 1;
            pair_pass2_fft(t, EX + h * kExSecondPair, TAB + kTabTw, regs[lane][0], regs[lane][1]);
        }
        lane_order(order_seed, ph + 3, ord);
        for (int li = 0, lane = ord[0]; li < 32; lane = ord[++li < 32 ? li : 0]) {  // pass 2b (P aliases EX) -- __syncwarp --
            const int h = lane & 1, t = lane >> 1;
            pair_power_store(t, h, regs[lane][0], regs[lane][1], EX);
        }
        lane_order(order_seed, ph + 4, ord);
        for (int li = 0, lane = ord[0]; li < 32; lane = ord[++li < 32 ? li : 0]) {  // mel                    -- __syncwarp --
            float acc[kMelSlots][kGroupFrames];
            if (spec) warp_mel<true>(lane, EX, mt, acc);
            else warp_mel<false>(lane, EX, mt, acc);
            for (int i = 0; i < kMelSlots; ++i)
                for (int f = 0; f < kGroupFrames; ++f) {
                    const int row = f0 + f, j = mt.bin[32 * i + lane];
                    if (row < m && j >= 0) out[row * num_bins + j] = logf(fmaxf(acc[i][f], eps));
                }
        }
    }
    return m;
}

// expose the native tables for table-parity tests: window [400] (not halved), dense mel, and the flat table buffer
int emu_tables(int num_bins, float* window, float* mel_dense, float* tab, int tab_cap, int* mel_spec) {
    build_window_half(window, kFrameLen);
    for (int j = 0; j < kFrameLen; ++j) window[j] *= 2.f;
    build_mel_dense(mel_dense, num_bins, kNfft, 16000.f, 20.f, 0.f);
    std::vector<float> buf;
    bool spec = false;
    if (!build_table_buffer(window, mel_dense, num_bins, &buf, &spec)) return -1;
    *mel_spec = spec ? 1 : 0;
    if (static_cast<int>(buf.size()) > tab_cap) return -2;
    std::memcpy(tab, buf.data(), buf.size() * sizeof(float));
    return static_cast<int>(buf.size());
}

// table offsets (floats) so the test can slice the buffer: win, tw, twz, mel_start, mel_bin, mel_info, mel_w, row strides
void emu_layout(int* v) {
    v[0] = kTabWin; v[1] = kTabTw; v[2] = kTabTwz; v[3] = kTabMelStart; v[4] = kTabMelBin; v[5] = kTabMelInfo;
    v[6] = kTabMelW; v[7] = kWinRow; v[8] = kTwRowFloats;
}

// the register FFT templates alone against a direct DFT (N = 16, 32) and the pruned 32-point variant (n = -32:
// inputs 25..31 are ignored / treated as zero)
void emu_fft(int n, const float* in_re, const float* in_im, float* out_re, float* out_im) {
    if (n == 32 || n == -32) {
        cf z[32];
        for (int i = 0; i < 32; ++i) z[i] = mk(in_re[i], in_im[i]);
        if (n == 32) fft_dit<32, 5>(z);
        else {
            for (int i = 25; i < 32; ++i) z[i] = mk(0.f, 0.f);
            fft_dit<32, 5, 25>(z);
        }
        for (int r = 0; r < 32; ++r) { out_re[brev<5>(r)] = z[r].x; out_im[brev<5>(r)] = z[r].y; }
    } else {
        cf z[16];
        for (int i = 0; i < 16; ++i) z[i] = mk(in_re[i], in_im[i]);
        fft_dit<16, 4>(z);
        for (int r = 0; r < 16; ++r) { out_re[brev<4>(r)] = z[r].x; out_im[brev<4>(r)] = z[r].y; }
    }
}
}
