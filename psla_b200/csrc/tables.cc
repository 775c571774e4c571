#include "tables.h"

#include <cmath>
#include <cstring>

namespace psla {

void build_window_half(float* w, int n) {
    // ATen hann_window == hamming_window(alpha=.5, beta=.5):
    //   arange(n).mul_(2*pi/(n-1)).cos_().mul_(-0.5).add_(0.5)   all in fp32
    const float step = static_cast<float>(3.14159265358979323846 * 2.0 / static_cast<double>(n - 1));
    for (int j = 0; j < n; ++j) {
        const float a = static_cast<float>(j) * step;
        const float c = cosf(a);
        const float h = c * -0.5f + 0.5f;
        w[j] = 0.5f * h;
    }
}

void build_mel_dense(float* mel, int num_bins, int nfft, float sample_rate, float low_freq, float high_freq) {
    const int num_fft_bins = nfft / 2;
    const double nyquist = 0.5 * sample_rate;
    double high = high_freq;
    if (high <= 0.0) high += nyquist;
    const double fft_bin_width = static_cast<double>(sample_rate) / nfft;
    const double mel_low = 1127.0 * std::log(1.0 + low_freq / 700.0);  // mel_scale_scalar, kaldi.py:326
    const double mel_high = 1127.0 * std::log(1.0 + high / 700.0);
    const double delta = (mel_high - mel_low) / (num_bins + 1);
    const float mel_low_f = static_cast<float>(mel_low), delta_f = static_cast<float>(delta);
    const float width_f = static_cast<float>(fft_bin_width);
    std::vector<float> melk(num_fft_bins);
    for (int k = 0; k < num_fft_bins; ++k) {  // mel_scale(), kaldi.py:330: 1127 * log(1 + f/700) in fp32
        const float f = width_f * static_cast<float>(k);
        const float t = 1.0f + f / 700.0f;
        melk[k] = 1127.0f * logf(t);
    }
    for (int b = 0; b < num_bins; ++b) {
        const float left = mel_low_f + static_cast<float>(b) * delta_f;
        const float center = mel_low_f + (static_cast<float>(b) + 1.0f) * delta_f;
        const float right = mel_low_f + (static_cast<float>(b) + 2.0f) * delta_f;
        for (int k = 0; k < num_fft_bins; ++k) {
            const float up = (melk[k] - left) / (center - left);
            const float down = (right - melk[k]) / (right - center);
            const float m = up < down ? up : down;
            mel[b * num_fft_bins + k] = m > 0.f ? m : 0.f;
        }
    }
}

bool build_table_buffer(const float* window, const float* mel_dense, int num_bins, std::vector<float>* buf, bool* mel_spec) {
    const int nb = kNfft / 2;
    if (num_bins > kMaxMel || num_bins < 1) return false;
    // ---- sparse structure of the mel matrix ----
    int start[kMaxMel] = {0}, count[kMaxMel] = {0};
    for (int j = 0; j < num_bins; ++j) {
        int first = -1, last = -1;
        for (int k = 0; k < nb; ++k)
            if (mel_dense[j * nb + k] != 0.f) {
                if (first < 0) first = k;
                last = k;
            }
        if (first < 0) continue;  // empty row (row 3 at 16 kHz / 128 bins): log floor only
        for (int k = first; k <= last; ++k)
            if (mel_dense[j * nb + k] == 0.f) return false;
        if (last - first + 1 > kMelMaxCount) return false;
        start[j] = first;
        count[j] = last - first + 1;
    }
    int slot_max[kMelSlots], slot_off[kMelSlots], rows = 0;
    for (int i = 0; i < kMelSlots; ++i) {
        int mx = 0;
        for (int l = 0; l < 32; ++l) {
            const int j = l + 32 * i;
            if (j < num_bins && count[j] > mx) mx = count[j];
        }
        slot_max[i] = mx;
    }
    const int spec[kMelSlots] = {kMelSpec0, kMelSpec1, kMelSpec2, kMelSpec3};
    bool fits = num_bins == kMaxMel;  // the unrolled path also assumes that every (slot, lane) owns a mel bin
    for (int i = 0; i < kMelSlots; ++i) fits = fits && slot_max[i] <= spec[i];
    if (fits)  // the unrolled path always runs its compile-time tap counts (weights are zero padded)
        for (int i = 0; i < kMelSlots; ++i) slot_max[i] = spec[i];
    for (int i = 0; i < kMelSlots; ++i) {
        slot_off[i] = rows;
        rows += slot_max[i];
    }
    *mel_spec = fits;
    buf->assign(static_cast<size_t>(kTabMelW) + static_cast<size_t>(rows) * 32 + 3, 0.f);
    buf->resize((buf->size() / 4) * 4);
    float* b = buf->data();
    // ---- window, transposed for 128-bit row reads ----
    for (int n2 = 0; n2 < 16; ++n2)
        for (int n1 = 0; n1 < 25; ++n1) b[kTabWin + kWinRow * n2 + n1] = 0.5f * window[16 * n1 + n2];
    // ---- twiddles (double -> float) ----
    const double two_pi = 2.0 * 3.14159265358979323846;
    for (int t = 0; t < 16; ++t)
        for (int n2 = 0; n2 < 16; ++n2) {
            const double a = (t == 0) ? -two_pi * n2 / 32.0 : -two_pi * static_cast<double>((t * n2) % 512) / 512.0;
            b[kTabTw + kTwRowFloats * t + 2 * n2] = static_cast<float>(std::cos(a));
            b[kTabTw + kTwRowFloats * t + 2 * n2 + 1] = static_cast<float>(std::sin(a));
        }
    for (int n2 = 0; n2 < 16; ++n2) {
        const double a = two_pi * n2 / 32.0;
        b[kTabTwz + 2 * n2] = static_cast<float>(std::cos(a));
        b[kTabTwz + 2 * n2 + 1] = static_cast<float>(std::sin(a));
    }
    // ---- mel: slot i owns mel bins 32 i .. 32 i + 31, one per lane ----
    // A tap is one 128-bit shared-memory read of P[bin][4 frames]; the 8 lanes of a quarter-warp are served together
    // and collide when their bins differ but are congruent mod 8.  Two freedoms remove the collisions: WHICH lane of
    // the slot owns a mel bin (the output row is one 128-byte line either way), and WHERE a row's taps start (rows
    // narrower than the slot's tap count can begin earlier, on zero weights).  Rows pick a start residue mod 8 such
    // that every residue is used at most 4 times (bipartite matching, capacity 4), then each quarter-warp takes one
    // row of every residue.
    int* istart = reinterpret_cast<int*>(b + kTabMelStart);
    int* ibin = reinterpret_cast<int*>(b + kTabMelBin);
    int* info = reinterpret_cast<int*>(b + kTabMelInfo);
    for (int i = 0; i < kMelSlots; ++i) {
        info[i] = slot_max[i];
        info[4 + i] = slot_off[i];
        const int R = slot_max[i];
        int nrows = 0, rows_j[32], lo_start[32], hi_start[32];  // allowed starts of a row: [lo_start, hi_start]
        for (int l = 0; l < 32; ++l) {
            const int j = l + 32 * i;
            if (j >= num_bins) break;
            rows_j[nrows] = j;
            if (count[j] == 0) {  // empty row: all weights zero, any start inside the spectrum
                lo_start[nrows] = 0;
                hi_start[nrows] = R > 0 ? (nb - R < 7 ? nb - R : 7) : 0;
            } else {
                int hi = start[j], lo = start[j] - (R - count[j]);
                if (hi + R > nb) hi = nb - R;  // keep every tap inside the bins the kernel writes (0 .. nfft/2 - 1)
                if (lo < 0) lo = 0;
                lo_start[nrows] = lo;
                hi_start[nrows] = hi;
            }
            ++nrows;
        }
        // residue -> rows matching (capacity 4 per residue)
        int owner[8][4], used[8] = {0}, res_of[32];
        for (int r = 0; r < 8; ++r)
            for (int c = 0; c < 4; ++c) owner[r][c] = -1;
        for (int n = 0; n < nrows; ++n) res_of[n] = -1;
        auto allowed = [&](int n, int r) {  // some start in [lo, hi] congruent to r mod 8?
            const int first = lo_start[n] + ((r - lo_start[n]) % 8 + 8) % 8;
            return first <= hi_start[n];
        };
        bool seen[8];
        auto try_assign = [&](auto&& self, int n) -> bool {
            for (int r = 0; r < 8; ++r) {
                if (seen[r] || !allowed(n, r)) continue;
                seen[r] = true;
                if (used[r] < 4) {
                    owner[r][used[r]++] = n;
                    res_of[n] = r;
                    return true;
                }
                for (int c = 0; c < 4; ++c) {
                    const int other = owner[r][c];
                    if (self(self, other)) {
                        owner[r][c] = n;
                        res_of[n] = r;
                        return true;
                    }
                }
            }
            return false;
        };
        bool perfect = true;
        for (int n = 0; n < nrows; ++n) {
            for (bool& v : seen) v = false;
            if (!try_assign(try_assign, n)) perfect = false;
        }
        // lanes: quarter-warp q takes the q-th row of every residue; rows left without a residue fill the gaps
        int lane_row[32];
        for (int l = 0; l < 32; ++l) lane_row[l] = -1;
        if (perfect) {
            int fill_q[4] = {0, 0, 0, 0};
            for (int r = 0; r < 8; ++r)
                for (int c = 0; c < used[r]; ++c) lane_row[8 * c + fill_q[c]++] = owner[r][c];
        } else {  // could not avoid every collision: keep the natural order and the latest starts
            for (int n = 0; n < nrows; ++n) {
                lane_row[n] = n;
                res_of[n] = -1;
            }
        }
        for (int l = 0; l < 32; ++l) {
            const int n = lane_row[l];
            ibin[32 * i + l] = n >= 0 ? rows_j[n] : -1;
            int s0 = 0;
            if (n >= 0) {
                const int j = rows_j[n];
                s0 = hi_start[n];
                if (res_of[n] >= 0) s0 -= ((hi_start[n] - res_of[n]) % 8 + 8) % 8;  // latest allowed start with that residue
                const int shift = count[j] > 0 ? start[j] - s0 : 0;
                for (int r = 0; r < count[j]; ++r)
                    b[kTabMelW + 32 * (slot_off[i] + shift + r) + l] = mel_dense[j * nb + start[j] + r];
            }
            istart[32 * i + l] = s0;
        }
    }
    return true;
}

}  // namespace psla
