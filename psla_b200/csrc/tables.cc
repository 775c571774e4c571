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

void build_twiddles(float* tw) {
    for (int k1 = 0; k1 < 32; ++k1)
        for (int n2 = 0; n2 < 16; ++n2) {
            const double a = -2.0 * 3.14159265358979323846 * static_cast<double>((k1 * n2) % 512) / 512.0;
            tw[2 * (k1 * 16 + n2)] = static_cast<float>(std::cos(a));
            tw[2 * (k1 * 16 + n2) + 1] = static_cast<float>(std::sin(a));
        }
}

bool pack_mel(const float* dense, int num_bins, int nfft, PackedMel* out) {
    const int nb = nfft / 2;
    if (num_bins > kMaxMel || num_bins < 1) return false;
    std::memset(out->start, 0, sizeof(out->start));
    std::memset(out->count, 0, sizeof(out->count));
    for (int j = 0; j < num_bins; ++j) {
        int first = -1, last = -1;
        for (int k = 0; k < nb; ++k)
            if (dense[j * nb + k] != 0.f) {
                if (first < 0) first = k;
                last = k;
            }
        if (first < 0) continue;  // empty row (row 3 at 16 kHz / 128 bins): log floor only
        for (int k = first; k <= last; ++k)
            if (dense[j * nb + k] == 0.f) return false;
        if (last - first + 1 > kMelMaxCount) return false;
        out->start[j] = static_cast<short>(first);
        out->count[j] = static_cast<short>(last - first + 1);
    }
    int off = 0;
    for (int i = 0; i < kMelSlots; ++i) {
        int mx = 0;
        for (int l = 0; l < 32; ++l) {
            const int j = l + 32 * i;
            if (j < num_bins && out->count[j] > mx) mx = out->count[j];
        }
        out->slot_max[i] = mx;
        out->slot_off[i] = off;
        off += mx * 32;
    }
    out->w.assign(off, 0.f);
    for (int i = 0; i < kMelSlots; ++i)
        for (int l = 0; l < 32; ++l) {
            const int j = l + 32 * i;
            if (j >= num_bins) continue;
            for (int r = 0; r < out->count[j]; ++r)
                out->w[out->slot_off[i] + 32 * r + l] = dense[j * nb + out->start[j] + r];
        }
    return true;
}

}  // namespace psla
