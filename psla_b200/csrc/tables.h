// Host-side constant tables of the fbank frontend (pure C++, no CUDA): Hann window, HTK-mel
// triangle weights in the fp32 op order of torchaudio/compliance/kaldi.py:436-511, inter-pass
// FFT twiddles, and the packing of the sparse mel matrix into the per-lane layout the kernel reads.
#pragma once

#include <vector>

#include "fbank_tile.cuh"

namespace psla {

// 0.5 * torch.hann_window(n, periodic=False) in ATen's fp32 op order (kaldi.py:94-95); the 0.5 is
// the 1/2 of the two-real-FFTs-in-one-complex-FFT separation (exact power-of-two scaling).
void build_window_half(float* w, int n);
// dense [num_bins][nfft/2] weights, kaldi.py:436-511 with vtln_warp == 1
void build_mel_dense(float* mel, int num_bins, int nfft, float sample_rate, float low_freq, float high_freq);
// TW[k1][n2] = exp(-2*pi*i*k1*n2/512) as (re, im) pairs, k1 < 32, n2 < 16
void build_twiddles(float* tw);

struct PackedMel {
    short start[kMaxMel];
    short count[kMaxMel];
    int slot_max[kMelSlots];
    int slot_off[kMelSlots];
    std::vector<float> w;  // [sum slot_max][32]
};
// Returns false if the matrix cannot be represented (non-contiguous row support, a row wider than
// kMelMaxCount, weight on bin >= nfft/2, num_bins > 128).
bool pack_mel(const float* dense, int num_bins, int nfft, PackedMel* out);

}  // namespace psla
