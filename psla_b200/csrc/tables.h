// Host-side constant tables of the fbank frontend (pure C++, no CUDA): Hann window, HTK-mel
// triangle weights in the fp32 op order of torchaudio/compliance/kaldi.py:436-511, inter-pass
// FFT twiddles, and the packing of the sparse mel matrix into the per-lane layout the kernel reads.
#pragma once

#include <vector>

#include "fbank_stream.cuh"

namespace psla {

// 0.5 * torch.hann_window(n, periodic=False) in ATen's fp32 op order (kaldi.py:94-95); the 0.5 is
// the 1/2 of the two-real-FFTs-in-one-complex-FFT separation (exact power-of-two scaling).
void build_window_half(float* w, int n);
// dense [num_bins][nfft/2] weights, kaldi.py:436-511 with vtln_warp == 1
void build_mel_dense(float* mel, int num_bins, int nfft, float sample_rate, float low_freq, float high_freq);
// The flat constant-table buffer the kernel copies into shared memory (layout: kTab* in fbank_stream.cuh):
//   WINT[n2][28]  0.5 * window[16*n1 + n2]  (the 0.5 is the 1/2 of the two-real-FFTs-in-one separation)
//   TW[t][n2]     t >= 1: exp(-2*pi*i*t*n2/512);  t == 0: exp(-2*pi*i*n2/32)   (re, im), row stride 36 floats
//   TWZ[n2]       exp(+2*pi*i*n2/32)
//   mel           per (slot, lane): first P bin, owned mel bin, zero-padded weights [tap][lane]
// `window` is the analysis window itself [400]; `mel_dense` is [num_bins][nfft/2].
// Returns false if the mel matrix cannot be represented (non-contiguous row support, a row wider than kMelMaxCount,
// num_bins > 128).  *mel_spec is set when the tap counts fit the fully unrolled path (kMelSpec*).
bool build_table_buffer(const float* window, const float* mel_dense, int num_bins, std::vector<float>* buf, bool* mel_spec);

}  // namespace psla
