/* psla_frontend.h -- C ABI of the B200-native PSLA feature frontend (libpsla_frontend.so).
 *
 * The reference (haoheliu/psla) has no FFI / plugin registry: its boundary for this path is the
 * Python Dataset protocol, `AudiosetDataset.__getitem__` (src/dataloaders/audioset_dataset.py:139-202),
 * which runs everything below on CPU dataloader workers.  Each entry point here replaces a piece of
 * that method for a whole BATCH of clips on one GPU; the reference-side binding is the ctypes stub in
 * INTEGRATION.md (and psla_b200/_lib.py is the one this repo ships).
 *
 * Conventions
 *   - plain C types only; `stream` is a CUstream / cudaStream_t passed as void*.
 *   - "device" pointers must be valid on the handle's device; the caller owns every buffer.
 *   - forward() is asynchronous on `stream`.  Metadata arrays (offsets, indices, lambda, masks) are HOST
 *     arrays: they are validated synchronously (so errors surface like the reference's Python exceptions)
 *     and staged to the device by the library.
 *   - every function returns PSLA_OK (0) or a negative psla_status; nothing is written to `out` on error.
 *   - a handle may be used from one thread at a time; use one handle per stream for pipelining.
 */
#ifndef PSLA_FRONTEND_H_
#define PSLA_FRONTEND_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct psla_frontend psla_frontend_t;

typedef enum psla_status {
    PSLA_OK = 0,
    PSLA_E_INVALID = -1,      /* null pointer, negative size, index out of range, misaligned buffer */
    PSLA_E_UNSUPPORTED = -2,  /* a frame whose FFT size falls outside [64, 4096] points, > 128 mel bins, or an option that exists
                                 for the 16 kHz streaming kernel only (output layouts, timeline) on the generic path */
    PSLA_E_SHORT_CLIP = -3,   /* a primary waveform has fewer samples than one frame; the reference raises
                                 AssertionError at torchaudio/compliance/kaldi.py:142 */
    PSLA_E_MASK = -4,         /* mask offsets outside [0, num_mel_bins] x [0, target_length] */
    PSLA_E_WORKSPACE = -5,    /* workspace too small / misaligned */
    PSLA_E_CUDA = -6,         /* a CUDA runtime call failed; psla_last_cuda_error() has the code */
    PSLA_E_NOMEM = -7,
    PSLA_E_STATIC_LIMIT = -8  /* psla_frontend_forward_static: more than PSLA_STATIC_MAX_CLIPS clips in one call, or the handle is in a
                                 mode whose launches allocate (set_param_meta(0), set_timing(1)) */
} psla_status;

/* Replaces: the constants of the `kaldi.fbank(...)` call at audioset_dataset.py:119-120 plus the tables that
 * call rebuilds every time (torchaudio/compliance/kaldi.py:94-95 Hann window, :436-511 get_mel_banks).
 * high_freq <= 0 means "offset from Nyquist" exactly like kaldi.py:457-458.
 * Geometry (kaldi.py:138-140): frame length int(sample_rate * frame_length_ms / 1000), shift int(sample_rate *
 * frame_shift_ms / 1000), FFT size the next power of two.  400 / 160 / 512 (16 kHz, 25 / 10 ms: every recipe of the reference)
 * runs the fused streaming kernel; any other geometry with an FFT of 64..4096 points (8 kHz ... 48 kHz and beyond at 25 / 10 ms)
 * runs a geometry-generic kernel: same results contract, clip means always from the statistics kernel, [B, T, F] fp32 output
 * only, not tuned for the roofline. */
int psla_frontend_create(psla_frontend_t** handle, int device, int sample_rate, int num_mel_bins,
                         float frame_length_ms, float frame_shift_ms, float preemph, float low_freq,
                         float high_freq);
void psla_frontend_destroy(psla_frontend_t* handle);

/* Optional: replace the natively built tables (host pointers).  `window` is [frame_length] = the analysis
 * window itself (NOT pre-halved); `mel` is the dense [num_mel_bins][nfft/2] weight matrix (psla_frontend_get_geometry).  Either may be
 * NULL.  The Python host uses this to install tables built with torch ops in torchaudio's op order. */
int psla_frontend_set_tables(psla_frontend_t* handle, const float* window, const float* mel);
/* Read the tables back (host pointers, same shapes; either may be NULL). */
int psla_frontend_get_tables(const psla_frontend_t* handle, float* window, float* mel);

/* Bytes of device scratch forward() needs for a batch of `n_clips` (256-byte aligned pointer required). */
size_t psla_frontend_workspace_bytes(const psla_frontend_t* handle, int n_clips);

/* Replaces: AudiosetDataset._wav2fbank + the SpecAugment / normalisation tail of __getitem__
 * (audioset_dataset.py:88-137, :177-192) for B clips.
 *
 *   wav         DEVICE  packed fp32 waveforms in [-1, 1] (what torchaudio.load returns, channel 0), 16-byte aligned
 *   wav_samples         number of samples `wav` holds: every (offset, length) pair is checked against it, a descriptor that
 *                       reaches past the buffer is PSLA_E_INVALID (never an out-of-bounds device read)
 *   wav_offset  HOST    [n_wav] start sample of every waveform inside `wav` (a multiple of 4 enables 128-bit loads)
 *   wav_length  HOST    [n_wav] number of samples of every waveform
 *   src1        HOST    [B] primary waveform index of each output clip
 *   src2        HOST    [B] mixup partner waveform index, or -1 (no mixup)                       (:147-156)
 *   lambda      HOST    [B] float64 mixup lambda (np.random.beta draw, :114); ignored where src2 < 0
 *   mask        HOST    [B][4] = f_start, f_width, t_start, t_width, drawn on the host exactly like
 *                       torchaudio/functional/functional.py:939-943; width 0 = no mask; may be NULL   (:177-188)
 *   norm_mean, norm_div  fbank = (fbank - norm_mean) / norm_div unless skip_norm                 (:191-192)
 *   out         DEVICE  [B][target_length][num_mel_bins] fp32, 16-byte aligned
 */
int psla_frontend_forward(psla_frontend_t* handle, const float* wav, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length, int n_wav,
                          const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask,
                          int n_clips, int target_length, float norm_mean, float norm_div, int skip_norm,
                          float* out, void* workspace, size_t workspace_bytes, void* stream);

/* Same, with HOST waveform and output buffers (pinned memory makes the copies asynchronous).  The library
 * owns the device staging buffers (grown on demand, kept in the handle).  `out_host` may be NULL to leave the
 * features on the device; psla_frontend_device_out() then returns the device pointer (valid until the next
 * forward_host call on the same handle).  Asynchronous on `stream`. */
int psla_frontend_forward_host(psla_frontend_t* handle, const float* wav_host, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length,
                               int n_wav,
                               const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask,
                               int n_clips, int target_length, float norm_mean, float norm_div, int skip_norm,
                               float* out_host, void* stream);
float* psla_frontend_device_out(const psla_frontend_t* handle);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * psla_frontend_forward() with NO workspace argument and NO allocation: the device scratch of the call -- the per-clip partial
 * sums `clip_stats_kernel` writes and `fbank_stream_kernel` reads (whole-clip means of exact_clip_mean mode and of mixup partners
 * shorter than their clip) -- lives in a file-scope `static __device__` array of the library, sized for at most
 * PSLA_STATIC_MAX_CLIPS clips per call (PSLA_STATIC_SCRATCH_BYTES, held by every GPU that loads the library until the process
 * exits).  Inputs and outputs are the caller's, with the meaning and checks of psla_frontend_forward(); results are bit-identical
 * to it.  The call makes no cudaMalloc / cudaFree / cudaHostAlloc and calls no allocating library routine: descriptors travel in
 * the kernels' parameter space, the array's device address was resolved (cudaGetSymbolAddress) when the handle was created.
 *
 *   n_clips > PSLA_STATIC_MAX_CLIPS                           -> PSLA_E_STATIC_LIMIT (split the batch, or use forward())
 *   handle with set_param_meta(0), set_timing(1) or on the
 *   geometry-generic path (not 16 kHz / set_generic_path)    -> PSLA_E_STATIC_LIMIT (those modes allocate per launch)
 *   the symbol's address could not be resolved at create()    -> PSLA_E_CUDA (psla_last_cuda_error() has the code)
 *
 * ONE array per GPU, shared by every handle and every call: it keeps the last call's contents (nothing in it has to start at
 * zero -- every partial sum a call reads is written by that call's own statistics kernel first), and calls that OVERLAP ON THE
 * GPU RACE on it.  Calls through one handle on one stream are safe (consecutive calls rotate through three partial-sum
 * regions exactly like forward() rotates through its workspace); calls from two handles, or on two streams, that need the
 * statistics pass must be ordered by the caller.  Batches without a statistics pass never touch the array. */
#define PSLA_STATIC_MAX_CLIPS 128
#define PSLA_STATIC_SCRATCH_BYTES (512 * 1024)
int psla_frontend_forward_static(psla_frontend_t* handle, const float* wav, int64_t wav_samples, const int64_t* wav_offset,
                                 const int64_t* wav_length, int n_wav, const int32_t* src1, const int32_t* src2,
                                 const double* lambda, const int32_t* mask, int n_clips, int target_length, float norm_mean,
                                 float norm_div, int skip_norm, float* out, void* stream);

/* The same two calls for 16-bit PCM waveforms as they come out of a wav file (the reference decodes them with
 * torchaudio.load, audioset_dataset.py:91,95-96, whose normalisation is value / 32768; the kernel applies exactly that
 * scale, so the features equal those of the fp32 entry points fed with the decoded floats bit for bit).  Half the input
 * bytes over PCIe and HBM.  Offsets and lengths are in samples; `wav` 16-byte aligned. */
int psla_frontend_forward_pcm16(psla_frontend_t* handle, const int16_t* wav, int64_t wav_samples, const int64_t* wav_offset, const int64_t* wav_length,
                                int n_wav, const int32_t* src1, const int32_t* src2, const double* lambda, const int32_t* mask,
                                int n_clips, int target_length, float norm_mean, float norm_div, int skip_norm,
                                float* out, void* workspace, size_t workspace_bytes, void* stream);
int psla_frontend_forward_host_pcm16(psla_frontend_t* handle, const int16_t* wav_host, int64_t wav_samples, const int64_t* wav_offset,
                                     const int64_t* wav_length, int n_wav, const int32_t* src1, const int32_t* src2,
                                     const double* lambda, const int32_t* mask, int n_clips, int target_length, float norm_mean,
                                     float norm_div, int skip_norm, float* out_host, void* stream);

/* Replaces: the noise / roll augmentation of audioset_dataset.py:197-199 (`noise == True`; off in every recipe of the
 * reference, run.py:77,80) for B clips:   out[b][(t + shift[b]) mod T][f] = in[b][t][f] + (noise[b][t][f] * scale[b]) / 10
 *   in, out   DEVICE  [B][target_length][num_mel_bins] fp32, distinct buffers (a roll cannot run in place)
 *   noise     DEVICE  the U[0,1) field the host drew with torch.rand(T, F) per clip (:198) -- bit-exact parity -- or NULL:
 *                     the kernel then draws it itself (Philox-4x32-10 keyed by `seed`, clip and cell; same distribution)
 *   scale     HOST    [B] fp32(np.random.rand()) per clip (:198)
 *   shift     HOST    [B] np.random.randint(-10, 10) per clip (:199), any value in (-T, T)
 * Asynchronous on `stream`; one extra pass over the features (read + write). */
int psla_frontend_noise_roll(psla_frontend_t* handle, const float* in, float* out, const float* noise, const float* scale,
                             const int32_t* shift, int n_clips, int target_length, uint64_t seed, void* stream);

/* Number of frames kaldi.fbank produces for n_samples (snip_edges=True, kaldi.py:63-67); < 0 if too short. */
int psla_frontend_num_frames(const psla_frontend_t* handle, int64_t n_samples);
/* Frame length, frame shift and FFT size of the handle in samples (kaldi.py:138-140), and whether forward() runs the
 * geometry-generic kernel (any pointer may be NULL). */
int psla_frontend_get_geometry(const psla_frontend_t* handle, int* frame_length, int* frame_shift, int* fft_size, int* generic_path);
/* 1: run the 16 kHz geometry through the geometry-generic kernel as well (an independent second implementation of the
 * recipe: the tests compare the streaming kernel with it); 0 (default): the streaming kernel.  No effect on other geometries. */
int psla_frontend_set_generic_path(psla_frontend_t* handle, int enable);
/* Kernels launched by this handle since creation (bench.py reports it as gpu_launches). */
int64_t psla_frontend_launch_count(const psla_frontend_t* handle);
/* Select the log implementation: 0 = logf (default), 1 = lg2.approx based (abs err < 4e-6 on log-mel). */
int psla_frontend_set_fast_log(psla_frontend_t* handle, int enable);

/* Output layout of the forward entry points (the consumer, src/models/Models.py:104-107, turns the
 * dataset's [B, T, F] into [B, 1, F, T] with unsqueeze(1).transpose(2, 3) before its first convolution; under autocast it
 * is then cast to a 16-bit type).  With a frequency-major layout the kernel writes what that convolution reads:
 *   PSLA_OUT_TF_F32   fp32 [n_clips, target_length, num_mel_bins]   the reference's (default)
 *   PSLA_OUT_FT_F32   fp32 [n_clips, num_mel_bins, target_length]   == out_tf.transpose(1, 2), bit for bit
 *   PSLA_OUT_FT_BF16  bf16 [n_clips, num_mel_bins, target_length]   == the same, rounded to nearest-even
 * The frequency-major layouts need num_mel_bins == 128, target_length % 4 == 0 and a 16-byte aligned `out`, and are not
 * available together with exact_clip_mean or the timeline (PSLA_E_UNSUPPORTED at the forward call otherwise).  `out` /
 * `out_host` keep their `float*` type in the signatures; for PSLA_OUT_FT_BF16 they point to n_clips * 128 *
 * target_length 16-bit elements.  psla_frontend_noise_roll works on the default layout only. */
#define PSLA_OUT_TF_F32 0
#define PSLA_OUT_FT_F32 1
#define PSLA_OUT_FT_BF16 2
int psla_frontend_set_output_layout(psla_frontend_t* handle, int layout);

/* Whole-clip DC removal (audioset_dataset.py:92 / :98-99 / :117).  Every constant the reference subtracts from a whole
 * clip is removed again by kaldi.fbank's per-frame DC removal (kaldi.py:183-186), so the default (0) subtracts a
 * five-sample median of the clip (first, quartile and last samples: robust against a click or a loud onset) instead of
 * its mean -- same features up to fp32 rounding, no extra pass over the waveform; clip statistics are then computed
 * only for mixup clips whose partner is shorter (its mean fills the zero-padded tail).
 * 1 = compute the three means with a statistics kernel and evaluate the reference's fp32 expression verbatim. */
int psla_frontend_set_exact_clip_mean(psla_frontend_t* handle, int enable);

/* Per-kernel device timing for bench.py's roofline: while enabled, forward() brackets its kernels with CUDA
 * events on `stream`; get_timing() synchronises those events, returns the accumulated milliseconds of the
 * clip-statistics kernel (0 when it was not needed) and of the fused streaming kernel plus the number of forward() calls, and resets. */
int psla_frontend_set_timing(psla_frontend_t* handle, int enable);
int psla_frontend_get_timing(psla_frontend_t* handle, double* stats_ms, double* tile_ms, int64_t* n_forward);

/* Optional per-warp timeline of the streaming kernel (tracing aid; the reference has no counterpart).  While a DEVICE
 * buffer of psla_frontend_trace_records() records (one per warp) x psla_frontend_trace_words() int64 is installed,
 * every warp of every forward() writes
 *   [0] globaltimer ns at warp start   [1] globaltimer ns at warp end
 *   [2] cycles before its first group | total cycles << 32
 *   [3] smid | groups computed << 16 | constant (masked / padded) groups << 32 | cold starts << 48
 *   [4 + i]: i-th event: cycles since warp start | kind << 40 | groups in the step << 44; kinds 1..3 open a loop step
 *            (1 computed group, 2 computed group that staged its whole window, 3 run of constant groups), 4..9 mark
 *            phases inside it (4 clip entered, 5 window staged, 6 pass 1, 7 pass 2 + power, 8 mel, 9 epilogue done)
 * Pass NULL to switch it off.  The timeline is compiled into separate kernel variants (the production kernels carry none
 * of its checks) that exist for the default clip-mean mode only: forward() returns PSLA_E_UNSUPPORTED while a trace buffer is
 * installed together with psla_frontend_set_exact_clip_mean(1). */
int psla_frontend_set_trace(psla_frontend_t* handle, int64_t* device_buf, size_t capacity_records);
int psla_frontend_trace_records(const psla_frontend_t* handle);
int psla_frontend_trace_words(void);

/* Weights of the work partition between the warps of the streaming kernel (performance tuning only; results do not
 * depend on them): cost of a computed group of 4 rows, of a group of 4 constant rows, and the surcharge of the first
 * computed group after constant rows / at a clip start.  Defaults are measured on B200. */
int psla_frontend_set_cost_model(psla_frontend_t* handle, int live, int dead, int cold);

/* Where the per-call clip descriptors (a few KB) travel to the device: 1 (default) on an internal copy stream, ordered
 * before the kernel by an event, so that the kernels of consecutive forward() calls run back to back on `stream`;
 * 0: as a copy on `stream` itself.  Results do not depend on it. */
int psla_frontend_set_meta_stream(psla_frontend_t* handle, int enable);

/* Overlap of consecutive launches of the streaming kernel (programmatic dependent launch).  The kernel signals at its
 * start that the next kernel on the stream may begin, and joins its predecessor before it exits.  A launch made with
 * programmatic stream serialization therefore starts its thread blocks on SM resources the previous launch has already
 * left (the tail of one batch overlaps the head of the next), yet never COMPLETES before that launch: whatever follows
 * on the stream (a model step, a copy, an event) still sees both finished, and at most two launches are in flight.
 * The kernel does NOT wait before reading its inputs, so the library attaches the attribute only behind a launch it
 * knows to be its own:
 *   PSLA_OVERLAP_OFF     never.
 *   PSLA_OVERLAP_CALL    (default) between the launches of ONE forward() call: a batch of more than 128 clips goes out as
 *                        chunks of at most 128 clips, and chunk k + 1 overlaps chunk k.  Needs nothing from the caller.
 *   PSLA_OVERLAP_STREAM  also between consecutive forward() calls of this handle on the same stream.  CALLER CONTRACT: any
 *                        work enqueued on that stream between two such calls must not itself release its dependents
 *                        early (griddepcontrol.launch_dependents / cudaTriggerProgrammaticLaunchCompletion) before it
 *                        has written this call's inputs -- some library GEMMs do.  Plain kernels, copies, memsets and
 *                        event waits in between are fine: the launch simply waits for them in full (no overlap then).
 * In every mode a launch only overlaps when it is at most 128 clips with its descriptors in parameter space (see
 * psla_frontend_set_param_meta), needs no statistics pass, and writes an output range disjoint from those of the
 * previous three launches of the handle; otherwise it is a plain launch.  Results do not depend on the mode. */
#define PSLA_OVERLAP_OFF 0
#define PSLA_OVERLAP_CALL 1
#define PSLA_OVERLAP_STREAM 2
int psla_frontend_set_launch_overlap(psla_frontend_t* handle, int mode);

/* 1 (default): batches of at most 128 clips that need no statistics pass carry their clip descriptors in the kernel's
 * parameter space (no upload; required for the launch overlap above), and larger batches that need no statistics pass
 * are validated as a whole and then enqueued as balanced chunks of at most 128 clips, each such a launch
 * (psla_frontend_launch_count advances by the number of chunks).  0: always upload them, one launch per call.  Results
 * do not depend on it. */
int psla_frontend_set_param_meta(psla_frontend_t* handle, int enable);

const char* psla_strerror(int status);
int psla_last_cuda_error(const psla_frontend_t* handle);
const char* psla_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PSLA_FRONTEND_H_ */
