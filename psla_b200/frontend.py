"""Batch-level GPU frontend: the host-side mirror of `AudiosetDataset._wav2fbank` + the SpecAugment /
normalisation tail of `__getitem__` (src/dataloaders/audioset_dataset.py:88-137, 177-192), for a
whole batch on one B200 through the C ABI in include/psla_frontend.h.

    fe = GpuFbankFrontend(audio_conf)                 # same dict the reference's dataset takes
    fbank = fe(waves, src1, src2, lam, masks)         # cuda FloatTensor [B, target_length, num_mel_bins]

`fbank` is exactly the `[batch, time_frame_num, frequency_bins]` tensor `EffNetAttention.forward`
expects (src/models/Models.py:104-107).  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _lib, tables

SAMPLE_RATE = 16000
OUTPUT_LAYOUTS = {"tf": 0, "ft": 1, "ft_bf16": 2}  # PSLA_OUT_TF_F32 / PSLA_OUT_FT_F32 / PSLA_OUT_FT_BF16


@dataclass
class PackedWaves:
    """Waveforms packed back to back in one buffer, every start a multiple of 4 samples: fp32 in [-1, 1] (what
    torchaudio.load returns) or int16 PCM (what the wav file holds; the kernel scales by 1 / 32768 like the decoder)."""
    wav: torch.Tensor  # 1-D float32 or int16 (cuda, or pinned / pageable host memory)
    offsets: np.ndarray  # int64 [n_wav]; waveform i = wav[offsets[i] : offsets[i] + lengths[i]]
    lengths: np.ndarray  # int64 [n_wav]

    @property
    def n_wav(self) -> int:
        return len(self.lengths)


def pack_waveforms(waves: Sequence[torch.Tensor], device: Union[str, torch.device, None] = None,
                   pinned: bool = False, out: Optional[torch.Tensor] = None) -> PackedWaves:
    """Pack [1, n] / [n] waveforms (channel 0, as kaldi.fbank selects it, kaldi.py:136) with each start rounded up to
    a multiple of 4 samples so the kernels can use vector loads.  All-int16 input is packed as int16 PCM, anything else
    as fp32.  `out`: pack into this 1-D host tensor (e.g. a slot of page-locked shared memory) instead of a new one."""
    lens = np.array([int(w.shape[-1]) for w in waves], dtype=np.int64)
    starts = np.zeros(len(waves), dtype=np.int64)
    pos = 0
    for i, n in enumerate(lens):
        starts[i] = pos
        pos += (int(n) + 3) // 4 * 4
    dtype = torch.int16 if len(waves) > 0 and all(w.dtype == torch.int16 for w in waves) else torch.float32
    if out is not None:
        if out.dtype != dtype or out.dim() != 1 or out.numel() < max(pos, 8):
            raise ValueError("out must be a 1-D tensor of the packed dtype with room for the batch")
        flat = out[:max(pos, 8)]
        for i in range(len(waves)):  # the gaps between waveforms (start alignment) are never read as samples of a clip
            flat[starts[i] + lens[i]:(starts[i + 1] if i + 1 < len(waves) else max(pos, 8))] = 0
    else:
        flat = torch.zeros(max(pos, 8), dtype=dtype, pin_memory=pinned)
    for i, w in enumerate(waves):
        if hasattr(w, "read_into"):  # samples still in their file (dataloaders.LazyPcm16): read them into place
            if dtype == torch.int16:
                w.read_into(flat[starts[i]:starts[i] + lens[i]])
                continue
            w = w.materialize()
        w = w.detach()
        w = w[0] if w.dim() == 2 else w
        if dtype == torch.float32 and w.dtype == torch.int16:  # a batch that mixes decoded floats and 16-bit PCM: the
            w = w.to(torch.float32) / 32768.0                  # decoder's normalisation (torchaudio.load: value / 32768)
        flat[starts[i]:starts[i] + lens[i]] = w.to(dtype)
    if device is not None:
        flat = flat.to(device, non_blocking=pinned)
    return PackedWaves(flat, starts, lens)


def to_pcm16(wave: torch.Tensor) -> torch.Tensor:
    """fp32 waveform in [-1, 1) -> the int16 PCM a 16-bit wav file of it holds (round to nearest, saturate)."""
    return torch.clamp(torch.round(wave.to(torch.float32) * 32768.0), -32768, 32767).to(torch.int16)


def _raw_stream(device: torch.device) -> int:
    """cudaStream_t of torch's current stream on `device` as an integer (the private fast path when it exists: the public
    accessor builds a Stream object per call)."""
    try:
        return torch._C._cuda_getCurrentRawStream(device.index)
    except AttributeError:  # pragma: no cover
        return torch.cuda.current_stream(device).cuda_stream


def _as_c(arr: Optional[np.ndarray], dtype, ctype=None):
    """(contiguous array of `dtype` -- keep it alive across the call --, its address as an int for a void* argument)."""
    if arr is None:
        return None, None
    a = np.ascontiguousarray(arr, dtype=dtype)
    return a, a.ctypes.data


class GpuFbankFrontend:
    """Kaldi-compatible log-mel + SpecAugment + normalisation on the GPU.

    audio_conf keys (same as the reference, audioset_dataset.py:60-80,122): num_mel_bins, target_length,
    mean, std, skip_norm (optional).  freqm / timem / mixup are consumed by the host-side parameter draw
    (psla_b200.dataloaders), the kernel receives the drawn offsets.

    exact_clip_mean: see psla_frontend_set_exact_clip_mean in include/psla_frontend.h.
    norm_std_scale: 1.0 reproduces this reference (`/ norm_std`, :192); 2.0 gives the `/ (norm_std * 2)`
    form of upstream PSLA/AST that BASELINE.json's north_star sentence quotes.
    """

    def __init__(self, audio_conf: dict, device: Union[int, str, torch.device, None] = None,
                 norm_std_scale: float = 1.0, fast_log: bool = True, tables_from: str = "torch",
                 sample_rate: int = SAMPLE_RATE, exact_clip_mean: bool = False):
        if not torch.cuda.is_available():
            raise RuntimeError("GpuFbankFrontend needs a CUDA device; psla_b200 has no CPU path")
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.type != "cuda":
            raise ValueError("device must be a CUDA device")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.audio_conf = dict(audio_conf)
        self.num_mel_bins = int(audio_conf.get("num_mel_bins"))
        self.target_length = int(audio_conf.get("target_length"))
        self.skip_norm = bool(audio_conf.get("skip_norm")) if audio_conf.get("skip_norm") else False
        self.norm_mean = float(audio_conf.get("mean")) if audio_conf.get("mean") is not None else 0.0
        std = audio_conf.get("std")
        self.norm_std = float(std) if std is not None else 1.0
        self.norm_std_scale = float(norm_std_scale)
        # python-float arithmetic first, one rounding to fp32 when it meets the tensor (as torch does)
        self.norm_div = self.norm_std * self.norm_std_scale
        self.sample_rate = int(sample_rate)
        h = ctypes.c_void_p()
        st = self.lib.psla_frontend_create(ctypes.byref(h), self.device.index, self.sample_rate, self.num_mel_bins,
                                           25.0, 10.0, 0.97, 20.0, 0.0)
        _lib.check(st, "psla_frontend_create")
        self._h = h
        fl, fs, nfft, gen = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.psla_frontend_get_geometry(self._h, ctypes.byref(fl), ctypes.byref(fs), ctypes.byref(nfft), ctypes.byref(gen)),
                   "psla_frontend_get_geometry", self._h)
        # kaldi.py:138-140: 400 / 160 / 512 at 16 kHz (the fused streaming kernel); other sample rates run the generic kernel
        self.frame_length, self.frame_shift, self.fft_size, self.generic_path = fl.value, fs.value, nfft.value, bool(gen.value)
        if tables_from == "torch":
            win = tables.hann_window(self.frame_length).numpy()
            mel = tables.mel_banks(self.num_mel_bins, self.fft_size, float(self.sample_rate)).numpy()
            st = self.lib.psla_frontend_set_tables(self._h, win.ctypes.data, np.ascontiguousarray(mel).ctypes.data)
            _lib.check(st, "psla_frontend_set_tables", self._h)
        elif tables_from != "native":
            raise ValueError("tables_from must be 'torch' or 'native'")
        self.set_fast_log(fast_log)
        self.set_exact_clip_mean(exact_clip_mean)
        self.output_layout = "tf"
        self.launch_overlap = "call"
        self._ws = {}

    # ---- lifetime ----
    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self.lib.psla_frontend_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- knobs / introspection ----
    def set_fast_log(self, enable: bool):
        self.fast_log = bool(enable)
        _lib.check(self.lib.psla_frontend_set_fast_log(self._h, int(self.fast_log)), "set_fast_log", self._h)

    def set_exact_clip_mean(self, enable: bool):
        """False (default): the whole-clip DC removal is folded into the per-frame one (no pass over the clip);
        True: clip means from a statistics kernel, the reference's fp32 expression verbatim."""
        self.exact_clip_mean = bool(enable)
        _lib.check(self.lib.psla_frontend_set_exact_clip_mean(self._h, int(self.exact_clip_mean)), "set_exact_clip_mean", self._h)

    def set_generic_path(self, enable: bool):
        """Run the 16 kHz geometry through the geometry-generic kernel too (psla_frontend_set_generic_path; tests)."""
        _lib.check(self.lib.psla_frontend_set_generic_path(self._h, int(bool(enable))), "set_generic_path", self._h)
        self.generic_path = bool(enable) or (self.frame_length, self.frame_shift, self.fft_size) != (400, 160, 512)

    def set_output_layout(self, layout: str = "tf"):
        """Layout of the features the forward calls write (psla_frontend_set_output_layout):
          "tf"       float32  [B, T, F]  what the reference's dataset returns (default)
          "ft"       float32  [B, F, T]  == tf.transpose(1, 2) bit for bit: `x.unsqueeze(1)` is the [B, 1, F, T] the model's
                                         first convolution reads (Models.py:106-107 does unsqueeze(1).transpose(2, 3))
          "ft_bf16"  bfloat16 [B, F, T]  the same, rounded to nearest-even (what autocast would cast it to)
        The frequency-major layouts need 128 mel bins and target_length % 4 == 0."""
        if layout not in OUTPUT_LAYOUTS:
            raise ValueError(f"layout must be one of {sorted(OUTPUT_LAYOUTS)}")
        _lib.check(self.lib.psla_frontend_set_output_layout(self._h, OUTPUT_LAYOUTS[layout]), "set_output_layout", self._h)
        self.output_layout = layout

    def out_shape_dtype(self, B: int, T: int):
        if self.output_layout == "tf":
            return (B, T, self.num_mel_bins), torch.float32
        return (B, self.num_mel_bins, T), (torch.bfloat16 if self.output_layout == "ft_bf16" else torch.float32)

    def set_timing(self, enable: bool):
        _lib.check(self.lib.psla_frontend_set_timing(self._h, int(bool(enable))), "set_timing", self._h)

    def get_timing(self):
        """(stats_kernel_ms, tile_kernel_ms, n_forward) accumulated since the last call; synchronises."""
        a, b, n = ctypes.c_double(), ctypes.c_double(), ctypes.c_int64()
        _lib.check(self.lib.psla_frontend_get_timing(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(n)),
                   "get_timing", self._h)
        return a.value, b.value, n.value

    def set_cost_model(self, live: int, dead: int, cold: int):
        """Work-partition weights of the streaming kernel (tuning only; see psla_frontend_set_cost_model)."""
        _lib.check(self.lib.psla_frontend_set_cost_model(self._h, int(live), int(dead), int(cold)), "set_cost_model", self._h)

    def set_meta_stream(self, enable: bool):
        """Descriptor upload on the library's copy stream (default) or on the launch stream (psla_frontend_set_meta_stream)."""
        _lib.check(self.lib.psla_frontend_set_meta_stream(self._h, int(bool(enable))), "set_meta_stream", self._h)

    def set_launch_overlap(self, mode="call"):
        """Programmatic overlap of consecutive launches (psla_frontend_set_launch_overlap):
          "off" / False   plain launches
          "call" / True   (default) between the chunks of one forward call (batches above 128 clips)
          "stream"        also between consecutive forward calls on the same stream -- only if nothing enqueued on that
                          stream in between releases its own dependents early (see include/psla_frontend.h)."""
        modes = {"off": 0, False: 0, 0: 0, "call": 1, True: 1, 1: 1, "stream": 2, 2: 2}
        if mode not in modes:
            raise ValueError('mode must be "off", "call" or "stream"')
        self.launch_overlap = ("off", "call", "stream")[modes[mode]]
        _lib.check(self.lib.psla_frontend_set_launch_overlap(self._h, modes[mode]), "set_launch_overlap", self._h)

    def set_param_meta(self, enable: bool):
        """Clip descriptors of small batches as kernel parameters (default) or uploaded (psla_frontend_set_param_meta)."""
        _lib.check(self.lib.psla_frontend_set_param_meta(self._h, int(bool(enable))), "set_param_meta", self._h)

    def set_trace(self, enable: bool) -> Optional[torch.Tensor]:
        """Per-warp timeline of the streaming kernel (see psla_frontend_set_trace).  Returns the int64 [warps, words]
        device tensor the next forward calls fill, or None when switched off."""
        if not enable:
            _lib.check(self.lib.psla_frontend_set_trace(self._h, None, 0), "set_trace", self._h)
            self._trace = None
            return None
        n = int(self.lib.psla_frontend_trace_records(self._h))
        self._trace = torch.zeros((n, int(self.lib.psla_frontend_trace_words())), dtype=torch.int64, device=self.device)
        _lib.check(self.lib.psla_frontend_set_trace(self._h, self._trace.data_ptr(), n), "set_trace", self._h)
        return self._trace

    @property
    def launch_count(self) -> int:
        return int(self.lib.psla_frontend_launch_count(self._h))

    def num_frames(self, n_samples: int) -> int:
        return int(self.lib.psla_frontend_num_frames(self._h, int(n_samples)))

    def get_tables(self):
        win = np.zeros(self.frame_length, np.float32)
        mel = np.zeros((self.num_mel_bins, self.fft_size // 2), np.float32)
        _lib.check(self.lib.psla_frontend_get_tables(self._h, win.ctypes.data, mel.ctypes.data), "get_tables")
        return win, mel

    def _workspace(self, n_clips: int) -> torch.Tensor:
        need = self._ws.get(n_clips)
        if need is None:
            need = self._ws[n_clips] = int(self.lib.psla_frontend_workspace_bytes(self._h, n_clips))
        ws = self._ws.get("buf")
        if ws is None or ws.numel() < need:
            ws = torch.empty(max(need, 1 << 16), dtype=torch.uint8, device=self.device)
            self._ws["buf"] = ws
        return ws

    # ---- the hot path ----
    def forward_packed(self, packed: PackedWaves, src1, src2=None, lam=None, masks=None,
                       out: Optional[torch.Tensor] = None, target_length: Optional[int] = None) -> torch.Tensor:
        """Device-resident waveforms -> [B, T, F] on the current CUDA stream (asynchronous)."""
        T = self.target_length if target_length is None else int(target_length)
        wav = packed.wav
        if not wav.is_cuda:
            raise ValueError("forward_packed needs device-resident waveforms; use forward_host for host buffers")
        B = len(src1)
        shape, dtype = self.out_shape_dtype(B, T)
        if out is None:
            out = torch.empty(shape, dtype=dtype, device=self.device)
        elif tuple(out.shape) != shape or out.dtype != dtype or not out.is_contiguous() or not out.is_cuda:
            raise ValueError(f"out must be a contiguous {dtype} {list(shape)} cuda tensor (output layout '{self.output_layout}')")
        if B == 0:
            return out
        ao, p_off = _as_c(packed.offsets, np.int64, ctypes.c_int64)
        an, p_len = _as_c(packed.lengths, np.int64, ctypes.c_int64)
        a1, p1 = _as_c(np.asarray(src1), np.int32, ctypes.c_int32)
        a2, p2 = _as_c(None if src2 is None else np.asarray(src2), np.int32, ctypes.c_int32)
        al, pl = _as_c(None if lam is None else np.asarray(lam), np.float64, ctypes.c_double)
        am, pm = _as_c(None if masks is None else np.asarray(masks).reshape(B, 4), np.int32, ctypes.c_int32)
        ws = self._workspace(B)
        stream = _raw_stream(self.device)
        if wav.dtype not in (torch.float32, torch.int16):
            raise ValueError("waveforms must be float32 or int16 PCM")
        fwd = self.lib.psla_frontend_forward_pcm16 if wav.dtype == torch.int16 else self.lib.psla_frontend_forward
        st = fwd(
            self._h, wav.data_ptr(), wav.numel(), p_off, p_len, packed.n_wav, p1, p2, pl, pm, B, T,
            self.norm_mean, self.norm_div, int(self.skip_norm), out.data_ptr(), ws.data_ptr(), ws.numel(), stream)
        _lib.check(st, "psla_frontend_forward", self._h)
        return out

    def noise_roll(self, fbank: torch.Tensor, scale, roll, noise: Optional[torch.Tensor] = None, seed: int = 0,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """The noise / roll augmentation of audioset_dataset.py:197-199 on a batch of features:
        `roll(fbank[b] + noise[b] * scale[b] / 10, roll[b], dim=time)`.  `scale` / `roll`: the host's np.random.rand() /
        np.random.randint(-10, 10) draws per clip.  `noise`: the torch.rand(T, F) fields the host drew ([B, T, F], moved to
        the device here) for bit-exact parity, or None to let the kernel draw them (Philox, keyed by `seed`)."""
        if not fbank.is_cuda or fbank.dtype != torch.float32 or not fbank.is_contiguous() or fbank.dim() != 3:
            raise ValueError("fbank must be a contiguous float32 [B, T, F] cuda tensor")
        B, T, F = fbank.shape
        if F != self.num_mel_bins:
            raise ValueError("feature width differs from num_mel_bins")
        if out is None:
            out = torch.empty_like(fbank)
        if noise is not None:
            noise = noise.to(self.device, dtype=torch.float32, non_blocking=True).contiguous()
            if tuple(noise.shape) != (B, T, F):
                raise ValueError("noise must be [B, T, F]")
        asc = np.ascontiguousarray(np.asarray(scale, dtype=np.float64).astype(np.float32).reshape(B))
        ash = np.ascontiguousarray(np.asarray(roll, dtype=np.int32).reshape(B))
        st = self.lib.psla_frontend_noise_roll(
            self._h, fbank.data_ptr(), out.data_ptr(), None if noise is None else noise.data_ptr(),
            asc.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), ash.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
            B, T, int(seed) & (2**64 - 1), torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(st, "psla_frontend_noise_roll", self._h)
        return out

    def forward_host(self, packed: PackedWaves, src1, src2=None, lam=None, masks=None,
                     out_host: Optional[torch.Tensor] = None, target_length: Optional[int] = None,
                     stream: Optional[torch.cuda.Stream] = None) -> Optional[torch.Tensor]:
        """Host waveforms (ideally pinned) -> features through the host-buffer C entry point.  With `out_host`
        (a pinned float32 [B, T, F] tensor) the result is copied back; without it a cuda tensor view of the
        library-owned output buffer is returned (valid until the next forward_host on this frontend)."""
        T = self.target_length if target_length is None else int(target_length)
        wav = packed.wav
        if wav.is_cuda:
            raise ValueError("forward_host takes host waveforms")
        B = len(src1)
        ao, p_off = _as_c(packed.offsets, np.int64, ctypes.c_int64)
        an, p_len = _as_c(packed.lengths, np.int64, ctypes.c_int64)
        a1, p1 = _as_c(np.asarray(src1), np.int32, ctypes.c_int32)
        a2, p2 = _as_c(None if src2 is None else np.asarray(src2), np.int32, ctypes.c_int32)
        al, pl = _as_c(None if lam is None else np.asarray(lam), np.float64, ctypes.c_double)
        am, pm = _as_c(None if masks is None else np.asarray(masks).reshape(B, 4), np.int32, ctypes.c_int32)
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        out_ptr = None
        shape, dtype = self.out_shape_dtype(B, T)
        if out_host is not None:
            if tuple(out_host.shape) != shape or out_host.dtype != dtype or out_host.is_cuda or not out_host.is_contiguous():
                raise ValueError(f"out_host must be a contiguous host {dtype} {list(shape)} tensor (output layout '{self.output_layout}')")
            out_ptr = out_host.data_ptr()
        if wav.dtype not in (torch.float32, torch.int16):
            raise ValueError("waveforms must be float32 or int16 PCM")
        fwd = self.lib.psla_frontend_forward_host_pcm16 if wav.dtype == torch.int16 else self.lib.psla_frontend_forward_host
        st = fwd(
            self._h, wav.data_ptr(), wav.numel(), p_off, p_len, packed.n_wav, p1, p2, pl, pm, B, T,
            self.norm_mean, self.norm_div, int(self.skip_norm), out_ptr, s)
        _lib.check(st, "psla_frontend_forward_host", self._h)
        if out_host is not None:
            return out_host
        if dtype == torch.bfloat16:  # the array interface has no bf16: view the 16-bit words, reinterpret
            v = _DeviceView(self.lib.psla_frontend_device_out(self._h), shape, self.device, "<i2").tensor()
            return v.view(torch.bfloat16)
        return _DeviceView(self.lib.psla_frontend_device_out(self._h), shape, self.device).tensor()

    def __call__(self, waves, src1=None, src2=None, lam=None, masks=None, **kw) -> torch.Tensor:
        """waves: PackedWaves or a sequence of [1, n] / [n] tensors.  src1 defaults to one clip per waveform."""
        if not isinstance(waves, PackedWaves):
            waves = pack_waveforms(waves, device=self.device)
        elif not waves.wav.is_cuda:
            waves = PackedWaves(waves.wav.to(self.device, non_blocking=True), waves.offsets, waves.lengths)
        if src1 is None:
            src1 = np.arange(waves.n_wav, dtype=np.int32)
        return self.forward_packed(waves, src1, src2, lam, masks, **kw)


class _DeviceView:
    """Zero-copy torch view of a library-owned device buffer (via __cuda_array_interface__)."""

    def __init__(self, ptr: int, shape, device, typestr: str = "<f4"):
        self._device = device
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 3, "strides": None}

    def tensor(self) -> torch.Tensor:
        return torch.as_tensor(self, device=self._device)
