"""psla_b200 -- B200-native (sm_100a) feature frontend of haoheliu/psla.

Only the data-parallel hot path is here: the Kaldi-compatible log-mel filterbank + SpecAugment +
normalisation that the reference's `AudiosetDataset` computes on CPU dataloader workers, as hand-written
CUDA behind a C ABI (include/psla_frontend.h), plus the host-side mirror of the reference's
dataset/frontend interface.  The model and the training loop stay stock PyTorch.
"""
__version__ = "0.1.0"


def __getattr__(name):  # lazy: importing the package must not require the CUDA library (build() imports it)
    if name == "GpuFbankFrontend":
        from .frontend import GpuFbankFrontend
        return GpuFbankFrontend
    if name in ("PackedWaves", "pack_waveforms"):
        from . import frontend
        return getattr(frontend, name)
    raise AttributeError(name)
