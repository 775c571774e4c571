"""Drop-in for the reference's `dataloaders` package (src/dataloaders/__init__.py:8)."""
from .audioset_dataset import (AudiosetDataset, FrontendLoader, PinnedSlotPool, RawBatch, RawItem, default_loader, lookup_list,
                               make_index_dict, make_name_dict, pcm16_lazy_loader, pcm16_loader)
from .norm_stats import NormStats, compute_norm_stats

__all__ = ["NormStats", "compute_norm_stats", "AudiosetDataset", "RawBatch", "RawItem", "FrontendLoader", "PinnedSlotPool",
           "default_loader", "pcm16_loader", "pcm16_lazy_loader", "make_index_dict", "make_name_dict", "lookup_list"]
