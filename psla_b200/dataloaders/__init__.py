"""Drop-in for the reference's `dataloaders` package (src/dataloaders/__init__.py:8)."""
from .audioset_dataset import AudiosetDataset, RawBatch, RawItem, FrontendLoader, make_index_dict, make_name_dict, lookup_list

from .norm_stats import NormStats, compute_norm_stats

__all__ = ["NormStats", "compute_norm_stats", "AudiosetDataset", "RawBatch", "RawItem", "FrontendLoader", "make_index_dict", "make_name_dict", "lookup_list"]
