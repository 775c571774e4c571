"""Drop-in for the reference's `dataloaders` package (src/dataloaders/__init__.py:8)."""
from .audioset_dataset import AudiosetDataset, RawBatch, RawItem, FrontendLoader, make_index_dict, make_name_dict, lookup_list

__all__ = ["AudiosetDataset", "RawBatch", "RawItem", "FrontendLoader", "make_index_dict", "make_name_dict", "lookup_list"]
