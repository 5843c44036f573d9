from .datasets import CUBDataset, DatasetCollection, prepare_dataloader  # noqa: F401
from .synthetic import SyntheticImages, SyntheticTokens, pinned_batches  # noqa: F401
