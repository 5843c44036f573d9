from .synthetic import SyntheticImages, SyntheticTokens, pinned_batches  # noqa: F401
