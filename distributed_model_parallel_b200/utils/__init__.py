from .checkpoint import load_checkpoint, merge_pipeline_shards, save_best, save_pipeline_shard, stage_key_map  # noqa: F401
from .config import TrainConfig, build_parser, config_from_args  # noqa: F401
from .logging import EpochLogger, get_lr  # noqa: F401
from .metrics import AverageMeter, accuracy  # noqa: F401
from .timing import StepTimer  # noqa: F401
