"""Command line front-end.

    python -m distributed_model_parallel_b200.cli dp  [DIR] [--lr ...] [--resume]
    python -m distributed_model_parallel_b200.cli ddp [DIR] --world-size N [...]
    python -m distributed_model_parallel_b200.cli mp  [DIR] --world-size N --schedule 1f1b --microbatches 4 [...]

``dp`` / ``mp`` mirror the reference's two scripts (/root/reference/code/distributed_training/data_parallel.py,
model_parallel.py) including their flag names; ``ddp`` is the multi-process trainer the reference only describes."""
from __future__ import annotations

import sys

from .utils import build_parser, config_from_args


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] not in ("dp", "ddp", "mp"):
        print(__doc__)
        return 2
    mode, rest = argv[0], argv[1:]
    cfg = config_from_args(build_parser("distributed_model_parallel_b200 %s trainer" % mode).parse_args(rest))
    if mode == "dp":
        from .train import dp as t
    elif mode == "ddp":
        from .train import ddp as t
    else:
        from .train import mp as t
    t.main(cfg)
    return 0


if __name__ == "__main__":
    sys.exit(main())
