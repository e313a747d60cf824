#!/usr/bin/env python
"""DistributedDataParallel ImageNet training — same CLI, logs and checkpoints as the reference's
``distributed.py`` (/root/reference/distributed.py), running on the B200-native stack.

The epoch loop lives in ``pytorch_distributed_template_b200.trainer`` (one copy
for all three scripts); this file only owns the parser and ``main()``.
"""
from pytorch_distributed_template_b200 import trainer
from pytorch_distributed_template_b200.config import build_parser
from pytorch_distributed_template_b200.utils import reduce_mean  # noqa: F401  (module-level helper of the reference script)

MODE = "ddp"
parser = build_parser(MODE)


def main(argv=None):
    args = parser.parse_args(argv)
    return trainer.run(args, MODE)


if __name__ == "__main__":
    main()
