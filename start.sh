#!/usr/bin/env bash
# Same three commands as the reference's start.sh, through torchrun (the modern
# spelling of `python -m torch.distributed.launch`), LF line endings, rendezvous on
# 127.0.0.1.  NGPU defaults to every visible GPU.
set -e
NGPU=${NGPU:-$(python -c 'import torch; print(max(1, torch.cuda.device_count()))')}
python dataparallel.py "$@"
python -m torch.distributed.run --nnodes=1 --nproc-per-node="$NGPU" --master-addr 127.0.0.1 --master-port=23334 distributed.py "$@"
python -m torch.distributed.run --nnodes=1 --nproc-per-node="$NGPU" --master-addr 127.0.0.1 --master-port=23334 distributed_syncBN_amp.py "$@"
