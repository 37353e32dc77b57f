#!/bin/bash
# usage: tools/run_n.sh N script args...   (torchrun on one node, rendezvous on 127.0.0.1)
n=$1; shift
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port ${PORT:-29511} "$@"
