#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun on N GPUs of one box (kept for the round scripts; the check itself
lives in tests/shard_parity.py, which the `-m gpu` test tests/test_gpu_sharded.py and bench.py also use):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/shard_check.py [--points 20000]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))

from shard_parity import main  # noqa: E402

if __name__ == "__main__":
    main()
