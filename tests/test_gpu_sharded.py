"""`-m gpu` test of the product's sharded path (viso_ba_shard_init, the in-place ncclAllReduce of the reduced
camera system, the grouped read-back): two ranks under torch.distributed.run, one process per GPU, against the CPU
oracle, the one-GPU run and each other (tests/shard_parity.py).  Needs two GPUs on the box: skipped on a
one-GPU box, where the same exchange is covered on CPU by tests/test_sharding_gloo.py (gloo, world size 2) and the
multi-GPU numbers come from bench.py's `sharded` section in the driver's scaling run."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpu_count():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=60).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("config,poses,points", [(3, 60, 20000), (4, 150, 15000)], ids=["dense_S", "banded_S"])
def test_two_ranks_match_oracle_one_gpu_and_each_other(vb, oracle, tmp_path, config, poses, points):
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs on one box (the driver's GPU test box has one)")
    out = str(tmp_path / "shard.json")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "shard_parity.py"), "--config", str(config),
           "--poses", str(poses), "--points", str(points), "--out", out]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    r = json.load(open(out))
    assert r["ok"] and r["ranks_bit_identical"] and r["world"] == 2
    assert r["cost_rel_vs_oracle"] < 1e-9 and r["pose_abs_vs_oracle"] < 1e-8 and r["point_abs_vs_oracle"] < 1e-7
