"""Multi-GPU parity (needs >= 2 GPUs on the box): the node-range sharded path, one rank per GPU under
torchrun, against the CPU oracle evaluated on the whole cluster."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world, nodes, *extra):
    env = dict(os.environ)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "tests", "sharded_worker.py"), str(nodes), *extra]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert "SHARDED OK" in res.stdout, res.stdout[-2000:] + res.stderr[-3000:]


def test_sharded_two_ranks(built):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, 3000)


def test_sharded_two_ranks_with_nodes_of_9_to_12_gpus(built):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, 2000, "wide")


def test_sharded_all_ranks(built):
    import torch

    n = torch.cuda.device_count()
    if n < 4:
        pytest.skip("needs >= 4 GPUs")
    _run(n, 1500)
