"""Two ranks, two GPUs, NCCL: column shards computed per rank and gathered through the C ABI equal the single-GPU result bit for
bit (tests/multigpu_worker.py).  Skipped on a box with one GPU; the CPU suite covers the partition logic with gloo
(tests/test_shard_gloo.py) and the plan's host logic (tests/test_cabi_cpu.py)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("mode", ["nccl", "p2p"])
@pytest.mark.parametrize("world", [2])
def test_c_abi_gather_matches_single_gpu(world, mode):
    """mode nccl: one message per field and peer + layout copies on the root; p2p: rrtmgp_b200_shard_plan_register, every rank
    writes its slabs into the root's arrays over NVLink (CUDA IPC)."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(HERE, "multigpu_worker.py")], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, RB200_TEST_GATHER=mode))
    assert r.returncode == 0 and "MULTIGPU_GATHER_OK" in r.stdout, (r.stdout[-2000:], r.stderr[-3000:])
