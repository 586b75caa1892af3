"""Multi-rank exchange (one process per GPU in production) on CPU: world_size-2/3
`gloo` runs of DeviceExchange.sort_distributed over the host-memory test double.
Covers the host logic of the N>1 path: splitter choice, per-row destination,
all-to-all-v of every column kind, receive-side sort and block cuts."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,extra", [(2, []), (3, []), (2, ["--empty-rank"])])
def test_sort_distributed_gloo(world, extra):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "_dist_worker.py")] + extra
    env = dict(os.environ, OMP_NUM_THREADS="1")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert p.returncode == 0 and f"DIST-OK {world}" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]


@pytest.mark.gpu
def test_sort_shuffle_aggregate_nccl_two_gpus():
    """The same worker on real GPUs: CUDA kernels, NCCL collectives and the NVLink
    peer-memory exchange (needs two visible GPUs; the round-end single-GPU run skips it,
    the forced single-rank tests cover the same code there)."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "_dist_worker.py"), "--gpu"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert p.returncode == 0 and "DIST-OK 2" in p.stdout, p.stdout[-2000:] + p.stderr[-6000:]
