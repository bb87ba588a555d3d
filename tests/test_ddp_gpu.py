"""GPU, two or more devices: kernel-level data-parallel equivalence.  Spawns
``tests/ddp_equivalence.py`` under torchrun, one process per GPU over NCCL; skipped below two devices."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_sharded_step_equals_single_gpu_step():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two CUDA devices (gpurun --gpus 2)")
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "ddp_equivalence.py")]
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    print(proc.stdout[-6000:])
    assert proc.returncode == 0, proc.stdout[-4000:]
    assert "DDP-OK all checks passed" in proc.stdout
