"""GPU (-m gpu), needs >= 2 GPUs (skipped on a 1-GPU box): one process per GPU under torchrun, halo rows
stored peer-to-peer over NVLink; the ring must equal the single-GPU result bit for bit."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ring_over_gpus_equals_single_gpu():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tools", "mgpu_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    import re
    m = re.search(r"\[mgpu\] (\d+)/(\d+) cases BIT-EXACT", out.stdout)
    assert m and m.group(1) == m.group(2) and int(m.group(2)) >= 15, out.stdout[-3000:]
    assert "MISMATCH" not in out.stdout


def test_ring_of_two_processes_on_one_gpu():
    """What a one-GPU box can check of the multi-process path: two ranks share GPU 0, the slabs find each other through CUDA IPC
    handles (shipped over gloo), order their steps with the device-flag barrier (fused into the seam launches of the in-place
    engine) and must reproduce the single-slab lattice bit for bit; then one rank stops and the other must get an error after
    one time-out instead of hanging (the ring's abort word)."""
    import re
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29519", os.path.join(ROOT, "tools", "mgpu_check.py"), "--same-gpu", "--abort"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    m = re.search(r"\[mgpu\] (\d+)/(\d+) cases BIT-EXACT", out.stdout)
    assert m and m.group(1) == m.group(2) and int(m.group(2)) >= 14, out.stdout[-3000:]
    assert "MISMATCH" not in out.stdout and "abort: rank 0 got" in out.stdout and "-> OK" in out.stdout, out.stdout[-2000:]
