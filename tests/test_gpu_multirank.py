"""Multi-GPU parity through the C-ABI (needs >= 2 GPUs on the box; skipped otherwise): tools/mgpu_parity.py under
torchrun -- hex8 / tet4 / quad4 slabs on 2 ranks, interface rows moved by NCCL, by femx's peer push kernel and by
the assembly kernel's own remote stores; merged owned rows vs the 1-rank oracle (pattern bit-exact, values and
internal force <= 1e-12)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_parity_all_exchanges():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "mgpu_parity.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MGPU PARITY OK" in r.stdout and "MISMATCH" not in r.stdout
