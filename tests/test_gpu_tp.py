"""Tensor-parallel GPU test: needs two GPUs on the box; skipped otherwise. The single-GPU emulation of
the fused reduction is tests/test_gpu_kernels.py::test_tp_fused_reduce_add_norm; the host-side sharding logic is
covered on CPU with gloo in tests/test_host_logic.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("scatter_min", [str(1 << 30), "1"])
def test_tp2_fused_reduce_matches_nccl_allreduce(scatter_min):
    """scatter_min = 1: every exchange uses the two-step form (qspec_tp_rs_add_norm_*: per-row owners, pushes and per-row flags
    over NVLink); the default is the one-step form."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(here, "tp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env=dict(os.environ, QSPEC_TP_SCATTER_MIN=scatter_min))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "TP_OK" in r.stdout, r.stdout[-3000:]
