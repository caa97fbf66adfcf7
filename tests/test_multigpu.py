"""Multi-GPU tier (needs >= 2 B200s): peer-memory kernels vs NCCL, fused vs nccl training backends."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def test_fused_backend_all_gpus():
    """Every visible GPU takes part (2, 4 or 8 ranks): rank rotation, slot strides and barrier fan-in differ."""
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (K1 / K9 / peer-memory kernels are multi-GPU by construction)")
    r = subprocess.run([sys.executable, "-m", "dlti_b200.launch", f"--num_gpus={n}", "--master_port=29731",
                        "--no_local_rank", os.path.join(ROOT, "tests", "mp_fused_worker.py")],
                       capture_output=True, text=True, timeout=900, cwd=ROOT,
                       env=dict(os.environ, PYTHONPATH=ROOT))
    print(r.stdout[-4000:])
    assert r.returncode == 0 and "ALL MULTI-GPU CHECKS PASSED" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
