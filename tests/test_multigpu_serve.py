"""Multi-GPU serving tier (needs >= 2 B200s): TP=N greedy decoding == TP=1 on nccl / fused / fused+graphs."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def test_tp_serving_all_gpus():
    """TP over every visible GPU (2, 4 or 8 ranks)."""
    import torch

    n = torch.cuda.device_count()
    n = 8 if n >= 8 else (4 if n >= 4 else n)
    if n < 2:
        pytest.skip("needs >= 2 GPUs (the TP kernels are multi-GPU by construction)")
    r = subprocess.run([sys.executable, "-m", "dlti_b200.launch", f"--num_gpus={n}", "--master_port=29741",
                        "--no_local_rank", os.path.join(ROOT, "tests", "mp_serve_worker.py")],
                       capture_output=True, text=True, timeout=900, cwd=ROOT, env=dict(os.environ, PYTHONPATH=ROOT))
    print(r.stdout[-3000:])
    assert r.returncode == 0 and "ALL TP SERVING CHECKS PASSED" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
