"""N > 1 on real GPUs (skipped on a single-GPU box; `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`):
the path's cross-rank exchange over NVLink peer memory (pbhc_peer_allreduce_sum) against the rank-ordered fp64 sum, under
CUDA-graph replay, and the sharded TrackingStep.finalize against the single-process step - tools/peer_check.py under torchrun."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.timeout(600)
def test_peer_allreduce_and_sharded_finalize_two_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port", "29631",
           os.path.join(ROOT, "tools", "peer_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=580, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["world"] == 2 and "sharded_finalize" in line
