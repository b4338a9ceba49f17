"""GPU, two or more devices: the gradient exchange fused into the backward (sl_backward_batch_push,
splatloc_b200.exchange) must leave, on every GPU, the same sums an NCCL all-reduce of the per-GPU
gradients gives (float atomics: within the suite's gradient tolerance; the visibility counts exactly).
Runs tools/check_exchange.py under torchrun with two ranks; skipped on a one-GPU box."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fused_exchange_equals_nccl_all_reduce_on_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, CHECK_P="100000")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29577", os.path.join(ROOT, "tools", "check_exchange.py")]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "fused exchange == NCCL all-reduce" in r.stdout
    # the same sums through the split-phase barrier (arrive / wait inside the replayed step), ranks skewed
    assert "split-phase exchange == NCCL all-reduce" in r.stdout
    # and with the rows sent by sl_exchange_push beside the next step's forward
    assert "deferred-push exchange == NCCL all-reduce" in r.stdout
    # several views per GPU per step: summed gradients, and the rows' visibility slot counts views
    assert "two views per GPU per step == NCCL all-reduce" in r.stdout
