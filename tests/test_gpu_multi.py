"""Multi-GPU correctness on hardware: runs scripts/dp_check.py under torchrun on 2 GPUs of the box.
Skipped on a one-GPU box (the driver's round-end GPU tier has one); `gpurun --gpus 2` runs it for real."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mode", ["", "p2p"])
def test_dp_training_and_sharded_render_on_two_gpus(mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, NERF_B200_DP=mode)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "scripts", "dp_check.py")], env=env, capture_output=True, text=True, timeout=600)
    line = [x for x in r.stdout.splitlines() if x.startswith("{")]
    print(r.stdout[-1500:], r.stderr[-1500:])
    assert r.returncode == 0 and line, r.stderr[-3000:]
    res = json.loads(line[-1])
    assert res["ok"], res
