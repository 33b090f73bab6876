"""Two ranks on two GPUs (NCCL): the sharded photon pass and the banded frame must equal the single-GPU results byte for
byte.  Runs tools/mgpu_check.py under torchrun; skipped on a box with fewer than two GPUs."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_two_rank_photon_maps_and_frame_match_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tools", "mgpu_check.py"), "3000000", "480", "270"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    assert "photon maps identical to single-GPU on all ranks: True" in out
    assert "BVH bytes identical across ranks: True" in out
    assert "banded frame identical to single-GPU frame: True" in out
    assert "final gathering: direct / radiance photons (with Lo) and banded frame identical to single-GPU: True" in out
