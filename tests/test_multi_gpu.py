"""Multi-GPU correctness on hardware (BASELINE.json configs[3]): rank r steps its shard_range slice on
cuda:r under torchrun + NCCL; rank 0 gathers and compares bit for bit with the single-GPU run of the same global seed,
and the NCCL reduction of the episode statistics equals the sum of the per-rank structs.  Needs >= 2 GPUs
(tools/gpu_multi.sh runs it under `gpurun --gpus 2`; the log is committed under profiles/)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("env_id,n_total", [("usv-asmc-simple", 4099), ("usv-simple", 2048), ("usv-asmc-v0", 4096)])
def test_shards_on_their_own_devices_equal_the_single_gpu_run(env_id, n_total):
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 CUDA devices")
    world = 2
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "_multi_gpu_worker.py"), str(n_total), "60",
           env_id]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0 and "MULTI_GPU_OK" in r.stdout, (r.stdout[-3000:], r.stderr[-3000:])
