"""Real multi-GPU run (skipped on a 1-GPU box): torchrun with one process per GPU, CUDA IPC
peer mapping, flag barrier and the fused push/apply/pull kernel over NVLink."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("layout", ["balanced", "logical"])
def test_two_ranks_match_oracle_and_each_other(lib, layout):
    n = lib.device_count_or_zero()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "scripts", "mgpu_check.py"),
           layout]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "MGPU_CHECK PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_two_ranks_sparse_embedding_matches_oracle(lib):
    """The sparse path over real peers: CUDA-IPC mapped shards and staging buffers, stream-ordered barrier."""
    if lib.device_count_or_zero() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29613", os.path.join(ROOT, "scripts", "mgpu_sparse_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "MGPU_SPARSE_CHECK PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
