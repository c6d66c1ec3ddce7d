"""N=2 over NCCL against the undivided mesh on one GPU (needs two GPUs; the driver's single-GPU run skips it)."""
import os, subprocess, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["RANS", "NAVIER_STOKES", "EULER", "EULER_JST", "NAVIER_STOKES_JST",
                                  "ITER_RANS", "ITER_EULER", "ITER_TET_EULER", "ITER_TET_RANS", "ITER_SST", "ITER_MIXED"])
def test_two_slabs_match_the_undivided_mesh(kind):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "check_multigpu.py"), kind]
    p = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "multigpu parity" in p.stdout
