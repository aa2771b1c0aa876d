"""One process per GPU over NCCL (the torchrun-style host): the shared fit must be bit-identical to the replicated one on
every rank, and a block that is not positive definite must fail every rank the same way.  Needs >= 2 GPUs on the box
(skipped otherwise; on one GPU the same schedule is covered through the group in tests/test_gpu_group.py and, on CPU, by
tests/test_host_logic.py::test_shared_fit_schedule_over_gloo).  Driver: scripts/dist_fit_check.py."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shared_fit_over_nccl():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ranks = min(n, 4)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={ranks}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "scripts", "dist_fit_check.py"), "300,1500,3300", ""],
                       capture_output=True, text=True, timeout=900)
    print(r.stdout[-3000:], r.stderr[-1500:])
    assert r.returncode == 0 and "dist_fit_check OK" in r.stdout
