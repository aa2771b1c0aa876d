"""The ROS-free C++ host (host_example/main.cpp) drives the C-ABI exactly as the node would: build it with g++
and run three BO steps (full, then two rank-1 appends + incremental steps) on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_cpp_host_runs_three_steps(pkg):
    pkg.build()
    subprocess.run(["make", "-C", os.path.join(ROOT, "host_example")], check=True, capture_output=True)
    r = subprocess.run([os.path.join(ROOT, "host_example", "host_example"), "300", "256", "192"], capture_output=True,
                       text=True, timeout=300)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "host_example OK" in r.stdout
    # the tcgen05 kernel served the full step, the thin incremental kernel the two rank-1 steps
    assert r.stdout.count("kernel=2") == 1 and r.stdout.count("kernel=3") == 2


def test_cpp_host_fails_loudly_without_gpu(pkg):
    """No GPU here: sbo_create must refuse (no CPU fallback) and the host exits non-zero."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    pkg.build()
    subprocess.run(["make", "-C", os.path.join(ROOT, "host_example")], check=True, capture_output=True)
    r = subprocess.run([os.path.join(ROOT, "host_example", "host_example")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "no CUDA device" in r.stderr
