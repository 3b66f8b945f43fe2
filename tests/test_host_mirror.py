"""The C++ mirror of the reference's Dealer interface (nano-gpu-scheduler_b200/host): its test program
re-runs the reference's own unit-test tables (pkg/dealer/rater_test.go, allocate_test.go) and
Dealer-level sequences through the C ABI.  tests/host/test_dealer_host.cpp holds the cases."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "nano-gpu-scheduler_b200", "host", "test_dealer_host")


def test_host_mirror_builds_and_fails_loudly_without_gpu(built):
    import torch

    assert os.path.exists(BIN)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    res = subprocess.run([BIN], capture_output=True, text=True, timeout=60)
    assert res.returncode == 2 and "no usable CUDA device" in res.stdout  # no CPU fallback


@pytest.mark.gpu
def test_reference_tests_through_the_host_mirror(built):
    res = subprocess.run([BIN], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0 and "HOST TESTS OK" in res.stdout, res.stdout[-3000:] + res.stderr[-1000:]
