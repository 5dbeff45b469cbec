"""The C++ host layer (bipedal-locomotion-framework_b200/host): the reference's class surface over the C ABI.
The C++ test program mirrors the reference's Catch2 tests; here it is built and run."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "bipedal-locomotion-framework_b200", "host")
BIN = os.path.join(HOST, "tests", "host_api_test")
MODEL = os.path.join(ROOT, "tests", "golden", "ergocub23.model.txt")


def _build():
    lib = os.path.join(ROOT, "bipedal-locomotion-framework_b200", "libblfik_cuda.so")
    assert os.path.exists(lib), "build the CUDA library first (__graft_entry__.build())"
    subprocess.check_call(["make", "-C", HOST, "-s"])


def test_model_file_matches_generator():
    import blf_b200  # noqa: F401
    from blf_b200 import models as M
    assert open(MODEL).read() == M.ergocub23().to_text()


def test_host_layer_cpu_checks():
    """Parameter handler, variables handler, weight provider, task initialisation keys, addTask / setTaskWeight rules,
    error conventions of the solver and of the integrator -- no GPU needed."""
    _build()
    r = subprocess.run([BIN, "cpu", MODEL], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failures" in r.stdout


@pytest.mark.gpu
def test_host_layer_gpu_checks():
    """Task A / b identities against the GPU kinematics, the ergoCub closed loop built from a parameter handler
    (N = 1 and N = 48, with and without limit tasks), setTaskWeight, and the integrator's closed form."""
    _build()
    r = subprocess.run([BIN, "gpu", MODEL], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failures" in r.stdout
