"""Builds and runs the C++ adapter test (tests/cpp/test_adapters.cpp) against libhg_b200.so."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_adapters.bin")


def _build():
    import hectorgrapher_b200 as hg
    if not os.path.exists(hg.capi.LIB_PATH):
        hg.build()
    lib_dir = os.path.dirname(hg.capi.LIB_PATH)
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-pthread", "-o", BIN,
                           os.path.join(ROOT, "tests", "cpp", "test_adapters.cpp"), "-L" + lib_dir, "-lhg_b200",
                           "-Wl,-rpath," + lib_dir, "-Wl,-rpath,/usr/local/cuda/lib64"])


def test_adapters_compile_and_link():
    _build()
    out = subprocess.check_output([BIN, "compile-only"], text=True)
    assert "sm_100a" in out


@pytest.mark.gpu
def test_adapters_on_gpu():
    _build()
    out = subprocess.run([BIN], text=True, capture_output=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ADAPTERS OK" in out.stdout


def test_host_pose_preparation():
    """tests/cpp/test_pose_prep.cpp: the product's host-side InterpolateTransform + slerp Jacobians
    against transform/timestamped_transform_test.cc and central differences (CPU only)."""
    exe = os.path.join(ROOT, "tests", "cpp", "test_pose_prep.bin")
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-ffp-contract=off", "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "test_pose_prep.cpp")])
    out = subprocess.run([exe], text=True, capture_output=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "POSE PREP OK" in out.stdout


def test_branch_free_lower_voxel_selection():
    """tests/cpp/test_lower_voxel.cpp: the arithmetic of the cost kernels' branch-free lower-voxel
    selection against the reference's division + lround, 20 M values incl. cell boundaries (CPU only)."""
    exe = os.path.join(ROOT, "tests", "cpp", "test_lower_voxel.bin")
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-ffp-contract=off", "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "test_lower_voxel.cpp")])
    out = subprocess.run([exe], text=True, capture_output=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "LOWER VOXEL OK" in out.stdout


def test_closed_form_rotation_terms():
    """tests/cpp/test_rotation_terms.cpp: the closed form of grad . d world / d q the kernels use,
    against central differences of Eigen's quaternion * vector formula (CPU only)."""
    exe = os.path.join(ROOT, "tests", "cpp", "test_rotation_terms.bin")
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-ffp-contract=off", "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "test_rotation_terms.cpp")])
    out = subprocess.run([exe], text=True, capture_output=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ROTATION TERMS OK" in out.stdout
