"""CPU: the one-FMA voxel keys of the fast keyframe path
(mast3r_slam_quality_dualtsdf_b200/csrc/march_key.h, shared by the CUDA kernels)
against the reference operation floor(fl(fl(o + fl(dist * d)) / vs))
(global_volume.py:63-64, :133-134) on millions of in-band samples: whenever the
fast path accepts a sample its three keys must be the reference keys, for
random rays and for adversarial ones (origins / surfaces on voxel faces,
axis-aligned rays, walls on voxel faces, a 40 m scene)."""
import ctypes as C
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("mk") / "libmkc.so")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fno-fast-math", "-mfma", "-fPIC",
                           "-shared", "-o", out,
                           os.path.join(ROOT, "oracle", "march_key_check.cpp")])
    L = C.CDLL(out)
    L.march_key_check.restype = C.c_int64
    L.march_key_check.argtypes = [C.c_int64, C.c_uint64, C.c_double, C.c_double, C.c_double,
                                  C.c_double, C.c_int, C.c_void_p, C.c_void_p]
    return L


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
@pytest.mark.parametrize("voxel,half", [(0.005, 2.4874), (0.005, 2.5), (0.01, 1.25), (0.03, 6.0),
                                        (0.005, 20.0)])
def test_fast_keys_equal_reference_keys(checker, mode, voxel, half):
    total, acc = C.c_int64(), C.c_int64()
    bad = checker.march_key_check(120000, 11 + mode, voxel, 4 * voxel, 0.5, half, mode,
                                  C.byref(total), C.byref(acc))
    assert bad == 0
    assert total.value > 1_000_000
    # the threshold must not reject the bulk: room-scale scenes keep > 90 % on the fast path
    if half < 10:
        assert acc.value > 0.9 * total.value
