"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol
include/b200tsdf.h declares (no compute calls -- there is no GPU here)."""
import ctypes
import os
import re

from mast3r_slam_quality_dualtsdf_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200tsdf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200tsdf_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _declared() == sorted(_ffi.SYMBOLS)


def test_library_exports_every_declared_symbol(built_lib):
    L = ctypes.CDLL(built_lib)
    for name in _declared():
        assert hasattr(L, name), f"{name} missing from libb200tsdf.so"
    L.b200tsdf_abi_version.restype = ctypes.c_int
    hdr = open(os.path.join(ROOT, "include", "b200tsdf.h")).read()
    assert L.b200tsdf_abi_version() == int(re.search(r"B200TSDF_ABI_VERSION (\d+)", hdr).group(1))


def test_config_struct_layout():
    # must match b200tsdf_config in include/b200tsdf.h: 4 doubles, 2 i32, 6 i64, 2 i32, 2 i32, 1 i64
    assert ctypes.sizeof(_ffi.Config) == 4 * 8 + 2 * 4 + 6 * 8 + 2 * 4 + 2 * 4 + 8


def test_kf_batch_struct_layout():
    # b200tsdf_kf_batch: 4 pointers, 6 i32, i64, f32, i32, f64, 2 i32, u64, pointer
    assert ctypes.sizeof(_ffi.KfBatch) == 4 * 8 + 6 * 4 + 8 + 4 + 4 + 8 + 2 * 4 + 8 + 8
    assert _ffi.KfBatch.seg_stride.offset == 56 and _ffi.KfBatch.step_scale.offset == 72


def test_create_without_gpu_fails_loudly(built_lib):
    import torch

    if torch.cuda.is_available():
        return
    cfg = _ffi.Config()
    cfg.voxel_size, cfg.truncation, cfg.max_weight, cfg.min_weight = 0.01, 0.04, 100.0, 1e-3
    cfg.capacity_log2 = 8
    h = ctypes.c_void_p()
    rc = _ffi.lib().b200tsdf_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc != 0 and _ffi.lib().b200tsdf_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "mast3r_slam_quality_dualtsdf_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "liboracle" not in txt, f
