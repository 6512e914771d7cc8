"""GPU: b200tsdf_extract_mesh_static -- the extraction pass that allocates nothing (its
neighbour table is a file-scope device array of the library) -- against the existing entry
point b200tsdf_extract_mesh, through the C ABI, on two calls in a row.  The entry has no
counterpart in the reference.  Triangle counts
bit-exact, triangle soups equal as multisets: the tolerance of tests/test_gpu_mesh.py.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from mast3r_slam_quality_dualtsdf_b200 import _ffi, synth

pytestmark = pytest.mark.gpu


def _canon(tri):
    t = np.ascontiguousarray(tri.reshape(-1, 9))
    return t[np.lexsort(t.T[::-1])]


def _fused_volume(dev, capacity_log2, scene="sphere", seed=4):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    kf = synth.make_keyframes(synth.SCENES[scene], 3, H=48, W=64, seed=seed, f=50.0, total=12)
    vol = TSDFVolume(0.03, 0.12, capacity_log2=capacity_log2, device=dev)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose")),
                            torch.from_numpy(kf["q"]).to(dev), 0.05, 48, 64)
    return vol


def _static(vol, which, buf, cnt):
    L = _ffi.lib()
    st = C.c_void_p(torch.cuda.current_stream(vol.device).cuda_stream)
    return L.b200tsdf_extract_mesh_static(vol._h, which, C.c_void_p(buf.data_ptr()),
                                          buf.shape[0] // 3, C.c_void_p(cnt.data_ptr()), st)


def test_static_extraction_equals_the_allocating_entry_twice(cuda_device):
    L = _ffi.lib()
    vol = _fused_volume(cuda_device, 14)
    other = _fused_volume(cuda_device, 13, scene="plane", seed=7)  # shares the static table
    st = C.c_void_p(torch.cuda.current_stream(cuda_device).cuda_stream)
    for which in (0, 1):
        n = C.c_int64()
        _ffi.check(L.b200tsdf_extract_mesh(vol._h, which, None, 0, C.byref(n), st), "count")
        want = torch.empty(n.value * 3, 3, dtype=torch.float32, device=cuda_device)
        n2 = C.c_int64()
        _ffi.check(L.b200tsdf_extract_mesh(vol._h, which, C.c_void_p(want.data_ptr()), n.value,
                                           C.byref(n2), st), "write")
        assert n2.value == n.value and n.value > 100
        want = _canon(want.cpu().numpy())
        # poison what the entry must clear or overwrite itself
        cnt = torch.full((1,), 12345, dtype=torch.int64, device=cuda_device)
        for call in range(2):
            buf = torch.full((n.value * 3 + 30, 3), float("nan"), dtype=torch.float32,
                             device=cuda_device)
            _ffi.check(_static(vol, which, buf, cnt), "extract_mesh_static")
            assert int(cnt.item()) == n.value, f"call {call}"
            got = buf[: n.value * 3].cpu().numpy()
            assert np.array_equal(_canon(got), want), f"call {call}"
            assert torch.isnan(buf[n.value * 3:]).all()
            # a different volume in between leaves its rows in the shared table
            ocnt = torch.zeros(1, dtype=torch.int64, device=cuda_device)
            obuf = torch.empty(30, 3, dtype=torch.float32, device=cuda_device)
            _ffi.check(_static(other, which, obuf, ocnt), "extract_mesh_static (other volume)")
            m = C.c_int64()
            _ffi.check(L.b200tsdf_extract_mesh(other._h, which, None, 0, C.byref(m), st), "count")
            assert int(ocnt.item()) == m.value  # count exact although only 10 triangles fit
        # count-only call (no output buffer)
        _ffi.check(L.b200tsdf_extract_mesh_static(vol._h, which, None, 0,
                                                  C.c_void_p(cnt.data_ptr()), st), "count only")
        assert int(cnt.item()) == n.value


def test_static_extraction_rejects_volumes_beyond_its_table(cuda_device):
    L = _ffi.lib()
    big = _fused_volume(cuda_device, 17)
    cnt = torch.zeros(1, dtype=torch.int64, device=cuda_device)
    buf = torch.empty(30, 3, dtype=torch.float32, device=cuda_device)
    rc = _static(big, 0, buf, cnt)
    assert rc == _ffi.E_INVALID
    msg = L.b200tsdf_last_error().decode()
    assert "65536" in msg and "hash slots" in msg
    with pytest.raises(_ffi.B200TSDFError):
        _ffi.check(rc, "extract_mesh_static")
    rc = L.b200tsdf_extract_mesh_static(big._h, 2, None, 0, C.c_void_p(cnt.data_ptr()), None)
    assert rc == _ffi.E_INVALID
