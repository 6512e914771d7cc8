"""Local (per-ROI, dense) TSDF of the refiner on the GPU.

Host-side mirror of the numeric methods of the reference TSDFRefiner
(mast3r_slam/tsdf_refine.py): `build_tsdf_robust` = _build_tsdf_robust
(:837-940), `extract_surface_safe` = _extract_surface_safe (:942-1021, which
samples through _sample_tsdf_trilinear :1023-1064).  Same argument meaning and
return values; CUDA through the C ABI (include/b200tsdf.h), no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import _ffi


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _cfg(cfg):
    c = _ffi.LocalCfg()
    c.voxel_size = float(cfg["voxel_size"])
    c.trunc_dist = float(cfg["trunc_dist"])
    c.min_confidence = float(cfg.get("min_confidence", 0.2))
    c.max_displacement = float(cfg.get("max_displacement", 0.015))
    c.max_grid_dim = int(cfg["max_grid_dim"])
    c.ray_samples = int(cfg.get("ray_samples", 64))
    return c


def _box(xyz_min, xyz_max):
    lo = torch.as_tensor(xyz_min).detach().float().cpu().numpy().reshape(3)
    hi = torch.as_tensor(xyz_max).detach().float().cpu().numpy().reshape(3)
    return np.ascontiguousarray(np.concatenate([lo, hi]), np.float32)


def _pose8(T_WC, device):
    data = T_WC.data if hasattr(T_WC, "data") else T_WC
    return torch.as_tensor(data).detach().reshape(-1)[:8].to(device=device, dtype=torch.float32) \
        .contiguous()


def grid_dims(cfg, xyz_min, xyz_max):
    """(nx, ny, nz) of the ROI grid (tsdf_refine.py:844-846)."""
    box = _box(xyz_min, xyz_max)
    dims = (C.c_int32 * 3)()
    c = _cfg(cfg)
    _ffi.check(_ffi.lib().b200tsdf_local_dims(box.ctypes.data_as(C.c_void_p), C.byref(c), dims),
               "local_dims")
    return int(dims[0]), int(dims[1]), int(dims[2])


def build_tsdf_robust(cfg, X_canon, C_flat, K, xyz_min, xyz_max, H, W, T_WC, device=None,
                      return_fused=False):
    """Reference signature minus `self` (tsdf_refine.py:837).  K, H, W are
    unused there as well.  Returns (tsdf, weights) float32 (nz, ny, nx) device
    tensors."""
    device = torch.device(device) if device is not None else X_canon.device
    if device.type != "cuda":
        raise _ffi.B200TSDFError("build_tsdf_robust needs a CUDA device (no CPU fallback)")
    X = X_canon.detach().to(device=device, dtype=torch.float32).reshape(-1, 3).contiguous()
    Cf = C_flat.detach().to(device=device, dtype=torch.float32)
    if Cf.ndim > 1:
        Cf = Cf.squeeze(-1)  # :854-855
    Cf = Cf.reshape(-1).contiguous()
    box = _box(xyz_min, xyz_max)
    nx, ny, nz = grid_dims(cfg, xyz_min, xyz_max)
    tsdf = torch.empty((nz, ny, nx), device=device, dtype=torch.float32)
    weights = torch.empty_like(tsdf)
    pose = _pose8(T_WC, device)
    c = _cfg(cfg)
    fused = C.c_int64(0)
    _ffi.check(_ffi.lib().b200tsdf_local_build(
        X.data_ptr(), Cf.data_ptr(), pose.data_ptr(), box.ctypes.data_as(C.c_void_p), X.shape[0],
        C.byref(c), tsdf.data_ptr(), weights.data_ptr(),
        C.byref(fused) if return_fused else None, _stream(device)), "local_build")
    if return_fused:
        return tsdf, weights, int(fused.value)
    return tsdf, weights


def extract_surface_safe(cfg, tsdf, xyz_min, xyz_max, K, mask, H, W, X_original, generator=None,
                         order=None):
    """Reference signature minus `self` (tsdf_refine.py:942).  The reference
    draws torch.randperm(len(pixel_indices))[:100]; pass `order` (positions in
    the mask's pixel list) to fix the choice, or a torch.Generator.
    Returns (X_refined (N,3), hits bool (mask.sum(),))."""
    device = tsdf.device
    X = X_original.detach().to(device=device, dtype=torch.float32).reshape(-1, 3).contiguous()
    X_refined = X.clone()
    mask = mask.to(device).reshape(-1)
    pixel_indices = torch.where(mask)[0]
    hits = torch.zeros(int(pixel_indices.numel()), dtype=torch.bool, device=device)
    if pixel_indices.numel() == 0:  # :949-950
        return X_refined, hits
    n_proc = min(int(pixel_indices.numel()), 100)  # :961
    if order is None:
        order = torch.randperm(int(pixel_indices.numel()), device=device, generator=generator)
    order = torch.as_tensor(order, device=device, dtype=torch.int64)[:n_proc]
    pix = pixel_indices[order].contiguous()
    nz, ny, nx = tsdf.shape
    dims = (C.c_int32 * 3)(nx, ny, nz)
    box = _box(xyz_min, xyz_max)
    c = _cfg(cfg)
    h8 = torch.zeros(n_proc, dtype=torch.uint8, device=device)
    t = tsdf.detach().to(torch.float32).contiguous()
    _ffi.check(_ffi.lib().b200tsdf_local_raycast(
        t.data_ptr(), dims, box.ctypes.data_as(C.c_void_p), X.data_ptr(), pix.data_ptr(), n_proc,
        C.byref(c), X_refined.data_ptr(), h8.data_ptr(), _stream(device)), "local_raycast")
    hits[order] = h8.bool()  # :1011 hits[pixel_array_idx] = True
    return X_refined, hits
