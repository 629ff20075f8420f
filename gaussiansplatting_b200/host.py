"""ctypes binding of libgs_host.so — the C++ host layer (gaussiansplatting_b200/host/: the reference's TiledRasterizer /
GPURadixSort / AdamOptimizer / DensityController class interfaces and the headless train loop of mtl_engine.mm) driven
from Python.  No CPU fallback: a missing library or CUDA device raises GsError."""
import ctypes as C
import os
import numpy as np
from .capi import GsError
from .layouts import GAUSSIAN

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgs_host.so")
_lib = None


class TrainConfig(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("num_views", C.c_uint32), ("epochs", C.c_uint32),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("scene_extent", C.c_float),
                ("densify_from", C.c_uint32), ("densify_until", C.c_uint32), ("densify_interval", C.c_uint32),
                ("opacity_reset_interval", C.c_uint32), ("seed", C.c_uint32), ("quantize_rgba8", C.c_int),
                ("log_loss_every_step", C.c_int), ("export_dir", C.c_char_p), ("load_checkpoint", C.c_char_p),
                ("save_checkpoint", C.c_char_p), ("export_ply", C.c_char_p), ("start_iteration", C.c_uint32),
                ("total_iterations", C.c_uint32), ("permute_moments", C.c_int)]


class TrainResult(C.Structure):
    _fields_ = [("seconds", C.c_double), ("densify_seconds", C.c_double), ("iterations", C.c_uint64),
                ("densify_events", C.c_uint64), ("opacity_resets", C.c_uint64), ("pruned", C.c_uint64),
                ("cloned", C.c_uint64), ("split", C.c_uint64), ("final_count", C.c_uint32), ("adam_timestep", C.c_uint32),
                ("last_pairs", C.c_uint64), ("first_loss", C.c_float), ("last_loss", C.c_float), ("views_exported", C.c_uint32)]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GsError(f"{LIB_PATH} is missing: run `python -m gaussiansplatting_b200.build`")
        _lib = C.CDLL(LIB_PATH)
        _lib.gs_host_last_error.restype = C.c_char_p
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def train(gaussians, gt_gaussians, quats, trans, cfg, device=0, want_output=False):
    """Engine::train (mtl_engine.mm:1047-1221) on synthetic views; returns (TrainResult dict, trained Gaussians or None)"""
    g = np.ascontiguousarray(gaussians, GAUSSIAN)
    gt = np.ascontiguousarray(gt_gaussians, GAUSSIAN)
    q = np.ascontiguousarray(quats, np.float32)
    t = np.ascontiguousarray(trans, np.float32)
    res = TrainResult()
    cap = 1500000 + len(g)
    out = np.zeros(cap, GAUSSIAN) if want_output else None
    rc = lib().gs_host_train(C.c_int(device), _p(g), C.c_uint32(len(g)), _p(gt), C.c_uint32(len(gt)), _p(q), _p(t),
                             C.byref(cfg), C.byref(res), _p(out) if want_output else None, C.c_uint32(cap))
    if rc != 0:
        raise GsError("gs_host_train: " + lib().gs_host_last_error().decode())
    return res.asdict(), (out[:res.final_count] if want_output else None)


def sort64(keys, vals, device=0):
    """GPURadixSort64::sort through the C++ class"""
    k = np.ascontiguousarray(keys, np.uint64)
    v = np.ascontiguousarray(vals, np.uint32)
    ko, vo = np.zeros_like(k), np.zeros_like(v)
    rc = lib().gs_host_sort64(C.c_int(device), _p(k), _p(v), C.c_uint64(len(k)), _p(ko), _p(vo))
    if rc != 0:
        raise GsError("gs_host_sort64: " + lib().gs_host_last_error().decode())
    return ko, vo


def render_view(gaussians, quat, trans, fx, fy, cx, cy, width, height, rgba8_target=False, device=0):
    """One viewer frame through the C++ Engine::render (MTLEngine::render, mtl_engine.mm:401-609, headless): returns
    (image HxWx4 float32, (tile, instance) pair count)"""
    g = np.ascontiguousarray(gaussians, GAUSSIAN)
    q = np.ascontiguousarray(quat, np.float32)
    t = np.ascontiguousarray(trans, np.float32)
    out = np.zeros((height, width, 4), np.float32)
    pairs = C.c_uint64()
    rc = lib().gs_host_render_view(C.c_int(device), _p(g), C.c_uint32(len(g)), _p(q), _p(t), C.c_float(fx), C.c_float(fy),
                                   C.c_float(cx), C.c_float(cy), C.c_uint32(width), C.c_uint32(height),
                                   C.c_int(1 if rgba8_target else 0), _p(out), C.byref(pairs))
    if rc != 0:
        raise GsError("gs_host_render_view: " + lib().gs_host_last_error().decode())
    return out, pairs.value


def ply_export(path, gaussians):
    """PLYExporter::exportPLY (ply_exporter.hpp:17-162)"""
    g = np.ascontiguousarray(gaussians, GAUSSIAN)
    if lib().gs_host_ply_export(str(path).encode(), _p(g), C.c_uint64(len(g))) != 0:
        raise GsError(f"cannot write {path}")


def ply_load(path):
    """load_ply (ply_loader.cpp:54-290) -> (Gaussians, info dict)"""
    n, info = C.c_uint64(), (C.c_uint64 * 4)()
    if lib().gs_host_ply_load(str(path).encode(), None, C.c_uint64(0), C.byref(n), info) != 0:
        raise GsError("load_ply: " + lib().gs_host_last_error().decode())
    out = np.zeros(n.value, GAUSSIAN)
    if lib().gs_host_ply_load(str(path).encode(), _p(out), C.c_uint64(len(out)), C.byref(n), info) != 0:
        raise GsError("load_ply: " + lib().gs_host_last_error().decode())
    return out, dict(vertices=info[0], skipped=info[1], linear_scales=bool(info[2]), has_sh_rest=bool(info[3]))


def checkpoint_resave(in_path, out_path, scene_extent=1.0, device=0):
    if lib().gs_host_checkpoint_resave(C.c_int(device), str(in_path).encode(), str(out_path).encode(), C.c_float(scene_extent)) != 0:
        raise GsError("checkpoint: " + lib().gs_host_last_error().decode())


def ply_device_roundtrip(in_path, out_path, device=0):
    if lib().gs_host_ply_device_roundtrip(C.c_int(device), str(in_path).encode(), str(out_path).encode()) != 0:
        raise GsError("ply roundtrip: " + lib().gs_host_last_error().decode())


def read_checkpoint(path):
    """parse Engine::saveCheckpoint's file (engine.hpp) -> dict"""
    raw = open(path, "rb").read()
    assert raw[:4] == b"GSCK"
    version, = np.frombuffer(raw, "<u4", 1, 4)
    n = int(np.frombuffer(raw, "<u8", 1, 8)[0])
    t = int(np.frombuffer(raw, "<u4", 1, 16)[0])
    off = 24
    g = np.frombuffer(raw, GAUSSIAN, n, off).copy(); off += n * 112
    m, v = [], []
    for w in (3, 3, 4, 1, 12):
        m.append(np.frombuffer(raw, "<f4", n * w, off).copy()); off += n * w * 4
        v.append(np.frombuffer(raw, "<f4", n * w, off).copy()); off += n * w * 4
    accum = np.frombuffer(raw, "<f4", n, off).copy(); off += 4 * n
    views = np.frombuffer(raw, "<u4", n, off).copy(); off += 4 * n
    assert off == len(raw)
    return dict(version=int(version), n=n, timestep=t, gaussians=g, m=m, v=v, accum=accum, views=views)
