"""ctypes loader for oracle/_ref/libref.so: the CPU halves of the reference's hot path compiled from the reference's OWN
sources (oracle/Makefile lists exactly which lines) — parallelRadixSort, DensityController::accumulateGradients / apply,
viewMatrixFromColmap / projectionFromColmap, exponentialLRDecay, load_ply / PLYExporter::exportPLY, and the wire-struct
layouts as the reference's headers give them.  TEST INFRASTRUCTURE ONLY: it pins the oracle (and, in the GPU tests, the
CUDA path) to reference object code for everything that is not Metal.  Built only where /root/reference exists; the
prebuilt library travels to the GPU box.
"""
import ctypes as C
import os
import subprocess
import numpy as np
from .orc import GAUSSIAN, GRADIENTS, _p

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libref.so")
REF_SRC = "/root/reference/GuassianSplatting"
_LIB = None


def available():
    return os.path.exists(SO) or os.path.isdir(REF_SRC)


def build():
    if os.path.isdir(REF_SRC):
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref"])
    return SO


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(SO):
            build()
        _LIB = C.CDLL(SO)
        _LIB.ref_densify_apply.restype = C.c_uint32
        _LIB.ref_lr_decay.restype = C.c_float
        _LIB.ref_ply_load.restype = C.c_int64
    return _LIB


LAYOUT_NAMES = ("Gaussian", "Gaussian.scale", "Gaussian.rotation", "Gaussian.opacity", "Gaussian.sh",
                "GaussianGradients", "GaussianGradients.rotation", "GaussianGradients.sh", "GaussianGradients.viewspace_grad_x",
                "ProjectedGaussian", "ProjectedGaussian.conic", "ProjectedGaussian.depth", "ProjectedGaussian.color",
                "ProjectedGaussian.radius", "ProjectedGaussian.tileMinX", "ProjectedGaussian.viewPos_xy", "ProjectedGaussian.cov2D",
                "TileRange", "TiledUniforms", "TiledUniforms.screenSize", "TiledUniforms.cameraPos", "TiledUniforms.numTilesX")


def layouts():
    """sizeof / offsetof of the wire structs as compiled from the reference's headers"""
    out = np.zeros(len(LAYOUT_NAMES), np.uint32)
    lib().ref_layouts(_p(out))
    return dict(zip(LAYOUT_NAMES, [int(x) for x in out]))


def sort_pairs(keys, vals):
    """parallelRadixSort (tiled_rasterizer.mm:27-102)"""
    k = np.ascontiguousarray(keys, np.uint64).copy()
    v = np.ascontiguousarray(vals, np.uint32).copy()
    lib().ref_parallel_radix_sort(_p(k), _p(v), C.c_uint64(len(k)))
    return k, v


def accumulate_gradients(grads, accum, counts):
    """DensityController::accumulateGradients (density_control.mm:121-185); accum / counts updated in place"""
    g = np.ascontiguousarray(grads, GRADIENTS)
    lib().ref_accumulate_gradients(_p(g), C.c_uint32(len(g)), _p(accum), _p(counts))


def densify_apply(g, accum, counts, iteration, scene_extent, focal, image_width, avg_depth, seed):
    """DensityController::apply (density_control.mm:188-501) after srand(seed): (new Gaussians, markers, stats)"""
    g = np.ascontiguousarray(g, GAUSSIAN)
    n = len(g)
    cap = 2 * n + 16
    out = np.zeros(cap, GAUSSIAN)
    markers, stats = np.zeros(n, np.uint32), np.zeros(3, np.uint32)
    a, c = np.ascontiguousarray(accum, np.float32), np.ascontiguousarray(counts, np.uint32)
    cnt = lib().ref_densify_apply(_p(g), C.c_uint32(n), _p(a), _p(c), C.c_uint32(iteration), C.c_float(scene_extent),
                                  C.c_float(focal), C.c_float(image_width), C.c_float(avg_depth), C.c_uint32(seed),
                                  _p(out), C.c_uint32(cap), _p(markers), _p(stats))
    return out[:cnt].copy(), markers, stats


def view_matrix(quat, trans):
    out = np.zeros(16, np.float32)
    lib().ref_view_matrix(_p(np.asarray(quat, np.float32)), _p(np.asarray(trans, np.float32)), _p(out))
    return out


def projection(fx, fy, cx, cy, w, h, near=0.1, far=1000.0):
    out = np.zeros(16, np.float32)
    lib().ref_projection(C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), C.c_uint32(w), C.c_uint32(h),
                         C.c_float(near), C.c_float(far), _p(out))
    return out


def lr_decay(lr_init, lr_final, it, max_it):
    return float(lib().ref_lr_decay(C.c_float(lr_init), C.c_float(lr_final), C.c_uint64(it), C.c_uint64(max_it)))


def ply_export(path, g):
    g = np.ascontiguousarray(g, GAUSSIAN)
    if lib().ref_ply_export(str(path).encode(), _p(g), C.c_uint64(len(g))) != 0:
        raise RuntimeError("reference exportPLY failed")


def ply_load(path):
    n = lib().ref_ply_load(str(path).encode(), None, C.c_uint64(0))
    if n < 0:
        raise RuntimeError("reference load_ply threw")
    out = np.zeros(n, GAUSSIAN)
    lib().ref_ply_load(str(path).encode(), _p(out), C.c_uint64(n))
    return out
