"""ctypes loader for the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: import this from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never from the product package.  PARITY UNPINNED (see
orc_types.h): the reference ships no golden vectors and cannot be built on Linux.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

# numpy mirrors of orc_types.h (byte-for-byte the reference structs)
GAUSSIAN = np.dtype([("position", "<f4", 3), ("_pad0", "<f4"), ("scale", "<f4", 3), ("_pad1", "<f4"),
                     ("rotation", "<f4", 4), ("opacity", "<f4"), ("sh", "<f4", 12), ("_pad2", "<f4", 3)])
GRADIENTS = np.dtype([("position", "<f4", 3), ("opacity", "<f4"), ("scale", "<f4", 3), ("_pad1", "<f4"),
                      ("rotation", "<f4", 4), ("sh", "<f4", 12), ("viewspace", "<f4", 2), ("_pad2", "<f4", 2)])
PROJECTED = np.dtype([("screenPos", "<f4", 2), ("conic", "<f4", 3), ("depth", "<f4"), ("opacity", "<f4"),
                      ("color", "<f4", 3), ("radius", "<f4"), ("tileMinX", "<u4"), ("tileMinY", "<u4"),
                      ("tileMaxX", "<u4"), ("tileMaxY", "<u4"), ("_pad1", "<f4"), ("viewPos_xy", "<f4", 2),
                      ("cov2D", "<f4", 3), ("_pad2", "<f4")])
TILERANGE = np.dtype([("start", "<u4"), ("count", "<u4")])
UNIFORMS = np.dtype([("viewMatrix", "<f4", 16), ("projectionMatrix", "<f4", 16),
                     ("viewProjectionMatrix", "<f4", 16), ("screenSize", "<f4", 2), ("focalLength", "<f4", 2),
                     ("cameraPos", "<f4", 4), ("numTilesX", "<u4"), ("numTilesY", "<u4"),
                     ("numGaussians", "<u4"), ("_pad2", "<u4")])
assert GAUSSIAN.itemsize == 112 and GRADIENTS.itemsize == 112 and PROJECTED.itemsize == 88
assert TILERANGE.itemsize == 8 and UNIFORMS.itemsize == 240
# viewer path (viewer_oracle.cpp): Uniforms of mtl_engine.hpp:30-38 and the per-instance record of the restated vertex stage
VIEW_UNIFORMS = np.dtype([("viewMatrix", "<f4", 16), ("projectionMatrix", "<f4", 16), ("viewProjectionMatrix", "<f4", 16),
                          ("screenSize", "<f4", 2), ("focalLength", "<f4", 2), ("cameraPos", "<f4", 4), ("_pad", "<f4", 4)])
VIEW_SPLAT = np.dtype([("center", "<f4", 2), ("conic", "<f4", 3), ("radius", "<f4"), ("opacity", "<f4"), ("color", "<f4", 3),
                       ("valid", "<u4"), ("tileMinX", "<u4"), ("tileMinY", "<u4"), ("tileMaxX", "<u4"), ("tileMaxY", "<u4"),
                       ("_pad", "<u4")])
assert VIEW_UNIFORMS.itemsize == 240 and VIEW_SPLAT.itemsize == 64


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("gs_oracle.cpp", "viewer_oracle.cpp", "orc_types.h", "orc_math.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    # oracle/_ref (the reference's own CPU code, compiled where /root/reference exists; oracle/ref.py)
    ref_so = os.path.join(_HERE, "_ref", "libref.so")
    if os.path.isdir("/root/reference/GuassianSplatting"):
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref"])
    elif not os.path.exists(ref_so):
        pass        # GPU box without a prebuilt copy: the tests that need it skip
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.orc_generate_pairs.restype = C.c_uint64
        _LIB.orc_densify_apply.restype = C.c_uint32
        _LIB.orc_split_rand_u32.restype = C.c_uint32
        _LIB.orc_lr_decay.restype = C.c_float
        _LIB.orc_test_expf.restype = C.c_float
        _LIB.orc_test_exp_blend.restype = C.c_float
        _LIB.orc_test_expf.argtypes = [C.c_float]
        _LIB.orc_test_exp_blend.argtypes = [C.c_float]
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def max_threads():
    return int(lib().orc_max_threads())


def build_uniforms(quat, trans, fx, fy, cx, cy, width, height, n, near=0.1, far=1000.0):
    u = np.zeros(1, UNIFORMS)
    q = np.asarray(quat, np.float32)
    t = np.asarray(trans, np.float32)
    lib().orc_build_uniforms(_p(q), _p(t), C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                             C.c_uint32(width), C.c_uint32(height), C.c_float(near), C.c_float(far),
                             C.c_uint32(n), _p(u))
    return u


def project(g, u, threads=1):
    out = np.zeros(len(g), PROJECTED)
    lib().orc_project(_p(g), C.c_uint32(len(g)), _p(u), _p(out), C.c_int(threads))
    return out


def generate_pairs(proj, num_tiles_x):
    n = C.c_uint32(len(proj))
    cnt = int(lib().orc_generate_pairs(_p(proj), n, C.c_uint32(num_tiles_x), None, None, C.c_uint64(0)))
    keys = np.zeros(cnt, np.uint64)
    vals = np.zeros(cnt, np.uint32)
    if cnt:
        got = int(lib().orc_generate_pairs(_p(proj), n, C.c_uint32(num_tiles_x), _p(keys), _p(vals),
                                           C.c_uint64(cnt)))
        assert got == cnt
    return keys, vals


def sort_pairs(keys, vals, threads=1):
    k = np.ascontiguousarray(keys, np.uint64).copy()
    v = np.ascontiguousarray(vals, np.uint32).copy()
    lib().orc_sort_pairs(_p(k), _p(v), C.c_uint64(len(k)), C.c_int(threads))
    return k, v


def depth_keys32(positions4, cam):
    p = np.ascontiguousarray(positions4, np.float32)
    n = p.shape[0]
    keys = np.zeros(n, np.uint32)
    vals = np.zeros(n, np.uint32)
    c = np.asarray(cam, np.float32)
    lib().orc_depth_keys32(_p(p), _p(c), C.c_uint32(n), _p(keys), _p(vals))
    return keys, vals


def sort_pairs32(keys, vals):
    k = np.ascontiguousarray(keys, np.uint32).copy()
    v = np.ascontiguousarray(vals, np.uint32).copy()
    lib().orc_sort_pairs32(_p(k), _p(v), C.c_uint64(len(k)))
    return k, v


def tile_ranges(sorted_keys, num_tiles):
    out = np.zeros(num_tiles, TILERANGE)
    k = np.ascontiguousarray(sorted_keys, np.uint64)
    lib().orc_tile_ranges(_p(k), C.c_uint32(len(k)), C.c_uint32(num_tiles), _p(out))
    return out


def forward(proj, sorted_vals, ranges, u, threads=1):
    W, H = int(u["screenSize"][0][0]), int(u["screenSize"][0][1])
    img = np.zeros((H, W, 4), np.float32)
    last = np.zeros((H, W), np.uint32)
    stats = np.zeros(2, np.uint64)
    sv = np.ascontiguousarray(sorted_vals, np.uint32)
    lib().orc_forward(_p(proj), _p(sv), _p(ranges), _p(u), _p(img), _p(last), _p(stats), C.c_int(threads))
    return img, last, (int(stats[0]), int(stats[1]))


def backward(g, proj, sorted_vals, ranges, u, last, rendered, gt, quantize=False, threads=1, want_tfinal=False, want_abs=False):
    """tiledBackward.  want_abs: also return (sum over the (pixel, Gaussian) terms of |term| per component, as a GRADIENTS
    array, and the number of terms per Gaussian) -- the per-component band a float-atomic accumulation in arbitrary order
    can move in, which the GPU gradient tolerance is tied to (oracle/parity.py)"""
    n = len(g)
    grads = np.zeros(n, GRADIENTS)
    W, H = int(u["screenSize"][0][0]), int(u["screenSize"][0][1])
    tf = np.zeros((H, W), np.float32) if want_tfinal else None
    sv = np.ascontiguousarray(sorted_vals, np.uint32)
    r = np.ascontiguousarray(rendered, np.float32)
    t = np.ascontiguousarray(gt, np.float32)
    if want_abs:
        ab, cnt = np.zeros(n, GRADIENTS), np.zeros(n, np.uint32)
        lib().orc_backward_ex(_p(g), _p(proj), _p(sv), _p(ranges), _p(u), _p(last), _p(r), _p(t),
                              C.c_int(int(quantize)), _p(grads), _p(tf), C.c_int(threads), _p(ab), _p(cnt))
        return (grads, tf, ab, cnt) if want_tfinal else (grads, ab, cnt)
    lib().orc_backward(_p(g), _p(proj), _p(sv), _p(ranges), _p(u), _p(last), _p(r), _p(t),
                       C.c_int(int(quantize)), _p(grads), _p(tf), C.c_int(threads))
    return (grads, tf) if want_tfinal else grads


def forward_half(proj, sorted_vals, ranges, u, threads=1):
    """tiledForward with the Metal source's `half` typing emulated in _Float16 (SURVEY note H): (image, lastContribIdx)"""
    W, H = int(u["screenSize"][0][0]), int(u["screenSize"][0][1])
    img = np.zeros((H, W, 4), np.float32)
    last = np.zeros((H, W), np.uint32)
    sv = np.ascontiguousarray(sorted_vals, np.uint32)
    lib().orc_forward_half(_p(proj), _p(sv), _p(ranges), _p(u), _p(img), _p(last), C.c_int(threads))
    return img, last


class AdamState:
    """The reference's 10 SoA moment buffers (optimizer.mm:34-60)."""
    def __init__(self, n):
        z = lambda k: np.zeros(n * k, np.float32)
        self.m = [z(3), z(3), z(4), z(1), z(12)]
        self.v = [z(3), z(3), z(4), z(1), z(12)]
        self.t = 0


def adam_step(g, grads, state, lrs, beta1=0.9, beta2=0.999, eps=1e-8, threads=1):
    state.t += 1
    l = np.asarray(lrs, np.float32)
    lib().orc_adam_step(_p(g), _p(grads), C.c_uint32(len(g)), *[_p(a) for a in state.m],
                        *[_p(a) for a in state.v], _p(l), C.c_float(beta1), C.c_float(beta2),
                        C.c_float(eps), C.c_uint32(state.t), C.c_int(threads))


def accumulate_gradients(grads, accum, counts):
    lib().orc_accumulate_gradients(_p(grads), C.c_uint32(len(grads)), _p(accum), _p(counts))


def densify_apply(g, accum, counts, iteration, scene_extent, focal, image_width, avg_depth, seed,
                  max_gaussians=1500000):
    n = len(g)
    out = np.zeros(max(max_gaussians, 2 * n), GAUSSIAN)
    markers = np.zeros(n, np.uint32)
    stats = np.zeros(3, np.uint32)
    cnt = lib().orc_densify_apply(_p(g), C.c_uint32(n), _p(accum), _p(counts), C.c_uint32(iteration),
                                  C.c_float(scene_extent), C.c_float(focal), C.c_float(image_width),
                                  C.c_float(avg_depth), C.c_uint32(seed), C.c_uint32(max_gaussians),
                                  _p(out), _p(markers), _p(stats))
    return out[:cnt].copy(), markers, stats


def render(g, u, threads=1):
    """project -> pairs -> sort -> ranges -> forward; returns a dict of every intermediate."""
    proj = project(g, u, threads)
    keys, vals = generate_pairs(proj, int(u["numTilesX"][0]))
    skeys, svals = sort_pairs(keys, vals, threads)
    ranges = tile_ranges(skeys, int(u["numTilesX"][0]) * int(u["numTilesY"][0]))
    img, last, stats = forward(proj, svals, ranges, u, threads)
    return dict(proj=proj, keys=keys, vals=vals, skeys=skeys, svals=svals, ranges=ranges,
                image=img, last=last, E=stats[0], C=stats[1])


def loss(rendered, gt, lambda_dssim=0.2, quantize=False):
    """MTLEngine::computeLoss (mtl_engine.mm:769-853): (1-lambda) L1 + lambda D-SSIM, mean over pixels"""
    r = np.ascontiguousarray(rendered, np.float32); g = np.ascontiguousarray(gt, np.float32)
    assert r.shape == g.shape and r.ndim == 3 and r.shape[2] == 4
    f = lib().orc_loss
    f.restype = C.c_float
    return float(f(_p(r), _p(g), C.c_uint32(r.shape[1]), C.c_uint32(r.shape[0]), C.c_float(lambda_dssim), C.c_int(1 if quantize else 0)))


def view_uniforms(u):
    """Uniforms (mtl_engine.hpp:30-38) of the viewer from a TiledUniforms record: same matrices, sizes and camera"""
    v = np.zeros(1, VIEW_UNIFORMS)
    for f in ("viewMatrix", "projectionMatrix", "viewProjectionMatrix", "screenSize", "focalLength", "cameraPos"):
        v[f] = u[f]
    return v


def viewer_order(g, cam):
    """GPURadixSort32::sort (gpu_sort.mm:196-338): back-to-front instance order of the viewer"""
    pos4 = np.zeros((len(g), 4), np.float32)
    pos4[:, :3] = g["position"]
    k, v = depth_keys32(pos4, cam)
    return sort_pairs32(k, v)[1]


def viewer_splats(g, vu, threads=1):
    out = np.zeros(len(g), VIEW_SPLAT)
    lib().orc_viewer_splats(_p(g), C.c_uint32(len(g)), _p(vu), _p(out), C.c_int(threads))
    return out


def viewer_render(g, order, vu, quantize=False, threads=1):
    """MTLEngine::render's draw (mtl_engine.mm:575-605) -> (image HxWx4 float32, fragments shaded, fragments blended)"""
    W, H = int(vu["screenSize"][0][0]), int(vu["screenSize"][0][1])
    img = np.zeros((H, W, 4), np.float32)
    stats = np.zeros(2, np.uint64)
    o = np.ascontiguousarray(order, np.uint32)
    lib().orc_viewer_render(_p(g), C.c_uint32(len(g)), _p(o), _p(vu), C.c_uint32(W), C.c_uint32(H),
                            C.c_int(1 if quantize else 0), _p(img), _p(stats), C.c_int(threads))
    return img, int(stats[0]), int(stats[1])
