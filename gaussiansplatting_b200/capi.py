"""ctypes binding of include/gs_abi.h (libgs_b200.so).

This is the reference-side stub a maintainer would write in Python; the C++ mirror of the reference's host classes
lives in gaussiansplatting_b200/host/.  There is NO CPU fallback: if the CUDA library is missing or a call fails, a
GsError is raised.
"""
import ctypes as C
import os
import numpy as np
from .layouts import GAUSSIAN, GRADIENTS, PROJECTED, TILERANGE, UNIFORMS, VIEW_UNIFORMS, VIEW_SPLAT

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GS_LIB", os.path.join(_HERE, "libgs_b200.so"))
_lib = None

GS_BWD_QUANTIZE_RGBA8 = 1
GS_BWD_ACCUMULATE = 2
GS_VIEW_QUANTIZE_RGBA8 = 1
GS_VIEW_COUNT_FRAGMENTS = 2
(VIEW_SPLATS, VIEW_SORTED_TILES, VIEW_SORTED_VALUES, VIEW_TILE_RANGES) = range(4)
(RAST_PROJECTED, RAST_KEYS, RAST_VALUES, RAST_SORTED_KEYS, RAST_SORTED_VALUES, RAST_TILE_RANGES, RAST_LAST_IDX,
 RAST_TFINAL) = range(8)


class GsError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GsError(f"{LIB_PATH} is missing: run `python -m gaussiansplatting_b200.build` "
                          "(there is no CPU fallback for the CUDA path)")
        L = C.CDLL(LIB_PATH)
        L.gs_last_error.restype = C.c_char_p
        L.gs_last_error.argtypes = [C.c_void_p]
        L.gs_ctx_stream.restype = C.c_void_p
        L.gs_image_devptr.restype = C.c_void_p
        L.gs_rast_pair_count.restype = C.c_uint64
        L.gs_rast_coarse_pair_count.restype = C.c_uint64
        L.gs_viewer_pair_count.restype = C.c_uint64
        for name in ("gs_sort64_sorted_keys", "gs_sort64_sorted_values", "gs_sort32_sorted_keys",
                     "gs_sort32_sorted_indices"):
            getattr(L, name).restype = C.c_void_p
        for name in ("gs_gaussians_count", "gs_gaussians_capacity", "gs_adam_timestep"):
            getattr(L, name).restype = C.c_uint32
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def tune(name, value):
    if lib().gs_tune(name.encode(), C.c_int(value)) != 0:
        raise GsError(f"unknown tuning knob {name}")


class Context:
    def __init__(self, device=0, stream=None):
        self.h = C.c_void_p()
        if stream is None:
            rc = lib().gs_ctx_create(C.c_int(device), C.byref(self.h))
        else:
            rc = lib().gs_ctx_create_on_stream(C.c_int(device), C.c_void_p(stream), C.byref(self.h))
        if rc != 0:
            raise GsError(f"gs_ctx_create failed ({rc}): {lib().gs_last_error(None).decode()}")

    def check(self, rc):
        if rc != 0:
            raise GsError(f"gs call failed ({rc}): {lib().gs_last_error(self.h).decode()}")

    def sync(self):
        self.check(lib().gs_ctx_sync(self.h))

    @property
    def stream(self):
        return lib().gs_ctx_stream(self.h)

    def alloc(self, nbytes):
        p = C.c_void_p()
        self.check(lib().gs_dev_alloc(self.h, C.c_size_t(nbytes), C.byref(p)))
        return p.value

    def free(self, dptr):
        self.check(lib().gs_dev_free(self.h, C.c_void_p(dptr)))

    def upload(self, dptr, arr):
        arr = np.ascontiguousarray(arr)
        self.check(lib().gs_dev_upload(self.h, C.c_void_p(dptr), _p(arr), C.c_size_t(arr.nbytes)))

    def download(self, dptr, shape, dtype):
        out = np.empty(shape, dtype)
        if out.nbytes:
            self.check(lib().gs_dev_download(self.h, _p(out), C.c_void_p(dptr), C.c_size_t(out.nbytes)))
        return out

    def close(self):
        if self.h:
            lib().gs_ctx_destroy(self.h)
            self.h = None


class Gaussians:
    def __init__(self, ctx, capacity):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_gaussians_create(ctx.h, C.c_uint32(capacity), C.byref(self.h)))

    def upload(self, g):
        g = np.ascontiguousarray(g, GAUSSIAN)
        self.ctx.check(lib().gs_gaussians_upload(self.h, _p(g), C.c_uint32(len(g))))

    def download(self, n=None):
        n = self.count if n is None else n
        out = np.zeros(n, GAUSSIAN)
        self.ctx.check(lib().gs_gaussians_download(self.h, _p(out), C.c_uint32(n)))
        return out

    @property
    def count(self):
        return int(lib().gs_gaussians_count(self.h))

    def clamp_opacity(self, max_raw):
        self.ctx.check(lib().gs_gaussians_clamp_opacity(self.h, C.c_float(max_raw)))

    def positions(self, dptr):
        self.ctx.check(lib().gs_gaussians_positions(self.h, C.c_void_p(dptr)))

    def close(self):
        lib().gs_gaussians_destroy(self.h)


class Grads:
    def __init__(self, ctx, capacity, external_ptr=None):
        """external_ptr: device memory the caller owns (16 * capacity floats, capacity a multiple of 64), e.g. a symmetric
        multicast-mapped allocation for the NVLS gradient sum (gs_grads_create_external)"""
        self.ctx = ctx
        self.h = C.c_void_p()
        if external_ptr is None:
            ctx.check(lib().gs_grads_create(ctx.h, C.c_uint32(capacity), C.byref(self.h)))
        else:
            ctx.check(lib().gs_grads_create_external(ctx.h, C.c_uint32(capacity), C.c_void_p(external_ptr), C.byref(self.h)))

    def zero(self, n):
        self.ctx.check(lib().gs_grads_zero(self.h, C.c_uint32(n)))

    def upload(self, g):
        g = np.ascontiguousarray(g, GRADIENTS)
        self.ctx.check(lib().gs_grads_upload(self.h, _p(g), C.c_uint32(len(g))))

    def download(self, n):
        out = np.zeros(n, GRADIENTS)
        self.ctx.check(lib().gs_grads_download(self.h, _p(out), C.c_uint32(n)))
        return out

    def set_stat_mode(self, per_view):
        self.ctx.check(lib().gs_grads_set_stat_mode(self.h, C.c_int(int(per_view))))

    def packed(self):
        """(device pointer, number of floats, plane stride) of the packed fp32 buffer (all-reduce payload)."""
        p, n, s = C.c_void_p(), C.c_uint64(), C.c_uint32()
        self.ctx.check(lib().gs_grads_packed(self.h, C.byref(p), C.byref(n), C.byref(s)))
        return p.value, int(n.value), int(s.value)

    def close(self):
        lib().gs_grads_destroy(self.h)


def loss(ctx, rendered, ground_truth, lambda_dssim=0.2, quantize=False):
    """MTLEngine::computeLoss: (1 - lambda) * L1 + lambda * D-SSIM, mean over pixels (synchronises)"""
    out = C.c_float(0.0)
    ctx.check(lib().gs_loss_compute(ctx.h, rendered.h, ground_truth.h, C.c_float(lambda_dssim), C.c_int(1 if quantize else 0), C.byref(out)))
    return float(out.value)


def loss_noalloc(ctx, rendered, ground_truth, d_out, lambda_dssim=0.2, quantize=False):
    """gs_loss_compute_noalloc: the same loss with static scratch and no allocation; enqueues only, the mean loss is written to
    the caller's DEVICE float at d_out (no counterpart in the reference; see include/gs_abi.h)"""
    ctx.check(lib().gs_loss_compute_noalloc(ctx.h, rendered.h, ground_truth.h, C.c_float(lambda_dssim), C.c_int(1 if quantize else 0),
                                            C.c_void_p(d_out)))


class Image:
    def __init__(self, ctx, width, height):
        self.ctx, self.w, self.hgt = ctx, width, height
        self.h = C.c_void_p()
        ctx.check(lib().gs_image_create(ctx.h, C.c_uint32(width), C.c_uint32(height), C.byref(self.h)))

    def upload(self, img):
        img = np.ascontiguousarray(img, np.float32)
        assert img.shape == (self.hgt, self.w, 4)
        self.ctx.check(lib().gs_image_upload_rgba32f(self.h, _p(img)))
        self.ctx.sync()

    def upload_rgba8_ptr(self, host_ptr):
        """enqueue an H2D of RGBA8 pixels from (pinned) host memory; does not synchronise"""
        self.ctx.check(lib().gs_image_upload_rgba8(self.h, C.c_void_p(host_ptr)))

    def upload_rgba8(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        self.ctx.check(lib().gs_image_upload_rgba8(self.h, _p(img)))
        self.ctx.sync()

    def download(self):
        out = np.empty((self.hgt, self.w, 4), np.float32)
        self.ctx.check(lib().gs_image_download_rgba32f(self.h, _p(out)))
        return out

    def download_rgba8(self, out=None):
        if out is None:
            out = np.empty((self.hgt, self.w, 4), np.uint8)
        self.ctx.check(lib().gs_image_download_rgba8(self.h, _p(out)))
        return out

    def download_rgba8_ptr(self, host_ptr):
        self.ctx.check(lib().gs_image_download_rgba8(self.h, C.c_void_p(host_ptr)))

    def upload_rgba8_async(self, host_ptr):
        """H2D of RGBA8 pixels from PINNED host memory on the context's copy-in stream; compute calls that use the image
        wait for it on the device"""
        self.ctx.check(lib().gs_image_upload_rgba8_async(self.h, C.c_void_p(host_ptr)))

    def download_rgba8_async(self, host_ptr):
        """D2H of the image as RGBA8 into PINNED host memory on the copy-out stream; wait() before reading the buffer"""
        self.ctx.check(lib().gs_image_download_rgba8_async(self.h, C.c_void_p(host_ptr)))

    def wait(self):
        self.ctx.check(lib().gs_image_wait(self.h))

    @property
    def devptr(self):
        return lib().gs_image_devptr(self.h)

    def close(self):
        lib().gs_image_destroy(self.h)


class Rasterizer:
    """TiledRasterizer (tiled_rasterizer.hpp:56-124)."""
    def __init__(self, ctx, max_gaussians):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_rast_create(ctx.h, C.c_uint32(max_gaussians), C.byref(self.h)))

    def forward(self, gaussians, uniforms, out_image):
        self.ctx.check(lib().gs_rast_forward(self.h, gaussians.h, _p(uniforms), out_image.h))
        self._shape = (out_image.hgt, out_image.w)
        self._n = gaussians.count
        self._tiles = ((out_image.w + 15) // 16) * ((out_image.hgt + 15) // 16)

    def backward(self, gaussians, grads, uniforms, rendered, gt, flags=0):
        self.ctx.check(lib().gs_rast_backward(self.h, gaussians.h, grads.h, _p(uniforms), rendered.h, gt.h, C.c_int(flags)))

    def set_profiling(self, on=True):
        self.ctx.check(lib().gs_rast_set_profiling(self.h, C.c_int(int(on))))

    STAGES = ("preprocess", "depth_sort", "scan", "emit", "sort", "ranges", "blend_fwd", "blend_bwd", "preprocess_bwd")

    def timings(self, include_backward=True):
        ms = (C.c_float * 9)()
        self.ctx.check(lib().gs_rast_get_timings(self.h, ms, C.c_int(int(include_backward))))
        return dict(zip(self.STAGES, [float(x) for x in ms]))

    def count_evals(self):
        e, c = C.c_uint64(), C.c_uint64()
        self.ctx.check(lib().gs_rast_count_evals(self.h, C.byref(e), C.byref(c)))
        return int(e.value), int(c.value)

    @property
    def pair_count(self):
        return int(lib().gs_rast_pair_count(self.h))

    @property
    def coarse_pair_count(self):
        return int(lib().gs_rast_coarse_pair_count(self.h))

    @property
    def sort_passes(self):
        return int(lib().gs_rast_sort_passes(self.h))

    def set_overlap(self, on=True):
        """the ordering stage on a high-priority side stream (see gs_rast_set_overlap: for alternating two rasterizers)"""
        self.ctx.check(lib().gs_rast_set_overlap(self.h, C.c_int(1 if on else 0)))

    @property
    def backward_used_visits(self):
        return bool(lib().gs_rast_backward_used_visits(self.h))

    def get(self, which):
        P = self.pair_count
        shape, dt = {RAST_PROJECTED: (self._n, PROJECTED), RAST_KEYS: (P, np.uint64), RAST_VALUES: (P, np.uint32),
                     RAST_SORTED_KEYS: (P, np.uint64), RAST_SORTED_VALUES: (P, np.uint32),
                     RAST_TILE_RANGES: (self._tiles, TILERANGE), RAST_LAST_IDX: (self._shape, np.uint32),
                     RAST_TFINAL: (self._shape, np.float32)}[which]
        out = np.zeros(shape, dt)
        self.ctx.check(lib().gs_rast_get(self.h, C.c_int(which), _p(out), C.c_size_t(out.nbytes)))
        return out

    def close(self):
        lib().gs_rast_destroy(self.h)


class Sort64:
    """GPURadixSort64 (gpu_sort.hpp:64-107)."""
    def __init__(self, ctx, max_elements):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_sort64_create(ctx.h, C.c_uint64(max_elements), C.byref(self.h)))

    def sort(self, d_keys, d_vals, n, begin_bit=0, end_bit=64):
        self.ctx.check(lib().gs_sort64_sort(self.h, C.c_void_p(d_keys), C.c_void_p(d_vals), C.c_uint64(n),
                                            C.c_int(begin_bit), C.c_int(end_bit)))

    @property
    def sorted_keys(self):
        return lib().gs_sort64_sorted_keys(self.h)

    @property
    def sorted_values(self):
        return lib().gs_sort64_sorted_values(self.h)

    @property
    def last_plan_hybrid(self):
        """True if the last sort completed on the prefix-passes + local-pass plan, False on the 8-pass LSD plan"""
        h = C.c_int()
        self.ctx.check(lib().gs_sort64_last_plan(self.h, C.byref(h)))
        return bool(h.value)

    def sort_host(self, keys, vals, begin_bit=0, end_bit=64):
        """convenience for tests: upload, sort, download"""
        keys = np.ascontiguousarray(keys, np.uint64)
        vals = np.ascontiguousarray(vals, np.uint32)
        n = len(keys)
        dk, dv = self.ctx.alloc(max(n, 1) * 8), self.ctx.alloc(max(n, 1) * 4)
        try:
            if n:
                self.ctx.upload(dk, keys)
                self.ctx.upload(dv, vals)
            self.sort(dk, dv, n, begin_bit, end_bit)
            self.ctx.sync()
            return (self.ctx.download(self.sorted_keys, n, np.uint64), self.ctx.download(self.sorted_values, n, np.uint32))
        finally:
            self.ctx.free(dk)
            self.ctx.free(dv)

    def close(self):
        lib().gs_sort64_destroy(self.h)


class Sort32:
    """GPURadixSort32 (gpu_sort.hpp:14-62)."""
    def __init__(self, ctx, max_elements):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_sort32_create(ctx.h, C.c_uint64(max_elements), C.byref(self.h)))

    def sort_depth(self, d_positions4, cam, n):
        cam = np.asarray(cam, np.float32)
        self.ctx.check(lib().gs_sort32_sort_depth(self.h, C.c_void_p(d_positions4), _p(cam), C.c_uint64(n)))

    def sort(self, d_keys, d_vals, n):
        self.ctx.check(lib().gs_sort32_sort(self.h, C.c_void_p(d_keys), C.c_void_p(d_vals), C.c_uint64(n)))

    @property
    def sorted_keys(self):
        return lib().gs_sort32_sorted_keys(self.h)

    @property
    def sorted_indices(self):
        return lib().gs_sort32_sorted_indices(self.h)

    def close(self):
        lib().gs_sort32_destroy(self.h)


class PeerReduce:
    """gs_dp_*: the per-step gradient sum of view-parallel training over NVLink peer memory (csrc/peer_reduce.cu).
    usage on every rank: pr = PeerReduce(ctx, rank, world); blob = pr.export(grads); pr.connect(all blobs in rank order);
    then pr.allreduce(grads, n) after the step's backward calls."""
    BLOB_BYTES = 192

    def __init__(self, ctx, rank, world):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_dp_create(ctx.h, C.c_int(rank), C.c_int(world), C.byref(self.h)))

    def export(self, grads):
        blob = (C.c_ubyte * self.BLOB_BYTES)()
        self.ctx.check(lib().gs_dp_export(self.h, grads.h, blob))
        return bytes(blob)

    def connect(self, blobs):
        raw = b"".join(blobs)
        buf = (C.c_ubyte * len(raw)).from_buffer_copy(raw)
        self.ctx.check(lib().gs_dp_connect(self.h, buf))

    def connect_ptrs(self, grads, peer_bufs, peer_flags, multicast_buf=None):
        """gs_dp_connect_ptrs: pointers from a symmetric-memory rendezvous; multicast_buf != None -> in-switch (NVLS) sum"""
        w = len(peer_bufs)
        pb = (C.c_void_p * w)(*[C.c_void_p(int(p)) for p in peer_bufs])
        pf = (C.c_void_p * w)(*[C.c_void_p(int(p)) for p in peer_flags])
        self.ctx.check(lib().gs_dp_connect_ptrs(self.h, grads.h, pb, pf, C.c_void_p(int(multicast_buf)) if multicast_buf else None))

    def allreduce(self, grads, n):
        self.ctx.check(lib().gs_dp_allreduce_grads(self.h, grads.h, C.c_uint32(n)))

    def check(self):
        self.ctx.check(lib().gs_dp_check(self.h))

    def close(self):
        lib().gs_dp_destroy(self.h)


class Viewer:
    """The draw of MTLEngine::render (mtl_engine.mm:401-609; vertexShader / fragmentShader, shaders.metal:160-317) as CUDA
    compute into a headless Image."""
    def __init__(self, ctx, max_gaussians):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_viewer_create(ctx.h, C.c_uint32(max_gaussians), C.byref(self.h)))

    def render(self, gaussians, d_sorted_indices, view_uniforms, out_image, flags=0):
        u = np.ascontiguousarray(view_uniforms, VIEW_UNIFORMS)
        self.ctx.check(lib().gs_viewer_render(self.h, gaussians.h, C.c_void_p(d_sorted_indices), _p(u), out_image.h, C.c_int(flags)))

    @property
    def pair_count(self):
        return int(lib().gs_viewer_pair_count(self.h))

    def fragments(self):
        a, b = C.c_uint64(), C.c_uint64()
        self.ctx.check(lib().gs_viewer_fragments(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def get(self, which, n=None, tiles=None):
        if which == VIEW_SPLATS:
            out = np.zeros(n, VIEW_SPLAT)
        elif which == VIEW_TILE_RANGES:
            out = np.zeros(tiles, TILERANGE)
        else:
            out = np.zeros(self.pair_count, np.uint32)
        self.ctx.check(lib().gs_viewer_get(self.h, C.c_int(which), _p(out), C.c_size_t(out.nbytes)))
        return out

    def close(self):
        lib().gs_viewer_destroy(self.h)


class Adam:
    """AdamOptimizer (optimizer.hpp:23-95)."""
    WIDTH = (3, 3, 4, 1, 12)

    def __init__(self, ctx, num_gaussians):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_adam_create(ctx.h, C.c_uint32(num_gaussians), C.byref(self.h)))

    def step(self, gaussians, grads, lrs):
        l = np.asarray(lrs, np.float32)
        self.ctx.check(lib().gs_adam_step(self.h, gaussians.h, grads.h, _p(l)))

    def reset(self):
        self.ctx.check(lib().gs_adam_reset(self.h))

    def resize_if_needed(self, n):
        self.ctx.check(lib().gs_adam_resize_if_needed(self.h, C.c_uint32(n)))

    def reset_opacity_momentum(self):
        self.ctx.check(lib().gs_adam_reset_opacity_momentum(self.h))

    def reset_scale_momentum(self):
        self.ctx.check(lib().gs_adam_reset_scale_momentum(self.h))

    def permute_after_densify(self, density, gaussians, old_n):
        """opt-in correction of SURVEY note D: moments follow the Gaussians through Density.apply"""
        self.ctx.check(lib().gs_adam_permute_after_densify(self.h, density.h, gaussians.h, C.c_uint32(old_n)))

    def reset_state_for_new(self, start):
        self.ctx.check(lib().gs_adam_reset_state_for_new(self.h, C.c_uint32(start)))

    @property
    def timestep(self):
        return int(lib().gs_adam_timestep(self.h))

    def download(self, group, which, n):
        out = np.zeros(n * self.WIDTH[group], np.float32)
        self.ctx.check(lib().gs_adam_download(self.h, C.c_int(group), C.c_int(which), _p(out), C.c_uint32(n)))
        return out

    def upload(self, group, which, arr, n):
        arr = np.ascontiguousarray(arr, np.float32)
        self.ctx.check(lib().gs_adam_upload(self.h, C.c_int(group), C.c_int(which), _p(arr), C.c_uint32(n)))

    def close(self):
        lib().gs_adam_destroy(self.h)


class Density:
    """DensityController (density_control.hpp:20-62)."""
    def __init__(self, ctx, max_gaussians=1500000):
        self.ctx = ctx
        self.h = C.c_void_p()
        ctx.check(lib().gs_dens_create(ctx.h, C.c_uint32(max_gaussians), C.byref(self.h)))

    def set_scene_extent(self, e):
        self.ctx.check(lib().gs_dens_set_scene_extent(self.h, C.c_float(e)))

    def set_seed(self, s):
        self.ctx.check(lib().gs_dens_set_seed(self.h, C.c_uint32(s)))

    def accumulate(self, grads, n):
        self.ctx.check(lib().gs_dens_accumulate(self.h, grads.h, C.c_uint32(n)))

    def reset_accumulator(self, n):
        self.ctx.check(lib().gs_dens_reset_accumulator(self.h, C.c_uint32(n)))

    def apply(self, gaussians, iteration, focal, image_width, avg_depth):
        st = (C.c_uint32 * 3)()
        self.ctx.check(lib().gs_dens_apply(self.h, gaussians.h, C.c_uint32(iteration), C.c_float(focal),
                                           C.c_float(image_width), C.c_float(avg_depth), st))
        return tuple(int(x) for x in st)

    def download(self, n, markers=False):
        a, c = np.zeros(n, np.float32), np.zeros(n, np.uint32)
        m = np.zeros(n, np.uint32) if markers else None
        self.ctx.check(lib().gs_dens_download(self.h, _p(a), _p(c), _p(m), C.c_uint32(n)))
        return (a, c, m) if markers else (a, c)

    def upload(self, accum, count):
        a = np.ascontiguousarray(accum, np.float32)
        c = np.ascontiguousarray(count, np.uint32)
        self.ctx.check(lib().gs_dens_upload(self.h, _p(a), _p(c), C.c_uint32(len(a))))

    def close(self):
        lib().gs_dens_destroy(self.h)
