"""Synthetic scenes and COLMAP-convention cameras, shared by tests and bench.

All randomness is a counter-based SplitMix64 stream (never rand()/numpy Generator), so a
(seed, N) pair produces the same bytes on every machine and numpy version.

Camera maths follows the reference's caller (mtl_engine.mm:637-682 viewMatrixFromColmap /
projectionFromColmap, :912-924 uniform assembly).
"""
import numpy as np
from .layouts import GAUSSIAN, UNIFORMS

SH_C0 = 0.28209479177387814
TILE = 16

# per-config base scale s0, calibrated ONCE with the oracle so that the mean number of tiles per visible
# Gaussian is ~7-9 (the reference's sizing assumption AVG_TILES_PER_GAUSSIAN = 8, tiled_rasterizer.hpp:80) and frozen:
#   C1 256^2  orbit 3: s0 0.040 -> 6.7 tiles      C2 1080p orbit 3: s0 0.006 -> 7.9 tiles
#   C4 4K     orbit 3: s0 0.006 -> 20.8 tiles (log-scale is clamped at -5, tiled_shaders.metal:87,160, so no s0 goes lower)
#   C3/C5 (training) use s0 0.03 seen from orbit 12: Adam clamps log-scale to +-4 (shaders.metal:55,655), so a trainable
#   scene cannot hold scales below e^-4 = 0.018.
CONFIG_S0 = {"C1": 0.040, "C2": 0.006, "C3": 0.030, "C4": 0.006, "C5": 0.030}
CONFIG_ORBIT = {"C1": 3.0, "C2": 3.0, "C3": 12.0, "C4": 3.0, "C5": 12.0}
CONFIG_SEED = {"C1": 0x5EED0001, "C2": 0x5EED0002, "C3": 0x5EED0003, "C4": 0x5EED0004, "C5": 0x5EED0005}
CONFIG_N = {"C1": 10000, "C2": 1000000, "C3": 500000, "C4": 3000000, "C5": 1000000}
CONFIG_RES = {"C1": (256, 256), "C2": (1920, 1080), "C3": (1920, 1080), "C4": (3840, 2160), "C5": (1920, 1080)}


def splitmix64(seed, n, stream=0):
    """n uint64 values: SplitMix64 of counter (seed, stream, i)."""
    with np.errstate(over="ignore"):
        z = (np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + np.uint64(stream) * np.uint64(0xA0761D6478BD642F)
             + (np.arange(n, dtype=np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def uniform01(seed, n, stream):
    """float64 uniforms in [0,1) with 53 random bits."""
    return (splitmix64(seed, n, stream) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def normal(seed, n, stream):
    u1 = uniform01(seed, n, stream * 2 + 1000)
    u2 = uniform01(seed, n, stream * 2 + 1001)
    return np.sqrt(-2.0 * np.log(1.0 - u1)) * np.cos(2.0 * np.pi * u2)


def make_scene(n, seed, s0=0.010):
    """synth(N, seed): positions U[-1,1]^3, log-scale ln(s0)+U(-.5,.5), unit quats, raw opacity U(-2,4),
    SH-DC from U(.05,.95) colours, SH rest 0."""
    g = np.zeros(n, GAUSSIAN)
    for k in range(3):
        g["position"][:, k] = (uniform01(seed, n, k) * 2.0 - 1.0).astype(np.float32)
        g["scale"][:, k] = (np.log(s0) + uniform01(seed, n, 3 + k) - 0.5).astype(np.float32)
    q = np.stack([normal(seed, n, k) for k in range(4)], axis=1)
    q /= np.maximum(np.linalg.norm(q, axis=1, keepdims=True), 1e-12)
    g["rotation"] = q.astype(np.float32)
    g["opacity"] = (uniform01(seed, n, 6) * 6.0 - 2.0).astype(np.float32)
    for c in range(3):
        col = 0.05 + 0.9 * uniform01(seed, n, 7 + c)
        g["sh"][:, 4 * c] = ((col - 0.5) / SH_C0).astype(np.float32)
    return g


def _quat_from_rot(R):
    """(w,x,y,z) from a 3x3 rotation matrix (float64)."""
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        w, x, y, z = 0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = np.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        w, x, y, z = (R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s
    elif R[1, 1] > R[2, 2]:
        s = np.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        w, x, y, z = (R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s
    else:
        s = np.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        w, x, y, z = (R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s
    q = np.array([w, x, y, z])
    return q / np.linalg.norm(q)


def orbit_pose(i, num_views, radius=3.0):
    """COLMAP/OpenCV pose (quat wxyz world->camera, translation) of view i: eye on a sphere of the given
    radius at azimuth 2*pi*i/V, elevation 15deg*sin(i), looking at the origin; +z forward, y down."""
    az = 2.0 * np.pi * i / max(num_views, 1)
    el = np.deg2rad(15.0) * np.sin(float(i))
    eye = radius * np.array([np.cos(el) * np.cos(az), np.sin(el), np.cos(el) * np.sin(az)])
    zc = -eye / np.linalg.norm(eye)
    xc = np.cross(np.array([0.0, 1.0, 0.0]), zc)
    xc /= np.linalg.norm(xc)
    yc = np.cross(zc, xc)
    R = np.stack([xc, yc, zc], axis=0)          # rows: camera axes in world coords
    t = -R @ eye
    return _quat_from_rot(R).astype(np.float32), t.astype(np.float32)


def quat_to_mat32(q):
    """tiled_shaders.metal:91-99 / mtl_engine.mm:642-649, float32; returns M[row][col]."""
    w, x, y, z = [np.float32(v) for v in q]
    one, two = np.float32(1), np.float32(2)
    c0 = [one - two * (y * y + z * z), two * (x * y + w * z), two * (x * z - w * y)]
    c1 = [two * (x * y - w * z), one - two * (x * x + z * z), two * (y * z + w * x)]
    c2 = [two * (x * z + w * y), two * (y * z - w * x), one - two * (x * x + y * y)]
    return np.array([c0, c1, c2], np.float32).T


def build_uniforms(quat, trans, fx, fy, cx, cy, width, height, n, near=0.1, far=1000.0):
    """TiledUniforms exactly as trainStep assembles them (mtl_engine.mm:912-924)."""
    f32 = np.float32
    u = np.zeros(1, UNIFORMS)
    R = quat_to_mat32(quat)
    V = np.zeros((4, 4), f32)                    # V[col][row]
    for c in range(3):
        V[c, :3] = R[:, c]
    V[3, :3] = np.asarray(trans, f32)
    V[3, 3] = 1
    P = np.zeros((4, 4), f32)
    w, h, nz, fz = f32(width), f32(height), f32(near), f32(far)
    P[0, 0] = f32(2) * f32(fx) / w
    P[1, 1] = f32(2) * f32(fy) / h
    P[2, 0] = f32(2) * f32(cx) / w - f32(1)
    P[2, 1] = f32(2) * f32(cy) / h - f32(1)
    P[2, 2] = fz / (fz - nz)
    P[2, 3] = 1
    P[3, 2] = -(fz * nz) / (fz - nz)
    VP = np.zeros((4, 4), f32)
    for c in range(4):
        for r in range(4):
            s = f32(0)
            for k in range(4):
                s = f32(s + P[k, r] * V[c, k])
            VP[c, r] = s
    u["viewMatrix"][0] = V.reshape(16)
    u["projectionMatrix"][0] = P.reshape(16)
    u["viewProjectionMatrix"][0] = VP.reshape(16)
    u["screenSize"][0] = [width, height]
    u["focalLength"][0] = [fx, fy]
    t = np.asarray(trans, f32)
    cam = [-f32(f32(f32(R[0, r] * t[0]) + f32(R[1, r] * t[1])) + f32(R[2, r] * t[2])) for r in range(3)]
    u["cameraPos"][0][:3] = cam
    u["numTilesX"] = (width + TILE - 1) // TILE
    u["numTilesY"] = (height + TILE - 1) // TILE
    u["numGaussians"] = n
    return u


def make_camera(i, num_views, width, height, n, radius=3.0):
    """cam(i) of SURVEY §8d: fx = fy = 0.9 W, principal point at the centre, near .1, far 1000."""
    q, t = orbit_pose(i, num_views, radius)
    f = 0.9 * width
    return build_uniforms(q, t, f, f, width / 2.0, height / 2.0, width, height, n)


def scene_extent(radius=3.0):
    """computeSceneExtent would give 1.1 * max camera distance from the centroid (colmap_loader.cpp:232-264)."""
    return 1.1 * radius
