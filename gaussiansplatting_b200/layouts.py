"""numpy mirrors of the wire-format structs in include/gs_abi.h.

These are byte-for-byte the reference's host<->kernel structs:
  Gaussian           ply_loader.hpp:14-20            112 B
  GaussianGradients  gradients.hpp:11-31             112 B
  ProjectedGaussian  tiled_rasterizer.hpp:24-39       88 B
  TileRange          tiled_rasterizer.hpp:16-19        8 B
  TiledUniforms      tiled_rasterizer.hpp:42-53      240 B
"""
import numpy as np

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
# viewer draw (gs_viewer_*): Uniforms of mtl_engine.hpp:30-38 (host layout, 240 B) and the per-instance inspection record
VIEW_UNIFORMS = np.dtype([("viewMatrix", "<f4", 16), ("projectionMatrix", "<f4", 16), ("viewProjectionMatrix", "<f4", 16),
                          ("screenSize", "<f4", 2), ("focalLength", "<f4", 2), ("cameraPos", "<f4", 4), ("_pad", "<f4", 4)])
VIEW_SPLAT = np.dtype([("center", "<f4", 2), ("conic", "<f4", 3), ("radius", "<f4"), ("opacity", "<f4"), ("color", "<f4", 3),
                       ("valid", "<u4"), ("tileMinX", "<u4"), ("tileMinY", "<u4"), ("tileMaxX", "<u4"), ("tileMaxY", "<u4"),
                       ("_pad", "<u4")])
assert VIEW_UNIFORMS.itemsize == 240 and VIEW_SPLAT.itemsize == 64


def view_uniforms(u):
    """Uniforms of the viewer draw from a TiledUniforms record: same matrices, screen size, focal length and camera"""
    v = np.zeros(1, VIEW_UNIFORMS)
    for f in ("viewMatrix", "projectionMatrix", "viewProjectionMatrix", "screenSize", "focalLength", "cameraPos"):
        v[f] = u[f]
    return v
