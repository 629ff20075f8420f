/*
 * orc_types.h — POD layouts used by the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  PARITY UNPINNED: the reference (ctaylo41/GaussianSplatting) ships no
 * golden vectors, tests or fixtures for this path, and its Metal/Objective-C++ sources
 * cannot be compiled on Linux, so this restatement is pinned only by the self-checks in
 * tests/ (hand-computed cases, invariants, finite differences).
 *
 * Layouts restate (byte for byte) the reference structs:
 *   Gaussian           ply_loader.hpp:14-20, tiled_shaders.metal:11-22      112 B
 *   GaussianGradients  gradients.hpp:11-31,  tiled_shaders.metal:65-80      112 B
 *   ProjectedGaussian  tiled_rasterizer.hpp:24-39, tiled_shaders.metal:27-42 88 B
 *   TileRange          tiled_rasterizer.hpp:16-19                             8 B
 *   TiledUniforms      tiled_rasterizer.hpp:42-53, tiled_shaders.metal:51-62 240 B
 */
#ifndef ORC_TYPES_H
#define ORC_TYPES_H
#include <stdint.h>
#include <stddef.h>

typedef struct {
    float position[3]; float _pad0;
    float scale[3];    float _pad1;      /* LOG scale */
    float rotation[4];                   /* (w,x,y,z) */
    float opacity;                       /* raw, pre-sigmoid */
    float sh[12];                        /* R 0-3, G 4-7, B 8-11; DC at 0,4,8 */
    float _pad2[3];
} OrcGaussian;

typedef struct {
    float position_x, position_y, position_z;
    float opacity;
    float scale_x, scale_y, scale_z;
    float _pad1;
    float rotation[4];
    float sh[12];
    float viewspace_grad_x, viewspace_grad_y;
    float _pad2, _pad3;
} OrcGradients;

typedef struct {
    float screenPos[2];
    float conic[3];
    float depth;
    float opacity;
    float color[3];
    float radius;
    uint32_t tileMinX, tileMinY, tileMaxX, tileMaxY;
    float _pad1;
    float viewPos_xy[2];
    float cov2D[3];
    float _pad2;
} OrcProjected;

typedef struct { uint32_t start, count; } OrcTileRange;

typedef struct {
    float viewMatrix[16];            /* column-major: m[c*4+r] */
    float projectionMatrix[16];
    float viewProjectionMatrix[16];
    float screenSize[2];
    float focalLength[2];
    float cameraPos[4];              /* float3 + pad */
    uint32_t numTilesX, numTilesY, numGaussians, _pad2;
} OrcUniforms;

#ifdef __cplusplus
static_assert(sizeof(OrcGaussian) == 112, "Gaussian layout");
static_assert(offsetof(OrcGaussian, scale) == 16 && offsetof(OrcGaussian, rotation) == 32 &&
              offsetof(OrcGaussian, opacity) == 48 && offsetof(OrcGaussian, sh) == 52, "Gaussian offsets");
static_assert(sizeof(OrcGradients) == 112, "GaussianGradients layout");
static_assert(offsetof(OrcGradients, rotation) == 32 && offsetof(OrcGradients, sh) == 48 &&
              offsetof(OrcGradients, viewspace_grad_x) == 96, "GaussianGradients offsets");
static_assert(sizeof(OrcProjected) == 88, "ProjectedGaussian layout");
static_assert(offsetof(OrcProjected, conic) == 8 && offsetof(OrcProjected, depth) == 20 &&
              offsetof(OrcProjected, color) == 28 && offsetof(OrcProjected, radius) == 40 &&
              offsetof(OrcProjected, tileMinX) == 44 && offsetof(OrcProjected, viewPos_xy) == 64 &&
              offsetof(OrcProjected, cov2D) == 72, "ProjectedGaussian offsets");
static_assert(sizeof(OrcTileRange) == 8, "TileRange layout");
static_assert(sizeof(OrcUniforms) == 240, "TiledUniforms layout");
static_assert(offsetof(OrcUniforms, screenSize) == 192 && offsetof(OrcUniforms, cameraPos) == 208 &&
              offsetof(OrcUniforms, numTilesX) == 224, "TiledUniforms offsets");
#endif
#endif
