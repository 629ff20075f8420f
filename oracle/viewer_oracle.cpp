/*
 * viewer_oracle.cpp — scalar CPU restatement of the reference's interactive splat renderer.
 *
 * TEST INFRASTRUCTURE ONLY (see orc_types.h).  PARITY UNPINNED: the reference draws this through Metal's raster
 * pipeline (instanced triangle strips + fixed-function blending), which cannot run here (no graphics stack, no Apple GPU)
 * and for which it ships no golden images; tests/ pins this file with hand-computed cases and an independent float64
 * numpy statement.
 *
 * What is restated (paths relative to /root/reference/GuassianSplatting/):
 *   vertexShader      shaders.metal:160-288   per-instance: validity, clip position, cov3D (+-8 log-scale clamp, raw quaternion),
 *                                             cov2D (:96-139), radius / conic with the viewer's guards (:194-246), quad extent,
 *                                             SH-DC colour (:58-61), sigmoid opacity
 *   fragmentShader    shaders.metal:291-317   power, the two discards, premultiplied output
 *   blending          mtl_engine.mm:354-362   src ONE, dst ONE_MINUS_SRC_ALPHA on RGBA, clear (1,1,1,1) (:583)
 *   draw order        mtl_engine.mm:403,597   instance j draws Gaussian sortedIndices[j] (GPURadixSort32, far first)
 * The fixed-function parts are stated as the API defines them: a quad of one instance shares z and w, so it is clipped as a
 * whole (kept iff 0 <= z <= w, Metal's clip volume); viewport x_fb = (ndc.x/2 + 1/2) W, y_fb = (1/2 - ndc.y/2) H (NDC y up,
 * framebuffer y down); a pixel is shaded when its centre lies inside the quad with the top-left fill rule; the varying
 * quadOffset interpolates to d = (centre.x - c_fb.x, c_fb.y - centre.y).  uniforms.screenSize must equal the target size.
 * Deviations, stated: fp32 blending by default (quantize = 1 rounds the target to RGBA8Unorm after every blend, as the
 * reference's attachment stores it); a fragment whose power is NaN is discarded; arithmetic order pinned as in orc_math.h.
 */
#include "orc_types.h"
#include "orc_math.h"
#include <vector>
#include <algorithm>
#include <cmath>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {

typedef struct {            /* Uniforms — mtl_engine.hpp:30-38 (host layout, 240 B) */
    float viewMatrix[16], projectionMatrix[16], viewProjectionMatrix[16];
    float screenSize[2], focalLength[2];
    float cameraPos[4];
    float _pad[4];
} OrcViewUniforms;

typedef struct {            /* what one instance hands to the rasteriser (VertexOut, shaders.metal:37-44, per quad) */
    float center[2];        /* quad centre in framebuffer pixels */
    float conic[3];
    float radius;
    float opacity;
    float color[3];
    uint32_t valid;
    uint32_t tileMinX, tileMinY, tileMaxX, tileMaxY;   /* 16x16 tiles the quad's pixel box touches (valid only) */
    uint32_t _pad;
} OrcViewSplat;

static_assert(sizeof(OrcViewUniforms) == 240 && sizeof(OrcViewSplat) == 64, "viewer layouts");

}

namespace {

constexpr float V_SH_C0 = 0.28209479177387814f;   /* shaders.metal:47 */
constexpr float V_MAX_SCALE = 8.0f;               /* shaders.metal:52 */

inline void mulp4(const float* m, const float p[3], float out[4]) {
    for (int r = 0; r < 4; r++) out[r] = ((m[0 + r] * p[0] + m[4 + r] * p[1]) + m[8 + r] * p[2]) + m[12 + r];
}

void view_splat(const OrcGaussian& g, const OrcViewUniforms& u, OrcViewSplat& s) {
    memset(&s, 0, sizeof(s));
    const float* p = g.position;
    if (std::isnan(p[0]) || std::isnan(p[1]) || std::isnan(p[2]) ||
        fabsf(p[0]) > 1e6f || fabsf(p[1]) > 1e6f || fabsf(p[2]) > 1e6f) return;          /* :175-180 */
    float clip[4];
    mulp4(u.viewProjectionMatrix, p, clip);                                               /* :183 */
    if (clip[3] <= 0.0f) return;                                                          /* :186-190 */
    if (!(clip[2] >= 0.0f && clip[2] <= clip[3])) return;                                 /* whole-quad depth clip (0 <= z <= w) */

    /* computeCovariance3D, :79-93: raw quaternion (no normalisation), scale = exp(clamp(log scale, +-8)) */
    const float w = g.rotation[0], x = g.rotation[1], y = g.rotation[2], z = g.rotation[3];
    float R[3][3];                                                                        /* R[row][col], :65-76 */
    R[0][0] = 1.0f - 2.0f * (y * y + z * z); R[1][0] = 2.0f * (x * y + w * z); R[2][0] = 2.0f * (x * z - w * y);
    R[0][1] = 2.0f * (x * y - w * z); R[1][1] = 1.0f - 2.0f * (x * x + z * z); R[2][1] = 2.0f * (y * z + w * x);
    R[0][2] = 2.0f * (x * z + w * y); R[1][2] = 2.0f * (y * z - w * x); R[2][2] = 1.0f - 2.0f * (x * x + y * y);
    float sc[3];
    for (int i = 0; i < 3; i++) sc[i] = orc_expf(orc_clamp(g.scale[i], -V_MAX_SCALE, V_MAX_SCALE));
    float M[3][3];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) M[r][c] = R[r][c] * sc[c];
    float Sg[3][3];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Sg[i][j] = (M[i][0] * M[j][0] + M[i][1] * M[j][1]) + M[i][2] * M[j][2];

    /* computeCovariance2D, :96-139 */
    float t[4];
    mulp4(u.viewMatrix, p, t);
    float tz = t[2];
    if (tz < 0.001f) tz = 0.001f;                                                         /* :102 */
    const float fx = u.focalLength[0], fy = u.focalLength[1];
    const float limx = 1.3f * fx / tz, limy = 1.3f * fy / tz;                             /* :105-106 (sic) */
    const float txtz = orc_clamp(t[0] / tz, -limx, limx);
    const float tytz = orc_clamp(t[1] / tz, -limy, limy);
    const float J00 = fx / tz, J02 = -fx * txtz / tz, J11 = fy / tz, J12 = -fy * tytz / tz;   /* :112-115 */
    const float* V = u.viewMatrix;
    float T0[3], T1[3];
    for (int c = 0; c < 3; c++) {
        T0[c] = J00 * V[c * 4 + 0] + J02 * V[c * 4 + 2];
        T1[c] = J11 * V[c * 4 + 1] + J12 * V[c * 4 + 2];
    }
    float U0[3], U1[3];
    for (int l = 0; l < 3; l++) {
        U0[l] = (T0[0] * Sg[0][l] + T0[1] * Sg[1][l]) + T0[2] * Sg[2][l];
        U1[l] = (T1[0] * Sg[0][l] + T1[1] * Sg[1][l]) + T1[2] * Sg[2][l];
    }
    float ca = (U0[0] * T0[0] + U0[1] * T0[1]) + U0[2] * T0[2];
    float cb = (U0[0] * T1[0] + U0[1] * T1[1]) + U0[2] * T1[2];
    float cc = (U1[0] * T1[0] + U1[1] * T1[1]) + U1[2] * T1[2];
    ca += 0.3f; cc += 0.3f;                                                               /* :130-131 */

    /* vertexShader :194-246 */
    float a = fmaxf(ca, 0.3f), b = cb, c = fmaxf(cc, 0.3f);
    const float maxB = sqrtf(a * c) * 0.99f;
    b = orc_clamp(b, -maxB, maxB);
    const float det = a * c - b * b;
    float radius, k0, k1, k2;
    if (det > 0.0001f) {
        const float mid = 0.5f * (a + c);
        const float disc = fmaxf(0.0f, mid * mid - det);
        const float lambda = mid + sqrtf(disc);
        radius = ceilf(3.0f * sqrtf(fmaxf(lambda, 0.1f)));
        radius = orc_clamp(radius, 1.0f, 512.0f);
        const float invDet = 1.0f / det;
        k0 = c * invDet; k1 = -b * invDet; k2 = a * invDet;
        const float expected = 9.0f / (radius * radius);
        const float actual = 0.5f * (k0 + k2);
        if (actual < expected * 0.01f) {
            const float scale = expected / fmaxf(actual, 1e-10f);
            const float f = sqrtf(scale);
            k0 *= f; k1 *= f; k2 *= f;
        }
    } else {
        radius = 15.0f; k0 = 0.04f; k1 = 0.0f; k2 = 0.04f;
    }
    const float ndcx = clip[0] / clip[3], ndcy = clip[1] / clip[3];                       /* :258 */
    const float W = u.screenSize[0], H = u.screenSize[1];
    const float cx = (ndcx * 0.5f + 0.5f) * W;                                            /* viewport transform */
    const float cy = (0.5f - ndcy * 0.5f) * H;
    if (!(cx == cx) || !(cy == cy)) return;
    /* 16x16 tiles the quad's pixel box can touch: pixels floor(c - r) .. floor(c + r), clipped to the target */
    const float x0 = fmaxf(floorf(cx - radius), 0.0f), x1 = fminf(floorf(cx + radius), W - 1.0f);
    const float y0 = fmaxf(floorf(cy - radius), 0.0f), y1 = fminf(floorf(cy + radius), H - 1.0f);
    if (!(x0 <= x1) || !(y0 <= y1)) return;                                               /* off screen: nothing is shaded */
    s.center[0] = cx; s.center[1] = cy;
    s.conic[0] = k0; s.conic[1] = k1; s.conic[2] = k2;
    s.radius = radius;
    s.opacity = 1.0f / (1.0f + orc_expf(-orc_clamp(g.opacity, -8.0f, 8.0f)));             /* :276 */
    s.color[0] = orc_clamp(V_SH_C0 * g.sh[0] + 0.5f, 0.0f, 1.0f);                         /* evalSH, :58-61 */
    s.color[1] = orc_clamp(V_SH_C0 * g.sh[4] + 0.5f, 0.0f, 1.0f);
    s.color[2] = orc_clamp(V_SH_C0 * g.sh[8] + 0.5f, 0.0f, 1.0f);
    s.valid = 1;
    s.tileMinX = (uint32_t)x0 / 16u; s.tileMaxX = (uint32_t)x1 / 16u;
    s.tileMinY = (uint32_t)y0 / 16u; s.tileMaxY = (uint32_t)y1 / 16u;
}

inline float quant8(float v) { return rintf(fminf(fmaxf(v, 0.0f), 1.0f) * 255.0f) / 255.0f; }

/* one fragment of splat s at pixel (px,py) blended into dst[4]; returns 0 not covered, 1 shaded but discarded, 2 blended */
inline int shade(const OrcViewSplat& s, int px, int py, float* dst, int quantize) {
    const float dx = ((float)px + 0.5f) - s.center[0];
    const float dy = s.center[1] - ((float)py + 0.5f);
    const float r = s.radius;
    if (!(dx >= -r && dx < r && dy > -r && dy <= r)) return 0;                            /* top-left fill rule */
    const float t1 = (s.conic[0] * dx) * dx, t2 = ((2.0f * s.conic[1]) * dx) * dy, t3 = (s.conic[2] * dy) * dy;
    const float power = -0.5f * ((t1 + t2) + t3);                                         /* :296-298 */
    if (!(power <= 0.0f && power >= -4.5f)) return 1;                                     /* :301 (NaN discarded too) */
    const float alpha = s.opacity * orc_exp_blend(power);                                 /* :304-305 */
    if (alpha < 1.0f / 255.0f) return 1;                                                  /* :308 */
    const float om = 1.0f - alpha;
    for (int c = 0; c < 3; c++) dst[c] = (s.color[c] * alpha) + dst[c] * om;              /* :311 + blend state */
    dst[3] = alpha + dst[3] * om;
    if (quantize) for (int c = 0; c < 4; c++) dst[c] = quant8(dst[c]);
    return 2;
}

} // namespace

extern "C" {

void orc_viewer_splats(const OrcGaussian* g, uint32_t n, const OrcViewUniforms* u, OrcViewSplat* out, int threads) {
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#else
    threads = 1;
#endif
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) view_splat(g[i], *u, out[i]);
}

/* MTLEngine::render's draw (mtl_engine.mm:575-605): instance j = Gaussian order[j], blended in that order into an image
 * cleared to white.  image: H*W*4 floats.  stats[0] = fragments shaded (pixel centres inside a quad), stats[1] = fragments
 * blended.  Threads split the image into row bands; every band walks the instances in draw order, so the result does not
 * depend on the thread count. */
void orc_viewer_render(const OrcGaussian* g, uint32_t n, const uint32_t* order, const OrcViewUniforms* u, uint32_t width,
                       uint32_t height, int quantize, float* image, uint64_t* stats, int threads) {
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#else
    threads = 1;
#endif
    std::vector<OrcViewSplat> sp(n);
    orc_viewer_splats(g, n, u, sp.data(), threads);
    const int W = (int)width, H = (int)height;
    for (size_t i = 0; i < (size_t)W * H * 4; i++) image[i] = 1.0f;                       /* clear colour, :583 */
    uint64_t shaded = 0, blended = 0;
    const int bands = std::max(1, std::min(threads * 4, H));
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1) reduction(+ : shaded, blended)
    for (int band = 0; band < bands; band++) {
        const int yb0 = (int)((int64_t)H * band / bands), yb1 = (int)((int64_t)H * (band + 1) / bands);
        for (uint32_t j = 0; j < n; j++) {
            const uint32_t i = order[j];
            if (i >= n) continue;
            const OrcViewSplat& s = sp[i];
            if (!s.valid) continue;
            const int x0 = std::max(0, (int)floorf(s.center[0] - s.radius)), x1 = std::min(W - 1, (int)floorf(s.center[0] + s.radius));
            const int y0 = std::max(yb0, (int)floorf(s.center[1] - s.radius)), y1 = std::min(yb1 - 1, (int)floorf(s.center[1] + s.radius));
            for (int py = y0; py <= y1; py++)
                for (int px = x0; px <= x1; px++) {
                    const int k = shade(s, px, py, image + ((size_t)py * W + px) * 4, quantize);
                    shaded += (k != 0); blended += (k == 2);
                }
        }
    }
    if (stats) { stats[0] = shaded; stats[1] = blended; }
}

} /* extern "C" */
