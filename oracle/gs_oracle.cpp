/*
 * gs_oracle.cpp — scalar CPU restatement of the reference's tiled-splatting hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see orc_types.h).  PARITY UNPINNED: the reference has no golden
 * vectors and its Metal sources cannot be built here; tests/ pins this file with hand-computed
 * cases, invariants and finite differences instead.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/GuassianSplatting/).  Arithmetic is the pinned fp32 order of orc_math.h.
 * MSL indexing rule used throughout: m[i][j] is COLUMN i, ROW j; floatNxN(a,b,c) takes columns.
 *
 * Build: see oracle/Makefile (g++ -O2 -ffp-contract=off -fno-fast-math -mfma -fopenmp).
 */
#include "orc_types.h"
#include "orc_math.h"
#include <vector>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <cmath>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr float SH_C0 = 0.28209479177387814f;   /* tiled_shaders.metal:83 */
constexpr uint32_t TILE = 16;                   /* tiled_shaders.metal:84 */
constexpr float MAX_RADIUS = 512.0f;            /* tiled_shaders.metal:85 */
constexpr float MAX_SCALE = 5.0f;               /* tiled_shaders.metal:87 */
constexpr float MAX_SCALE_TRAIN = 4.0f;         /* shaders.metal:55 */
constexpr float GPU_MIN_OPACITY = 0.005f;       /* tiled_shaders.metal:742 */
constexpr uint32_t GPU_MAX_TILES = 256u;        /* tiled_shaders.metal:743 */
constexpr float ALPHA_MIN = 1.0f / 255.0f;      /* tiled_shaders.metal:366,454,495 */
constexpr float ALPHA_MAX = 0.99f;              /* tiled_shaders.metal:363 */
constexpr float T_EPS = 0.0001f;                /* tiled_shaders.metal:334,458 */

/* column-major 3x3, c[col][row], mirroring MSL float3x3 */
struct M3 { float c[3][3]; };
inline M3 m3_mul(const M3& A, const M3& B) {    /* (A*B)[col][row] = sum_k A[k][row]*B[col][k] */
    M3 R;
    for (int col = 0; col < 3; col++)
        for (int row = 0; row < 3; row++)
            R.c[col][row] = (A.c[0][row] * B.c[col][0] + A.c[1][row] * B.c[col][1]) + A.c[2][row] * B.c[col][2];
    return R;
}
inline M3 m3_abs(const M3& A) { M3 R; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R.c[i][j] = fabsf(A.c[i][j]); return R; }
inline M3 m3_T(const M3& A) { M3 R; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R.c[i][j] = A.c[j][i]; return R; }
inline void m3_mulv(const M3& A, const float v[3], float out[3]) { /* sum_c A[c]*v[c] */
    for (int r = 0; r < 3; r++) out[r] = (A.c[0][r] * v[0] + A.c[1][r] * v[1]) + A.c[2][r] * v[2];
}

/* tiled_shaders.metal:91-99 quatToMat; q = (w,x,y,z) in (.x,.y,.z,.w) */
inline M3 quat_to_mat(const float q[4]) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    M3 R;
    R.c[0][0] = 1.0f - 2.0f * (y * y + z * z); R.c[0][1] = 2.0f * (x * y + w * z); R.c[0][2] = 2.0f * (x * z - w * y);
    R.c[1][0] = 2.0f * (x * y - w * z); R.c[1][1] = 1.0f - 2.0f * (x * x + z * z); R.c[1][2] = 2.0f * (y * z + w * x);
    R.c[2][0] = 2.0f * (x * z + w * y); R.c[2][1] = 2.0f * (y * z - w * x); R.c[2][2] = 1.0f - 2.0f * (x * x + y * y);
    return R;
}

/* mat4 (column-major m[c*4+r]) * (p,1):  ((c0*x + c1*y) + c2*z) + c3 */
inline void m4_mulp(const float* m, const float p[3], float out[4]) {
    for (int r = 0; r < 4; r++)
        out[r] = ((m[0 + r] * p[0] + m[4 + r] * p[1]) + m[8 + r] * p[2]) + m[12 + r];
}

/* The pinned blend exponent.  Reference: power = -0.5*(A dx^2 + 2 B dx dy + C dy^2)
 * (tiled_shaders.metal:354-356, 442-444, 484-486).  Pinned as
 *   power = fma(dx, fma(Ah, dx, Bn*dy), (Ch*dy)*dy),  Ah=-0.5*A, Bn=-B, Ch=-0.5*C
 * (the scalings by -0.5 and -1 are exact). */
inline float blend_power(float Ah, float Bn, float Ch, float dx, float dy) {
    float u = orc_fma(Ah, dx, Bn * dy);
    float v = (Ch * dy) * dy;
    return orc_fma(dx, u, v);
}

inline int set_threads(int threads) {
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
    return threads;
#else
    (void)threads; return 1;
#endif
}

} // namespace

/* Self-check switch (tests/test_oracle_fd.py only; default 0 = the reference's arithmetic).  When 1, the
 * two factor-of-2 quirks of the reference's conic->cov2D step are removed (dL_dConic.y without the 2.0 of
 * :572, and dL_dCov2D.y halved on the off-diagonals of :624-628) and the quaternion formula is fed dL_dR
 * instead of its transpose (:675; the transpose flips the sign of every w-coupled term), which makes the analytic scale/rotation
 * gradients agree with finite differences — evidence that the rest of the chain is transliterated right. */
static int g_fd_selfcheck = 0;

extern "C" {

void orc_set_fd_selfcheck(int on) { g_fd_selfcheck = on; }

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

float orc_test_expf(float x) { return orc_expf(x); }
float orc_test_exp_blend(float x) { return orc_exp_blend(x); }

/* ------------------------------------------------------------------------------------------
 * A3  projectGaussians — tiled_shaders.metal:102-304
 * ---------------------------------------------------------------------------------------- */
static void project_one(const OrcGaussian& g, const OrcUniforms& u, OrcProjected& out) {
    OrcProjected proj;
    memset(&proj, 0, sizeof(proj));                       /* :112 */
    proj.radius = 0;
    proj.tileMinX = 0xFFFFFFFFu; proj.tileMaxX = 0;       /* :114-117 */
    proj.tileMinY = 0xFFFFFFFFu; proj.tileMaxY = 0;

    const float* p = g.position;
    const float* ls = g.scale;
    if (std::isnan(p[0]) || std::isnan(p[1]) || std::isnan(p[2]) ||
        std::isnan(ls[0]) || std::isnan(ls[1]) || std::isnan(ls[2]) ||
        fabsf(p[0]) > 1e6f || fabsf(p[1]) > 1e6f || fabsf(p[2]) > 1e6f) {   /* :120-125 */
        out = proj; return;
    }

    float viewPos[4], clipPos[4];
    m4_mulp(u.viewMatrix, p, viewPos);                    /* :129 */
    m4_mulp(u.viewProjectionMatrix, p, clipPos);          /* :130 */
    if (clipPos[3] <= 0.1f || viewPos[2] <= 0.1f) { out = proj; return; }   /* :135-138 */

    float ndcx = clipPos[0] / clipPos[3];                 /* :141 */
    float ndcy = clipPos[1] / clipPos[3];
    if (fabsf(ndcx) > 1.2f || fabsf(ndcy) > 1.2f) { out = proj; return; }   /* :144-147 */

    proj.screenPos[0] = (ndcx * 0.5f + 0.5f) * u.screenSize[0];   /* :150-153 */
    proj.screenPos[1] = (ndcy * 0.5f + 0.5f) * u.screenSize[1];
    proj.depth = viewPos[2];                              /* :156 */
    proj.viewPos_xy[0] = viewPos[0]; proj.viewPos_xy[1] = viewPos[1];   /* :157 */

    float s[3];
    for (int i = 0; i < 3; i++) s[i] = orc_expf(orc_clamp(ls[i], -MAX_SCALE, MAX_SCALE));  /* :160-161 */
    float maxS = fmaxf(fmaxf(s[0], s[1]), s[2]);
    float minS = fminf(fminf(s[0], s[1]), s[2]);
    if (maxS > 20.0f * minS) {                            /* :166-170 */
        float targetMax = 20.0f * minS;
        float k = targetMax / maxS;
        for (int i = 0; i < 3; i++) s[i] = s[i] * k;
    }

    float q[4] = { g.rotation[0], g.rotation[1], g.rotation[2], g.rotation[3] };
    float qLen = sqrtf(((q[0] * q[0] + q[1] * q[1]) + q[2] * q[2]) + q[3] * q[3]);   /* :174 */
    if (qLen > 0.001f) { for (int i = 0; i < 4; i++) q[i] = q[i] / qLen; }
    else { q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0; }     /* :175 */

    M3 R = quat_to_mat(q);                                /* :180 */
    /* M = R*S: column c of R scaled by s[c]  (:188).  Sigma = M * M^T  (:190), symmetric. */
    float Mrc[3][3];                                      /* Mrc[row][col] */
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Mrc[r][c] = R.c[c][r] * s[c];
    float Sg[3][3];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++)
        Sg[i][j] = (Mrc[i][0] * Mrc[j][0] + Mrc[i][1] * Mrc[j][1]) + Mrc[i][2] * Mrc[j][2];

    float z = viewPos[2];
    float fx = u.focalLength[0], fy = u.focalLength[1];
    float limx = 1.3f * fx / z;                           /* :198-201 (sic: 1.3*f/z) */
    float limy = 1.3f * fy / z;
    float txtz = orc_clamp(viewPos[0] / z, -limx, limx);
    float tytz = orc_clamp(viewPos[1] / z, -limy, limy);
    float J00 = fx / z;                                   /* :205-208 */
    float J02 = -fx * txtz / z;
    float J11 = fy / z;
    float J12 = -fy * tytz / z;

    /* W(r,c) = viewMatrix[c][r], upper 3x3 (:218-220).  T = J*W (:223): rows 0,1 only. */
    const float* V = u.viewMatrix;
    float T0[3], T1[3];
    for (int c = 0; c < 3; c++) {
        T0[c] = J00 * V[c * 4 + 0] + J02 * V[c * 4 + 2];
        T1[c] = J11 * V[c * 4 + 1] + J12 * V[c * 4 + 2];
    }
    /* cov = (T*Sigma)*T^T (:225) */
    float U0[3], U1[3];
    for (int l = 0; l < 3; l++) {
        U0[l] = (T0[0] * Sg[0][l] + T0[1] * Sg[1][l]) + T0[2] * Sg[2][l];
        U1[l] = (T1[0] * Sg[0][l] + T1[1] * Sg[1][l]) + T1[2] * Sg[2][l];
    }
    float a = (U0[0] * T0[0] + U0[1] * T0[1]) + U0[2] * T0[2];    /* cov2D_mat[0][0] :228 */
    float b = (U0[0] * T1[0] + U0[1] * T1[1]) + U0[2] * T1[2];    /* cov2D_mat[1][0] = row 0, col 1 :229 */
    float c = (U1[0] * T1[0] + U1[1] * T1[1]) + U1[2] * T1[2];    /* cov2D_mat[1][1] :230 */
    a += 0.3f; c += 0.3f;                                 /* :233-234 */
    proj.cov2D[0] = a; proj.cov2D[1] = b; proj.cov2D[2] = c;      /* :237 */

    float det = a * c - b * b;                            /* :240 */
    if (det < 0.0001f) { out = proj; return; }            /* :241-244 */
    float inv_det = 1.0f / det;                           /* :247-248 */
    proj.conic[0] = c * inv_det; proj.conic[1] = -b * inv_det; proj.conic[2] = a * inv_det;

    float mid = 0.5f * (a + c);                           /* :251-255 */
    float disc = mid * mid - det;
    float l1 = mid + sqrtf(fmaxf(0.1f, disc));
    float rawRadius = 3.0f * sqrtf(l1);
    proj.radius = fminf(ceilf(rawRadius), MAX_RADIUS);
    if (proj.radius <= 0) { out = proj; return; }         /* :258-261 */

    float r = proj.radius;                                /* :264-268 */
    int minX = std::max(0, (int)(proj.screenPos[0] - r));
    int minY = std::max(0, (int)(proj.screenPos[1] - r));
    int maxX = std::min((int)u.screenSize[0] - 1, (int)(proj.screenPos[0] + r));
    int maxY = std::min((int)u.screenSize[1] - 1, (int)(proj.screenPos[1] + r));
    if (minX > maxX || minY > maxY) { proj.radius = 0; out = proj; return; }   /* :271-275 */

    proj.tileMinX = (uint32_t)minX / TILE;                /* :278-281 */
    proj.tileMinY = (uint32_t)minY / TILE;
    proj.tileMaxX = std::min((uint32_t)maxX / TILE, u.numTilesX - 1);
    proj.tileMaxY = std::min((uint32_t)maxY / TILE, u.numTilesY - 1);
    uint32_t tilesX = proj.tileMaxX - proj.tileMinX + 1;  /* :284-290 */
    uint32_t tilesY = proj.tileMaxY - proj.tileMinY + 1;
    if (tilesX * tilesY > 256) { proj.radius = 0; out = proj; return; }

    float rawOpacity = orc_clamp(g.opacity, -8.0f, 8.0f); /* :293-294 */
    proj.opacity = 1.0f / (1.0f + orc_expf(-rawOpacity));
    proj.color[0] = orc_clamp(SH_C0 * g.sh[0] + 0.5f, 0.0f, 1.0f);   /* :297-301 */
    proj.color[1] = orc_clamp(SH_C0 * g.sh[4] + 0.5f, 0.0f, 1.0f);
    proj.color[2] = orc_clamp(SH_C0 * g.sh[8] + 0.5f, 0.0f, 1.0f);
    out = proj;                                           /* :303 */
}

void orc_project(const OrcGaussian* g, uint32_t n, const OrcUniforms* u, OrcProjected* out, int threads) {
    threads = set_threads(threads);
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) project_one(g[i], *u, out[i]);
}

/* ------------------------------------------------------------------------------------------
 * A4  generateTilePairs — tiled_shaders.metal:745-794
 * Order pinned (SURVEY note K): Gaussian-index order, ty-major, tx-minor; no drops (note C).
 * keys/vals may be NULL to count only.  Returns the number of pairs.
 * ---------------------------------------------------------------------------------------- */
static inline uint32_t pair_count(const OrcProjected& p) {
    if (p.radius <= 0) return 0;                                        /* :760 */
    if (p.tileMinX > p.tileMaxX || p.tileMinY > p.tileMaxY) return 0;   /* :761 */
    if (p.opacity < GPU_MIN_OPACITY) return 0;                          /* :762 */
    if (p.tileMinX > 10000 || p.tileMaxX > 10000 || p.tileMinY > 10000 || p.tileMaxY > 10000) return 0; /* :763 */
    uint32_t tc = (p.tileMaxX - p.tileMinX + 1) * (p.tileMaxY - p.tileMinY + 1);   /* :766-768 */
    if (tc > GPU_MAX_TILES) return 0;                                   /* :770 */
    return tc;
}
static inline uint32_t depth_key(float depth) {                         /* :773-774 */
    uint32_t k = orc_f2u(depth);
    return (k & 0x80000000u) ? ~k : (k | 0x80000000u);
}

uint64_t orc_generate_pairs(const OrcProjected* proj, uint32_t n, uint32_t numTilesX,
                            uint64_t* keys, uint32_t* vals, uint64_t capacity) {
    uint64_t w = 0;
    for (uint32_t i = 0; i < n; i++) {
        const OrcProjected& p = proj[i];
        uint32_t tc = pair_count(p);
        if (tc == 0) continue;
        if (keys) {
            if (w + tc > capacity) return (uint64_t)-1;
            uint32_t dk = depth_key(p.depth);
            uint64_t idx = w;
            for (uint32_t ty = p.tileMinY; ty <= p.tileMaxY; ty++)      /* :784-793 */
                for (uint32_t tx = p.tileMinX; tx <= p.tileMaxX; tx++) {
                    uint32_t tileIdx = ty * numTilesX + tx;
                    keys[idx] = ((uint64_t)tileIdx << 32) | (uint64_t)dk;
                    vals[idx] = i;
                    idx++;
                }
        }
        w += tc;
    }
    return w;
}

/* ------------------------------------------------------------------------------------------
 * A5  parallelRadixSort — tiled_rasterizer.mm:27-102 (what the reference ships: CPU LSD radix
 * sort, 8 passes x 8 bits, NUM_THREADS chunks with per-thread histograms and chunk-ordered
 * scatter, hence stable).  GCD dispatch_apply -> OpenMP.  Stable for all n (note J: the
 * reference falls back to unstable std::sort below 1000 pairs; the oracle does not).
 * ---------------------------------------------------------------------------------------- */
void orc_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t n, int threads) {
    if (n == 0) return;
    threads = set_threads(threads);
    if ((uint64_t)threads > n) threads = 1;
    struct KV { uint64_t k; uint32_t v; };
    std::vector<KV> a(n), b(n);
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) { a[i].k = keys[i]; a[i].v = vals[i]; }   /* :493-498 */
    std::vector<uint32_t> hist((size_t)threads * 256), off((size_t)threads * 256);
    KV* src = a.data(); KV* dst = b.data();
    const uint64_t chunk = (n + threads - 1) / threads;
    for (int pass = 0; pass < 8; pass++) {                /* :52 */
        const int shift = pass * 8;
#pragma omp parallel for num_threads(threads) schedule(static, 1)
        for (int t = 0; t < threads; t++) {               /* :58-68 */
            uint32_t* h = &hist[(size_t)t * 256];
            memset(h, 0, 256 * sizeof(uint32_t));
            uint64_t s = (uint64_t)t * chunk, e = std::min(s + chunk, n);
            for (uint64_t i = s; i < e; i++) h[(src[i].k >> shift) & 0xFF]++;
        }
        uint32_t sum = 0;                                 /* :71-79 */
        for (int d = 0; d < 256; d++)
            for (int t = 0; t < threads; t++) { off[(size_t)t * 256 + d] = sum; sum += hist[(size_t)t * 256 + d]; }
#pragma omp parallel for num_threads(threads) schedule(static, 1)
        for (int t = 0; t < threads; t++) {               /* :81-93 */
            uint32_t lo[256];
            memcpy(lo, &off[(size_t)t * 256], sizeof(lo));
            uint64_t s = (uint64_t)t * chunk, e = std::min(s + chunk, n);
            for (uint64_t i = s; i < e; i++) dst[lo[(src[i].k >> shift) & 0xFF]++] = src[i];
        }
        std::swap(src, dst);                              /* :96 */
    }
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) { keys[i] = src[i].k; vals[i] = src[i].v; }   /* :509-512 */
}

/* 32-bit viewer depth keys — sort.metal:28-33 (floatToSortable), :36-72 (computeDepths) */
void orc_depth_keys32(const float* positions4, const float cam[3], uint32_t n, uint32_t* keys, uint32_t* vals) {
    for (uint32_t i = 0; i < n; i++) {
        const float* p = positions4 + 4 * (size_t)i;
        vals[i] = i;
        if (std::isnan(p[0]) || std::isnan(p[1]) || std::isnan(p[2]) ||
            std::isinf(p[0]) || std::isinf(p[1]) || std::isinf(p[2])) { keys[i] = 0xFFFFFFFFu; continue; }
        float dx = p[0] - cam[0], dy = p[1] - cam[1], dz = p[2] - cam[2];
        float depth = (dx * dx + dy * dy) + dz * dz;
        if (std::isnan(depth) || std::isinf(depth) || depth < 0.0f) { keys[i] = 0xFFFFFFFFu; continue; }
        uint32_t u = orc_f2u(depth);
        uint32_t mask = (uint32_t)(-(int32_t)(u >> 31)) | 0x80000000u;
        keys[i] = ~(u ^ mask);
    }
}
/* stable LSD sort of (u32 key, u32 value), 4 passes x 8 bits — gpu_sort.mm:196-338 contract */
void orc_sort_pairs32(uint32_t* keys, uint32_t* vals, uint64_t n) {
    std::vector<uint32_t> k2(n), v2(n);
    uint32_t *ks = keys, *vs = vals, *kd = k2.data(), *vd = v2.data();
    for (int pass = 0; pass < 4; pass++) {
        uint32_t h[257] = {0};
        int shift = pass * 8;
        for (uint64_t i = 0; i < n; i++) h[((ks[i] >> shift) & 0xFF) + 1]++;
        for (int d = 0; d < 256; d++) h[d + 1] += h[d];
        for (uint64_t i = 0; i < n; i++) { uint32_t p = h[(ks[i] >> shift) & 0xFF]++; kd[p] = ks[i]; vd[p] = vs[i]; }
        std::swap(ks, kd); std::swap(vs, vd);
    }
    /* 4 passes: result is back in keys/vals */
}

/* ------------------------------------------------------------------------------------------
 * A7  buildTileRanges — sort.metal:553-589.  start = lower_bound(tile), count = run length.
 * Empty tiles keep start = insertion point (not 0).
 * ---------------------------------------------------------------------------------------- */
void orc_tile_ranges(const uint64_t* sortedKeys, uint32_t numPairs, uint32_t numTiles, OrcTileRange* out) {
    for (uint32_t id = 0; id < numTiles; id++) {
        uint32_t lo = 0, hi = numPairs;
        while (lo < hi) {                                 /* :568-577 */
            uint32_t mid = (lo + hi) / 2;
            uint32_t keyTile = (uint32_t)(sortedKeys[mid] >> 32);
            if (keyTile < id) lo = mid + 1; else hi = mid;
        }
        uint32_t start = lo, count = 0;
        for (uint32_t i = start; i < numPairs; i++) {     /* :581-586 */
            if ((uint32_t)(sortedKeys[i] >> 32) != id) break;
            count++;
        }
        out[id].start = start; out[id].count = count;
    }
}

/* ------------------------------------------------------------------------------------------
 * A8  tiledForward — tiled_shaders.metal:307-385, graded fp32 mode (SURVEY note H: the reference
 * accumulates in half; fp32 with the source's thresholds is the graded mode).
 * image: W*H*4 floats RGBA.  lastIdx: W*H.  tfinal (optional): the T_final the backward's first
 * pass (:430-460) would compute, for cross-checking a forward that exports it.
 * stats (optional): [0] = E (list entries visited), [1] = C (contributing evaluations).
 * ---------------------------------------------------------------------------------------- */
void orc_forward(const OrcProjected* proj, const uint32_t* sortedIdx, const OrcTileRange* ranges,
                 const OrcUniforms* u, float* image, uint32_t* lastIdxOut, uint64_t* stats, int threads) {
    const uint32_t W = (uint32_t)u->screenSize[0], H = (uint32_t)u->screenSize[1];
    threads = set_threads(threads);
    uint64_t E = 0, C = 0;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1) reduction(+:E,C)
    for (int64_t tile = 0; tile < (int64_t)(u->numTilesX * u->numTilesY); tile++) {
        uint32_t tX = (uint32_t)tile % u->numTilesX, tY = (uint32_t)tile / u->numTilesX;
        const OrcTileRange range = ranges[tile];          /* :320-323 */
        for (uint32_t py = tY * TILE; py < std::min(H, (tY + 1) * TILE); py++)
        for (uint32_t px = tX * TILE; px < std::min(W, (tX + 1) * TILE); px++) {
            float col[3] = {0, 0, 0};
            float T = 1.0f;
            float pixx = (float)px + 0.5f, pixy = (float)py + 0.5f;   /* :328 */
            uint32_t lastIdx = 0; bool hasContrib = false;
            for (uint32_t i = 0; i < range.count && T > T_EPS; i++) {   /* :334 */
                uint32_t sortIdx = range.start + i;
                uint32_t gIdx = sortedIdx[sortIdx];
                E++;
                if (gIdx >= u->numGaussians) continue;    /* :338 */
                const OrcProjected& p = proj[gIdx];
                if (p.radius <= 0) continue;              /* :344 */
                float dx = pixx - p.screenPos[0], dy = pixy - p.screenPos[1];   /* :347 */
                float conicMag = (fabsf(p.conic[0]) + fabsf(p.conic[1])) + fabsf(p.conic[2]);   /* :350-351 */
                if (conicMag < 0.0001f) continue;
                float power = blend_power(-0.5f * p.conic[0], -p.conic[1], -0.5f * p.conic[2], dx, dy);
                if (!(power <= 0.0f) || power < -4.5f) continue;   /* :359; NaN pinned as skip */
                float G = orc_exp_blend(power);           /* :362 */
                float alpha = fminf(p.opacity * G, ALPHA_MAX);   /* :363 */
                if (alpha < ALPHA_MIN) continue;          /* :366 */
                float w = alpha * T;                      /* :369  colour += c*alpha*T */
                col[0] = orc_fma(p.color[0], w, col[0]);
                col[1] = orc_fma(p.color[1], w, col[1]);
                col[2] = orc_fma(p.color[2], w, col[2]);
                T = T * (1.0f - alpha);                   /* :370 */
                lastIdx = sortIdx; hasContrib = true;     /* :372-373 */
                C++;
            }
            size_t pix = (size_t)py * W + px;
            image[pix * 4 + 0] = col[0] + T;              /* :377-378 white background */
            image[pix * 4 + 1] = col[1] + T;
            image[pix * 4 + 2] = col[2] + T;
            image[pix * 4 + 3] = 1.0f;                    /* :384 */
            lastIdxOut[pix] = hasContrib ? lastIdx : 0xFFFFFFFFu;   /* :382 */
        }
    }
    if (stats) { stats[0] = E; stats[1] = C; }
}

/* ------------------------------------------------------------------------------------------
 * A8 (literal precision)  tiledForward as the Metal source types it — tiled_shaders.metal:325-370: colour, T, power, G
 * and alpha are `half`.  Every half-typed operation is rounded to nearest-even _Float16 (g++'s native type); the
 * float expression of the power (:354-356) is evaluated in fp32 in the source's order and then converted (:354
 * `half(...)`); exp(half) is taken as fp32 libm expf rounded once to half (Apple's fast-math exp is not reproducible
 * off-device; SURVEY note H).  Thresholds are the half literals: 0.0001h, -4.5h, 0.99h, half(1.0/255.0).
 * This is NOT the graded mode (the graded mode is fp32, orc_forward): it exists to measure how far the reference's
 * precision choice moves the image and lastContribIdx from the fp32 statement (reported in BASELINE.md).
 * ---------------------------------------------------------------------------------------- */
void orc_forward_half(const OrcProjected* proj, const uint32_t* sortedIdx, const OrcTileRange* ranges,
                      const OrcUniforms* u, float* image, uint32_t* lastIdxOut, int threads) {
    typedef _Float16 half;
    const uint32_t W = (uint32_t)u->screenSize[0], H = (uint32_t)u->screenSize[1];
    threads = set_threads(threads);
    const half T_EPS_H = (half)0.0001f, PMIN_H = (half)-4.5f, AMAX_H = (half)0.99f, AMIN_H = (half)(1.0 / 255.0), ONE_H = (half)1.0f;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
    for (int64_t tile = 0; tile < (int64_t)(u->numTilesX * u->numTilesY); tile++) {
        uint32_t tX = (uint32_t)tile % u->numTilesX, tY = (uint32_t)tile / u->numTilesX;
        const OrcTileRange range = ranges[tile];
        for (uint32_t py = tY * TILE; py < std::min(H, (tY + 1) * TILE); py++)
        for (uint32_t px = tX * TILE; px < std::min(W, (tX + 1) * TILE); px++) {
            half col[3] = {(half)0.0f, (half)0.0f, (half)0.0f};    /* :325 */
            half T = ONE_H;                                          /* :326 */
            float pixx = (float)px + 0.5f, pixy = (float)py + 0.5f;
            uint32_t lastIdx = 0; bool hasContrib = false;
            for (uint32_t i = 0; i < range.count && T > T_EPS_H; i++) {   /* :334 */
                uint32_t sortIdx = range.start + i;
                uint32_t gIdx = sortedIdx[sortIdx];
                if (gIdx >= u->numGaussians) continue;
                const OrcProjected& p = proj[gIdx];
                if (p.radius <= 0) continue;
                float dx = pixx - p.screenPos[0], dy = pixy - p.screenPos[1];
                float conicMag = (fabsf(p.conic[0]) + fabsf(p.conic[1])) + fabsf(p.conic[2]);
                if (conicMag < 0.0001f) continue;
                /* :354-356, fp32 in source order, then one conversion to half */
                float pf = -0.5f * ((p.conic[0] * dx * dx + 2.0f * p.conic[1] * dx * dy) + p.conic[2] * dy * dy);
                half power = (half)pf;
                if (power > (half)0.0f || power < PMIN_H) continue;   /* :359 */
                if (power != power) continue;
                half G = (half)expf((float)power);                   /* :362 */
                half alpha = (half)((half)p.opacity * G);            /* :363 */
                if (alpha > AMAX_H) alpha = AMAX_H;
                if (alpha < AMIN_H) continue;                        /* :366 */
                for (int c = 0; c < 3; c++) {                        /* :369 color += half3(p.color) * alpha * T */
                    half t1 = (half)((half)p.color[c] * alpha);
                    half t2 = (half)(t1 * T);
                    col[c] = (half)(col[c] + t2);
                }
                T = (half)(T * (half)(ONE_H - alpha));               /* :370 */
                lastIdx = sortIdx; hasContrib = true;
            }
            size_t pix = (size_t)py * W + px;
            for (int c = 0; c < 3; c++) image[pix * 4 + c] = (float)(half)(col[c] + (half)(ONE_H * T));   /* :377-378 */
            image[pix * 4 + 3] = 1.0f;
            lastIdxOut[pix] = hasContrib ? lastIdx : 0xFFFFFFFFu;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * A10  tiledBackward — tiled_shaders.metal:388-738.
 * rendered / gt: W*H*4 float RGBA.  quantize: SURVEY note G (RGBA8 render target): when set, the
 * rendered colour is rounded to 8 bits before sign(diff).
 * grads: n x 112 B, zeroed here (tiled_rasterizer.mm:687) then accumulated.  Accumulation is done
 * in double and rounded once (the reference uses relaxed float atomics in arbitrary order, :699-736).
 * tfinal (optional, W*H): T_final of pass 1 per pixel (NaN where lastIdx == UINT_MAX).
 * ---------------------------------------------------------------------------------------- */
namespace {
struct GradAcc { double v[16]; };   /* pos3, opacity, scale3, rot4, shdc3, viewspace2 */
}
/* absOut / termCount (both optional, test diagnostics): per component the sum over the (pixel, Gaussian) terms the reference
 * adds with float atomics in arbitrary order (:699-736) of the term's MAGNITUDE CHAIN (the chain evaluated on magnitudes with
 * all-plus sums; equal to |term| where the term is a plain product, larger where the chain cancels internally), and the
 * number of such terms per Gaussian.  A float accumulation of k terms in any order stays within k * 2^-24 * sum|term| of the
 * exact sum, and each fp32-evaluated term carries a rounding error proportional to its magnitude chain: k * 2^-23 * this sum
 * is the per-component band the gradient tolerance of the GPU tests is tied to (oracle/parity.py). */
static void backward_impl(const OrcGaussian* gaussians, const OrcProjected* proj, const uint32_t* sortedIdx,
                  const OrcTileRange* ranges, const OrcUniforms* u, const uint32_t* lastContribIdx,
                  const float* rendered, const float* gt, int quantize,
                  OrcGradients* grads, float* tfinalOut, int threads, OrcGradients* absOut, uint32_t* termCount) {
    const uint32_t W = (uint32_t)u->screenSize[0], H = (uint32_t)u->screenSize[1];
    const uint32_t n = u->numGaussians;
    threads = set_threads(threads);
    if (threads > 32) threads = 32;
    std::vector<std::vector<GradAcc>> accs(threads), aaccs(threads);
    std::vector<std::vector<uint32_t>> cnts(threads);
    const bool wantAbs = absOut != nullptr || termCount != nullptr;
    const float fx = u->focalLength[0], fy = u->focalLength[1];
    M3 viewRot;                                           /* :558-562 */
    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) viewRot.c[c][r] = u->viewMatrix[c * 4 + r];
    const M3 viewRotT = m3_T(viewRot);

#pragma omp parallel num_threads(threads)
    {
#ifdef _OPENMP
        int tid = omp_get_thread_num();
#else
        int tid = 0;
#endif
        std::vector<GradAcc>& acc = accs[tid];
        acc.assign(n, GradAcc{});
        std::vector<GradAcc>& aacc = aaccs[tid];
        std::vector<uint32_t>& cnt = cnts[tid];
        if (wantAbs) { aacc.assign(n, GradAcc{}); cnt.assign(n, 0u); }
#pragma omp for schedule(dynamic, 1)
        for (int64_t tile = 0; tile < (int64_t)(u->numTilesX * u->numTilesY); tile++) {
            uint32_t tX = (uint32_t)tile % u->numTilesX, tY = (uint32_t)tile / u->numTilesX;
            const OrcTileRange range = ranges[tile];      /* :410-413 */
            for (uint32_t py = tY * TILE; py < std::min(H, (tY + 1) * TILE); py++)
            for (uint32_t px = tX * TILE; px < std::min(W, (tX + 1) * TILE); px++) {
                size_t pix = (size_t)py * W + px;
                uint32_t lastIdx = lastContribIdx[pix];   /* :404-407 */
                if (tfinalOut) tfinalOut[pix] = NAN;
                if (lastIdx == 0xFFFFFFFFu) continue;
                float pixx = (float)px + 0.5f, pixy = (float)py + 0.5f;   /* :416 */
                float dLdp[3];
                for (int c = 0; c < 3; c++) {             /* :418-423 */
                    float r = rendered[pix * 4 + c];
                    if (quantize) r = floorf(orc_clamp(r, 0.0f, 1.0f) * 255.0f + 0.5f) / 255.0f;
                    float diff = r - gt[pix * 4 + c];
                    float sg = (diff > 0.0f) ? 1.0f : ((diff < 0.0f) ? -1.0f : 0.0f);
                    dLdp[c] = sg / 3.0f;
                }
                uint32_t endIdx = std::min(lastIdx + 1, range.start + range.count);   /* :427 */

                float T_final = 1.0f;                     /* :430-460 pass 1 */
                for (uint32_t sortIdx = range.start; sortIdx < endIdx; sortIdx++) {
                    uint32_t gIdx = sortedIdx[sortIdx];
                    if (gIdx >= n) continue;
                    const OrcProjected& p = proj[gIdx];
                    if (p.radius <= 0) continue;
                    float dx = pixx - p.screenPos[0], dy = pixy - p.screenPos[1];
                    float power = blend_power(-0.5f * p.conic[0], -p.conic[1], -0.5f * p.conic[2], dx, dy);
                    if (!(power <= 0.0f) || power < -4.5f) continue;
                    float G = orc_exp_blend(power);
                    float alpha = fminf(p.opacity * G, ALPHA_MAX);
                    if (alpha < ALPHA_MIN) continue;
                    float test_T = T_final * (1.0f - alpha);
                    if (test_T < T_EPS) break;            /* :458 */
                    T_final = test_T;
                }
                if (tfinalOut) tfinalOut[pix] = T_final;

                float T = T_final;                        /* :464 */
                float accum_rec[3] = {1.0f, 1.0f, 1.0f};  /* :468-470 white */
                for (int64_t sortIdx = (int64_t)endIdx - 1; sortIdx >= (int64_t)range.start; sortIdx--) {   /* :473 */
                    uint32_t gIdx = sortedIdx[sortIdx];
                    if (gIdx >= n) continue;
                    const OrcProjected& p = proj[gIdx];
                    if (p.radius <= 0) continue;
                    float dx = pixx - p.screenPos[0], dy = pixy - p.screenPos[1];   /* :483 */
                    float power = blend_power(-0.5f * p.conic[0], -p.conic[1], -0.5f * p.conic[2], dx, dy);
                    if (!(power <= 0.0f) || power < -4.5f) continue;   /* :489 */
                    float G = orc_exp_blend(power);       /* :492 */
                    float alpha = fminf(p.opacity * G, ALPHA_MAX);
                    if (alpha < ALPHA_MIN) continue;      /* :495 */
                    T = T / fmaxf(1.0f - alpha, 0.0001f); /* :498 */
                    float weight = alpha * T;             /* :500 */
                    float dL_dColor[3];
                    for (int c = 0; c < 3; c++) {         /* :503-507 */
                        dL_dColor[c] = dLdp[c] * weight;
                        if (p.color[c] <= 0.01f || p.color[c] >= 0.99f) dL_dColor[c] = 0;
                    }
                    float dotv = (dLdp[0] * (p.color[0] - accum_rec[0]) + dLdp[1] * (p.color[1] - accum_rec[1]))
                                 + dLdp[2] * (p.color[2] - accum_rec[2]);
                    float dL_dAlpha = T * dotv;           /* :510 */
                    for (int c = 0; c < 3; c++)           /* :514 */
                        accum_rec[c] = alpha * p.color[c] + (1.0f - alpha) * accum_rec[c];
                    float sig = p.opacity;                /* :517-519 */
                    float dAlpha_dRawOp = sig * (1.0f - sig) * G;
                    float dL_dRawOpacity = dL_dAlpha * dAlpha_dRawOp;
                    float dL_dG = dL_dAlpha * sig;        /* :523 */
                    float gdx = G * dx, gdy = G * dy;     /* :528-529 */
                    float dG_ddelx = -gdx * p.conic[0] - gdy * p.conic[1];   /* :532-533 */
                    float dG_ddely = -gdy * p.conic[2] - gdx * p.conic[1];
                    float dLdS[2] = { dL_dG * -dG_ddelx, dL_dG * -dG_ddely };   /* :536 */
                    float z = p.depth;                    /* :540-554 */
                    float txtz = p.viewPos_xy[0] / z, tytz = p.viewPos_xy[1] / z;   /* unclamped (note I) */
                    float dLdV[3];
                    dLdV[0] = dLdS[0] * fx / z;
                    dLdV[1] = dLdS[1] * fy / z;
                    dLdV[2] = -dLdS[0] * fx * txtz / z - dLdS[1] * fy * tytz / z;
                    float dLdW[3];
                    m3_mulv(viewRotT, dLdV, dLdW);        /* :565 */
                    float dLdConic[3];                    /* :570-573 */
                    dLdConic[0] = -0.5f * dL_dG * G * dx * dx;
                    dLdConic[1] = -0.5f * dL_dG * G * 2.0f * dx * dy;
                    if (g_fd_selfcheck) dLdConic[1] = -0.5f * dL_dG * G * dx * dy;
                    dLdConic[2] = -0.5f * dL_dG * G * dy * dy;
                    float cov_a = p.cov2D[0], cov_b = p.cov2D[1], cov_c = p.cov2D[2];   /* :577-583 */
                    float denom = cov_a * cov_c - cov_b * cov_b;
                    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
                    float dLdCov[3];                      /* :588-596 */
                    dLdCov[0] = denom2inv * (-cov_c * cov_c * dLdConic[0] + 2.0f * cov_b * cov_c * dLdConic[1]
                                             + (denom - cov_a * cov_c) * dLdConic[2]);
                    dLdCov[2] = denom2inv * (-cov_a * cov_a * dLdConic[2] + 2.0f * cov_a * cov_b * dLdConic[1]
                                             + (denom - cov_a * cov_c) * dLdConic[0]);
                    dLdCov[1] = denom2inv * 2.0f * (cov_b * cov_c * dLdConic[0]
                                                    - (denom + 2.0f * cov_b * cov_b) * dLdConic[1]
                                                    + cov_a * cov_b * dLdConic[2]);
                    float tz = p.depth;                   /* :602-610 */
                    txtz = p.viewPos_xy[0] / tz; tytz = p.viewPos_xy[1] / tz;
                    float J00 = fx / tz, J02 = -fx * txtz / tz, J11 = fy / tz, J12 = -fy * tytz / tz;
                    M3 J = {{{J00, 0, 0}, {0, J11, 0}, {J02, J12, 0}}};   /* :613-617 columns */
                    M3 Tm = m3_mul(J, viewRot);           /* :620 */
                    M3 dCov2 = {{{dLdCov[0], dLdCov[1], 0}, {dLdCov[1], dLdCov[2], 0}, {0, 0, 0}}};   /* :624-628 */
                    if (g_fd_selfcheck) { dCov2.c[0][1] = 0.5f * dLdCov[1]; dCov2.c[1][0] = 0.5f * dLdCov[1]; }
                    M3 dCov3 = m3_mul(m3_mul(m3_T(Tm), dCov2), Tm);   /* :631 */
                    const OrcGaussian& go = gaussians[gIdx];   /* :635-646 */
                    float sc[3];
                    for (int k = 0; k < 3; k++) sc[k] = orc_expf(orc_clamp(go.scale[k], -MAX_SCALE, MAX_SCALE));
                    const float* q = go.rotation;         /* raw, un-normalised (note I) */
                    float rq = q[0], xq = q[1], yq = q[2], zq = q[3];
                    M3 R = quat_to_mat(q);
                    M3 S = {{{sc[0], 0, 0}, {0, sc[1], 0}, {0, 0, sc[2]}}};
                    M3 M = m3_mul(R, S);                  /* :655 */
                    M3 two = dCov3;                       /* :658  2.0f * dL_dCov3D * M */
                    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) two.c[c][r] = 2.0f * dCov3.c[c][r];
                    M3 dM = m3_mul(two, M);
                    M3 RtdM = m3_mul(m3_T(R), dM);        /* :661-663 */
                    float dLogS[3];
                    for (int k = 0; k < 3; k++) dLogS[k] = RtdM.c[k][k] * sc[k];   /* :666 */
                    M3 dR;                                /* :669-673 */
                    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) dR.c[c][r] = dM.c[c][r] * sc[c];
                    M3 D = m3_T(dR);                      /* :675  D.c[i][j] == dL_dMt_scaled[i][j] */
                    if (g_fd_selfcheck) D = dR;           /* third quirk: the transpose flips the w-coupled terms */
                    float dq[4];                          /* :678-696 */
                    dq[0] = 2.0f * (zq * (D.c[0][1] - D.c[1][0]) + yq * (D.c[2][0] - D.c[0][2]) + xq * (D.c[1][2] - D.c[2][1]));
                    dq[1] = 2.0f * (yq * (D.c[1][0] + D.c[0][1]) + zq * (D.c[2][0] + D.c[0][2]) + rq * (D.c[1][2] - D.c[2][1])
                                    - 2.0f * xq * (D.c[2][2] + D.c[1][1]));
                    dq[2] = 2.0f * (xq * (D.c[1][0] + D.c[0][1]) + rq * (D.c[2][0] - D.c[0][2]) + zq * (D.c[1][2] + D.c[2][1])
                                    - 2.0f * yq * (D.c[2][2] + D.c[0][0]));
                    dq[3] = 2.0f * (rq * (D.c[0][1] - D.c[1][0]) + xq * (D.c[2][0] + D.c[0][2]) + yq * (D.c[1][2] + D.c[2][1])
                                    - 2.0f * zq * (D.c[1][1] + D.c[0][0]));
                    GradAcc& A = acc[gIdx];               /* :699-736 */
                    A.v[0] += dLdW[0]; A.v[1] += dLdW[1]; A.v[2] += dLdW[2];
                    A.v[3] += dL_dRawOpacity;
                    A.v[4] += dLogS[0]; A.v[5] += dLogS[1]; A.v[6] += dLogS[2];
                    A.v[7] += dq[0]; A.v[8] += dq[1]; A.v[9] += dq[2]; A.v[10] += dq[3];
                    A.v[11] += dL_dColor[0] * SH_C0; A.v[12] += dL_dColor[1] * SH_C0; A.v[13] += dL_dColor[2] * SH_C0;
                    A.v[14] += dLdS[0]; A.v[15] += dLdS[1];
                    if (wantAbs) {
                        /* magnitude chain: the SAME chain with every operand replaced by its magnitude and every sum taken
                         * with all-plus signs, i.e. the quantity any fp32 evaluation of this term (the reference's included) has
                         * its rounding error proportional to.  a_* >= |*| component-wise. */
                        GradAcc& B = aacc[gIdx];
                        const float aG = fabsf(dL_dG);
                        float aS[2] = { aG * (fabsf(gdx * p.conic[0]) + fabsf(gdy * p.conic[1])), aG * (fabsf(gdy * p.conic[2]) + fabsf(gdx * p.conic[1])) };
                        float aV[3] = { aS[0] * fabsf(fx / z), aS[1] * fabsf(fy / z),
                                        aS[0] * fabsf(fx * txtz / z) + aS[1] * fabsf(fy * tytz / z) };
                        M3 aWt = m3_abs(viewRotT);
                        float aW[3]; m3_mulv(aWt, aV, aW);
                        float aC[3] = { fabsf(dLdConic[0]), fabsf(dLdConic[1]), fabsf(dLdConic[2]) };
                        float adn = fabsf(denom) + fabsf(cov_a * cov_c);
                        float aCov[3];
                        aCov[0] = denom2inv * (cov_c * cov_c * aC[0] + 2.0f * fabsf(cov_b * cov_c) * aC[1] + adn * aC[2]);
                        aCov[2] = denom2inv * (cov_a * cov_a * aC[2] + 2.0f * fabsf(cov_a * cov_b) * aC[1] + adn * aC[0]);
                        aCov[1] = denom2inv * 2.0f * (fabsf(cov_b * cov_c) * aC[0] + (fabsf(denom) + 2.0f * cov_b * cov_b) * aC[1] + fabsf(cov_a * cov_b) * aC[2]);
                        M3 aT = m3_mul(m3_abs(J), m3_abs(viewRot));
                        M3 aCov2 = {{{aCov[0], aCov[1], 0}, {aCov[1], aCov[2], 0}, {0, 0, 0}}};
                        M3 aCov3 = m3_mul(m3_mul(m3_T(aT), aCov2), aT);
                        M3 aR = m3_abs(R);
                        M3 aM = m3_mul(aR, S);
                        M3 a2 = aCov3;
                        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) a2.c[c][r] = 2.0f * aCov3.c[c][r];
                        M3 adM = m3_mul(a2, aM);
                        M3 aRtdM = m3_mul(m3_T(aR), adM);
                        float aLogS[3];
                        for (int k = 0; k < 3; k++) aLogS[k] = aRtdM.c[k][k] * sc[k];
                        M3 adR;
                        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) adR.c[c][r] = adM.c[c][r] * sc[c];
                        M3 aD = m3_T(adR);
                        const float ar = fabsf(rq), ax = fabsf(xq), ay = fabsf(yq), az = fabsf(zq);
                        float adq[4];
                        adq[0] = 2.0f * (az * (aD.c[0][1] + aD.c[1][0]) + ay * (aD.c[2][0] + aD.c[0][2]) + ax * (aD.c[1][2] + aD.c[2][1]));
                        adq[1] = 2.0f * (ay * (aD.c[1][0] + aD.c[0][1]) + az * (aD.c[2][0] + aD.c[0][2]) + ar * (aD.c[1][2] + aD.c[2][1])
                                         + 2.0f * ax * (aD.c[2][2] + aD.c[1][1]));
                        adq[2] = 2.0f * (ax * (aD.c[1][0] + aD.c[0][1]) + ar * (aD.c[2][0] + aD.c[0][2]) + az * (aD.c[1][2] + aD.c[2][1])
                                         + 2.0f * ay * (aD.c[2][2] + aD.c[0][0]));
                        adq[3] = 2.0f * (ar * (aD.c[0][1] + aD.c[1][0]) + ax * (aD.c[2][0] + aD.c[0][2]) + ay * (aD.c[1][2] + aD.c[2][1])
                                         + 2.0f * az * (aD.c[1][1] + aD.c[0][0]));
                        const float t16[16] = {aW[0], aW[1], aW[2], dL_dRawOpacity, aLogS[0], aLogS[1], aLogS[2], adq[0], adq[1], adq[2],
                                               adq[3], dL_dColor[0] * SH_C0, dL_dColor[1] * SH_C0, dL_dColor[2] * SH_C0, aS[0], aS[1]};
                        for (int k = 0; k < 16; k++) B.v[k] += fabs((double)t16[k]);
                        cnt[gIdx]++;
                    }
                }
            }
        }
    }
    memset(grads, 0, (size_t)n * sizeof(OrcGradients));   /* tiled_rasterizer.mm:687 */
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        double s[16] = {0};
        for (int t = 0; t < threads; t++) if (!accs[t].empty()) for (int k = 0; k < 16; k++) s[k] += accs[t][i].v[k];
        OrcGradients& g = grads[i];
        g.position_x = (float)s[0]; g.position_y = (float)s[1]; g.position_z = (float)s[2];
        g.opacity = (float)s[3];
        g.scale_x = (float)s[4]; g.scale_y = (float)s[5]; g.scale_z = (float)s[6];
        for (int k = 0; k < 4; k++) g.rotation[k] = (float)s[7 + k];
        g.sh[0] = (float)s[11]; g.sh[4] = (float)s[12]; g.sh[8] = (float)s[13];
        g.viewspace_grad_x = (float)s[14]; g.viewspace_grad_y = (float)s[15];
        if (wantAbs) {
            double a[16] = {0}; uint32_t c = 0;
            for (int t = 0; t < threads; t++) if (!aaccs[t].empty()) { for (int k = 0; k < 16; k++) a[k] += aaccs[t][i].v[k]; c += cnts[t][i]; }
            if (termCount) termCount[i] = c;
            if (absOut) {
                OrcGradients& o = absOut[i];
                memset(&o, 0, sizeof(o));
                o.position_x = (float)a[0]; o.position_y = (float)a[1]; o.position_z = (float)a[2];
                o.opacity = (float)a[3];
                o.scale_x = (float)a[4]; o.scale_y = (float)a[5]; o.scale_z = (float)a[6];
                for (int k = 0; k < 4; k++) o.rotation[k] = (float)a[7 + k];
                o.sh[0] = (float)a[11]; o.sh[4] = (float)a[12]; o.sh[8] = (float)a[13];
                o.viewspace_grad_x = (float)a[14]; o.viewspace_grad_y = (float)a[15];
            }
        }
    }
}
void orc_backward(const OrcGaussian* gaussians, const OrcProjected* proj, const uint32_t* sortedIdx,
                  const OrcTileRange* ranges, const OrcUniforms* u, const uint32_t* lastContribIdx,
                  const float* rendered, const float* gt, int quantize,
                  OrcGradients* grads, float* tfinalOut, int threads) {
    backward_impl(gaussians, proj, sortedIdx, ranges, u, lastContribIdx, rendered, gt, quantize, grads, tfinalOut, threads, nullptr, nullptr);
}
void orc_backward_ex(const OrcGaussian* gaussians, const OrcProjected* proj, const uint32_t* sortedIdx,
                     const OrcTileRange* ranges, const OrcUniforms* u, const uint32_t* lastContribIdx,
                     const float* rendered, const float* gt, int quantize,
                     OrcGradients* grads, float* tfinalOut, int threads, OrcGradients* absOut, uint32_t* termCount) {
    backward_impl(gaussians, proj, sortedIdx, ranges, u, lastContribIdx, rendered, gt, quantize, grads, tfinalOut, threads, absOut, termCount);
}

/* ------------------------------------------------------------------------------------------
 * A12  adamStep — shaders.metal:536-713; host side optimizer.mm:228-296.
 * Moments are the reference's 10 SoA buffers (optimizer.mm:34-60).  t is the timestep AFTER the
 * ++ in AdamOptimizer::step (optimizer.mm:237).  Runs over the live count (SURVEY note A).
 * ---------------------------------------------------------------------------------------- */
void orc_adam_step(OrcGaussian* gaussians, const OrcGradients* grads, uint32_t n,
                   float* m_position, float* m_scale, float* m_rotation, float* m_opacity, float* m_sh,
                   float* v_position, float* v_scale, float* v_rotation, float* v_opacity, float* v_sh,
                   const float lrs[5], float beta1, float beta2, float epsilon, uint32_t t, int threads) {
    threads = set_threads(threads);
    const float bc1 = 1.0f - powf(beta1, (float)t);       /* :579-580 */
    const float bc2 = 1.0f - powf(beta2, (float)t);
    const float clip = 0.5f;                              /* :582 */
    const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t tid = 0; tid < (int64_t)n; tid++) {
        const OrcGradients& g = grads[tid];
        OrcGaussian& G = gaussians[tid];
        if (std::isnan(g.position_x) || std::isnan(g.opacity) || std::isnan(g.sh[0]) ||
            std::isinf(g.position_x) || std::isinf(g.opacity)) continue;           /* :567-570 */
        if (std::isnan(G.position[0]) || std::isinf(G.position[0]) || fabsf(G.position[0]) > 1e6f) continue;   /* :573-576 */
#define ADAM_MV(gr, mo, vo, mn, vn) \
        float mn = beta1 * (mo) + omb1 * (gr); float vn = beta2 * (vo) + omb2 * (gr) * (gr);
        {   /* position :586-625 */
            float gr[3] = { orc_clamp(g.position_x, -clip, clip), orc_clamp(g.position_y, -clip, clip),
                            orc_clamp(g.position_z, -clip, clip) };
            float up[3];
            for (int k = 0; k < 3; k++) {
                ADAM_MV(gr[k], m_position[tid * 3 + k], v_position[tid * 3 + k], m, v)
                m_position[tid * 3 + k] = m; v_position[tid * 3 + k] = v;
                float m_hat = m / bc1, v_hat = v / bc2;
                up[k] = lrs[0] * m_hat / (sqrtf(v_hat) + epsilon);
            }
            float mag = sqrtf((up[0] * up[0] + up[1] * up[1]) + up[2] * up[2]);   /* :612-615 */
            if (mag > 0.1f) { float k = 0.1f / mag; for (int i = 0; i < 3; i++) up[i] = up[i] * k; }
            float np[3] = { G.position[0] - up[0], G.position[1] - up[1], G.position[2] - up[2] };
            if (!std::isnan(np[0]) && !std::isnan(np[1]) && !std::isnan(np[2]) &&
                fabsf(np[0]) < 1e6f && fabsf(np[1]) < 1e6f && fabsf(np[2]) < 1e6f) {   /* :621-624 */
                G.position[0] = np[0]; G.position[1] = np[1]; G.position[2] = np[2];
            }
        }
        {   /* scale :630-656 */
            float gr[3] = { orc_clamp(g.scale_x, -clip, clip), orc_clamp(g.scale_y, -clip, clip),
                            orc_clamp(g.scale_z, -clip, clip) };
            for (int k = 0; k < 3; k++) {
                ADAM_MV(gr[k], m_scale[tid * 3 + k], v_scale[tid * 3 + k], m, v)
                m_scale[tid * 3 + k] = m; v_scale[tid * 3 + k] = v;
                float m_hat = m / bc1, v_hat = v / bc2;
                float ns = G.scale[k] - lrs[1] * m_hat / (sqrtf(v_hat) + epsilon);
                G.scale[k] = orc_clamp(ns, -MAX_SCALE_TRAIN, MAX_SCALE_TRAIN);
            }
        }
        {   /* rotation :659-675 */
            float nr[4];
            for (int k = 0; k < 4; k++) {
                float gr = orc_clamp(g.rotation[k], -clip, clip);
                ADAM_MV(gr, m_rotation[tid * 4 + k], v_rotation[tid * 4 + k], m, v)
                m_rotation[tid * 4 + k] = m; v_rotation[tid * 4 + k] = v;
                float m_hat = m / bc1, v_hat = v / bc2;
                nr[k] = G.rotation[k] - lrs[2] * m_hat / (sqrtf(v_hat) + epsilon);
            }
            float len = sqrtf(((nr[0] * nr[0] + nr[1] * nr[1]) + nr[2] * nr[2]) + nr[3] * nr[3]);
            if (len > 0.001f) { for (int k = 0; k < 4; k++) G.rotation[k] = nr[k] / len; }
            else { G.rotation[0] = 1; G.rotation[1] = 0; G.rotation[2] = 0; G.rotation[3] = 0; }
        }
        {   /* opacity :678-692 */
            float gr = orc_clamp(g.opacity, -clip, clip);
            ADAM_MV(gr, m_opacity[tid], v_opacity[tid], m, v)
            m_opacity[tid] = m; v_opacity[tid] = v;
            float m_hat = m / bc1, v_hat = v / bc2;
            G.opacity = orc_clamp(G.opacity - lrs[3] * m_hat / (sqrtf(v_hat) + epsilon), -8.0f, 8.0f);
        }
        for (int i = 0; i < 12; i++) {   /* sh :695-712 */
            float gr = orc_clamp(g.sh[i], -clip, clip);
            size_t idx = (size_t)tid * 12 + i;
            ADAM_MV(gr, m_sh[idx], v_sh[idx], m, v)
            m_sh[idx] = m; v_sh[idx] = v;
            float m_hat = m / bc1, v_hat = v / bc2;
            float nsh = G.sh[i] - lrs[4] * m_hat / (sqrtf(v_hat) + epsilon);
            G.sh[i] = orc_clamp(nsh, -2.0f, 2.0f);
        }
#undef ADAM_MV
    }
}

/* ------------------------------------------------------------------------------------------
 * A13  DensityController::accumulateGradients — density_control.mm:121-185
 * ---------------------------------------------------------------------------------------- */
void orc_accumulate_gradients(const OrcGradients* grads, uint32_t n, float* accumGrad, uint32_t* counts) {
    for (uint32_t i = 0; i < n; i++) {
        float gx = grads[i].viewspace_grad_x, gy = grads[i].viewspace_grad_y;
        float gradMag = sqrtf(gx * gx + gy * gy);        /* :157-158 */
        gradMag = (gradMag < 1.0f) ? gradMag : ((1.0f < gradMag) ? 1.0f : gradMag);   /* std::min(gradMag, 1) :163 */
        if (!std::isnan(gradMag) && !std::isinf(gradMag) && gradMag > 0.0f) {   /* :166 */
            accumGrad[i] += gradMag;
            counts[i]++;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * A14  DensityController::apply — density_control.mm:188-501 (+ computeApproxScreenRadius :56-76).
 * libc rand() (:440-442) is replaced by a counter-based generator keyed on
 * (seed, iteration, parent index, component) — SURVEY note L; expf by orc_expf.
 * out must hold maxGaussians entries.  markers (optional, n) receives the final markers.
 * stats[3] = pruned, cloned, split.  Returns the new count.
 * ---------------------------------------------------------------------------------------- */
/* Test hook (tests/test_oracle_ref.py only): draw the split direction from libc rand() in index order, literally as
 * density_control.mm:440-442 does, instead of the counter-based generator (SURVEY note L).  With srand() called alike on
 * both sides this makes orc_densify_apply comparable bit for bit with the reference's own object code (oracle/_ref). */
static int g_split_rng_libc = 0;
void orc_set_split_rng_libc(int on) { g_split_rng_libc = on; }
static inline uint32_t split_rand_u32(uint32_t seed, uint32_t iteration, uint32_t parent, uint32_t comp) {
    uint64_t z = (((uint64_t)seed << 32) | iteration) + 0x9E3779B97F4A7C15ull * ((uint64_t)parent * 3 + comp + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (uint32_t)(z >> 32);
}
uint32_t orc_split_rand_u32(uint32_t seed, uint32_t iteration, uint32_t parent, uint32_t comp) {
    return split_rand_u32(seed, iteration, parent, comp);
}

uint32_t orc_densify_apply(const OrcGaussian* in, uint32_t n, const float* accumGrad, const uint32_t* counts,
                           uint32_t iteration, float sceneExtent, float focalLength, float imageWidth,
                           float avgDepth, uint32_t seed, uint32_t maxGaussians,
                           OrcGaussian* out, uint32_t* markersOut, uint32_t stats[3]) {
    const float GRAD_THRESHOLD = 0.0002f, OPACITY_PRUNE_THRESHOLD = 0.005f, PERCENT_DENSE = 0.01f;   /* :21-26 */
    const uint32_t DENSIFY_FROM_ITER = 500, DENSIFY_UNTIL_ITER = 15000, OPACITY_RESET_INTERVAL = 3000;   /* :29-35 */
    const float MAX_SCALE_LOG = 4.0f;                     /* :33 */
    stats[0] = stats[1] = stats[2] = 0;
    const bool canDensify = (iteration > DENSIFY_FROM_ITER && iteration < DENSIFY_UNTIL_ITER);   /* :213 */
    if (iteration >= DENSIFY_UNTIL_ITER) {                /* :216-220: unchanged */
        memcpy(out, in, (size_t)n * sizeof(OrcGaussian));
        if (markersOut) memset(markersOut, 0, (size_t)n * 4);
        return n;
    }
    const float maxScreenPixels = 40.0f;                  /* :232-233 */
    const bool enableScreenPruning = (iteration > OPACITY_RESET_INTERVAL);
    const float splitThreshold = PERCENT_DENSE * sceneExtent;   /* :246-247 */
    const float pruneThreshold = 0.1f * sceneExtent;
    std::vector<uint32_t> markers(n);
    for (uint32_t i = 0; i < n; i++) {                    /* :270-342 */
        const OrcGaussian& g = in[i];
        float opacity = 1.0f / (1.0f + orc_expf(-g.opacity));   /* :274 */
        float avgGrad = (counts[i] > 0) ? (accumGrad[i] / (float)counts[i]) : 0.0f;   /* :277 */
        float maxScaleVal = fmaxf(fmaxf(orc_expf(orc_clamp(g.scale[0], -MAX_SCALE_LOG, MAX_SCALE_LOG)),
                                        orc_expf(orc_clamp(g.scale[1], -MAX_SCALE_LOG, MAX_SCALE_LOG))),
                                  orc_expf(orc_clamp(g.scale[2], -MAX_SCALE_LOG, MAX_SCALE_LOG)));   /* :280-283 */
        bool shouldPrune = (opacity < OPACITY_PRUNE_THRESHOLD);   /* :289 */
        if (enableScreenPruning) {                        /* :295-307 */
            if (maxScaleVal > pruneThreshold) shouldPrune = true;
            float safeDepth = fmaxf(avgDepth, 0.1f);      /* :61-62 */
            float screenRadius = focalLength * maxScaleVal * 3.0f / safeDepth;   /* :72 */
            float screenFraction = screenRadius / imageWidth;   /* :75 */
            float screenRadiusPixels = screenFraction * imageWidth;   /* :303 */
            if (screenRadiusPixels > maxScreenPixels) shouldPrune = true;
        }
        if (shouldPrune) { markers[i] = 1; stats[0]++; }
        else if (canDensify && avgGrad > GRAD_THRESHOLD) {   /* :313 */
            if (maxScaleVal > splitThreshold) { markers[i] = 3; stats[2]++; }
            else { markers[i] = 2; stats[1]++; }
        } else markers[i] = 0;
    }
    uint64_t newCount = (uint64_t)n - stats[0] + stats[1] + stats[2];   /* :358 */
    if (newCount > maxGaussians) {                        /* :360-382 */
        uint64_t excess = newCount - maxGaussians;
        for (uint32_t i = 0; i < n && excess > 0; i++) if (markers[i] == 2) { markers[i] = 0; stats[1]--; excess--; }
        for (uint32_t i = 0; i < n && excess > 0; i++) if (markers[i] == 3) { markers[i] = 0; stats[2]--; excess--; }
        newCount = (uint64_t)n - stats[0] + stats[1] + stats[2];
    }
    const float logScaleFactor = -0.47000363f;            /* logf(1.0f/1.6f) :425-426 */
    uint32_t w = 0;
    for (uint32_t i = 0; i < n; i++) {                    /* :394-483 */
        const OrcGaussian& g = in[i];
        uint32_t marker = markers[i];
        if (marker == 1) continue;
        if (marker == 0) { out[w++] = g; }
        else if (marker == 2) { out[w++] = g; out[w++] = g; }   /* :408-421 identical clone, adjacent */
        else {
            float sc[3];
            for (int k = 0; k < 3; k++) sc[k] = orc_expf(orc_clamp(g.scale[k], -MAX_SCALE_LOG, MAX_SCALE_LOG));   /* :429-433 */
            float r[3];
            for (uint32_t k = 0; k < 3; k++) {            /* :440-442 with the counter-based generator */
                if (g_split_rng_libc) { r[k] = ((float)rand() / RAND_MAX - 0.5f) * 2.0f; continue; }
                uint32_t h = split_rand_u32(seed, iteration, i, k);
                r[k] = ((float)(h >> 8) * (1.0f / 16777216.0f) - 0.5f) * 2.0f;
            }
            float rNorm = sqrtf((r[0] * r[0] + r[1] * r[1]) + r[2] * r[2]);   /* :443-446 */
            if (rNorm > 0.001f) { r[0] = r[0] / rNorm; r[1] = r[1] / rNorm; r[2] = r[2] / rNorm; }
            float off[3] = { r[0] * sc[0], r[1] * sc[1], r[2] * sc[2] };   /* :448 */
            M3 R = quat_to_mat(g.rotation);               /* :452-458 raw quaternion */
            float ro[3];
            m3_mulv(R, off, ro);                          /* :459 */
            OrcGaussian c1 = g, c2 = g;
            for (int k = 0; k < 3; k++) {
                c1.position[k] = g.position[k] + ro[k];   /* :463 */
                c1.scale[k] = g.scale[k] + logScaleFactor;   /* :465-469 */
                c2.position[k] = g.position[k] - ro[k];   /* :477 */
                c2.scale[k] = c1.scale[k];
            }
            out[w++] = c1; out[w++] = c2;
        }
    }
    if (markersOut) memcpy(markersOut, markers.data(), (size_t)n * 4);
    (void)newCount;
    return w;
}

/* ------------------------------------------------------------------------------------------
 * A15  host helpers of the caller — mtl_engine.mm:637-682 (view/projection from COLMAP pose),
 * :912-924 (uniform assembly), :1039-1045 (LR decay).
 * quat = (w,x,y,z); translation t; intrinsics in pixels.
 * ---------------------------------------------------------------------------------------- */
void orc_build_uniforms(const float quat[4], const float trans[3], float fx, float fy, float cx, float cy,
                        uint32_t width, uint32_t height, float nearZ, float farZ, uint32_t numGaussians,
                        OrcUniforms* u) {
    memset(u, 0, sizeof(*u));
    M3 R = quat_to_mat(quat);                             /* :642-649 */
    float* V = u->viewMatrix;
    for (int c = 0; c < 3; c++) { for (int r = 0; r < 3; r++) V[c * 4 + r] = R.c[c][r]; V[c * 4 + 3] = 0; }
    V[12] = trans[0]; V[13] = trans[1]; V[14] = trans[2]; V[15] = 1;   /* :657 */
    float* P = u->projectionMatrix;                       /* :673-679 */
    float w = (float)width, h = (float)height;
    P[0 * 4 + 0] = 2.0f * fx / w;
    P[1 * 4 + 1] = 2.0f * fy / h;
    P[2 * 4 + 0] = 2.0f * cx / w - 1.0f;
    P[2 * 4 + 1] = 2.0f * cy / h - 1.0f;
    P[2 * 4 + 2] = farZ / (farZ - nearZ);
    P[2 * 4 + 3] = 1.0f;
    P[3 * 4 + 2] = -(farZ * nearZ) / (farZ - nearZ);
    float* VP = u->viewProjectionMatrix;                  /* :916 matrix_multiply(P, V) */
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) {
        float s = 0;
        for (int k = 0; k < 4; k++) s += P[k * 4 + r] * V[c * 4 + k];
        VP[c * 4 + r] = s;
    }
    u->screenSize[0] = w; u->screenSize[1] = h;           /* :917-918 */
    u->focalLength[0] = fx; u->focalLength[1] = fy;
    for (int r = 0; r < 3; r++)                           /* :921-925 cameraPos = -R^T t */
        u->cameraPos[r] = -((R.c[r][0] * trans[0] + R.c[r][1] * trans[1]) + R.c[r][2] * trans[2]);
    u->numTilesX = (width + TILE - 1) / TILE;             /* tiled_rasterizer.mm:212-213,309-311 */
    u->numTilesY = (height + TILE - 1) / TILE;
    u->numGaussians = numGaussians;
}

float orc_lr_decay(float lr_init, float lr_final, uint64_t iter, uint64_t maxIter) {   /* mtl_engine.mm:1039-1045 */
    if (iter >= maxIter) return lr_final;
    float t = (float)iter / (float)maxIter;
    return lr_init * powf(lr_final / lr_init, t);
}

/* MTLEngine::computeLoss (mtl_engine.mm:769-853): computeL1Loss (shaders.metal:320-340), computeSSIM (:377-466),
 * computeCombinedLoss (:470-498), reduceLoss (:343-358).  rendered / gt: RGBA float, row-major.  quantize: read the rendered
 * image through an RGBA8Unorm texture as the reference does (note G).  Returns sum / pixelCount. */
float orc_loss(const float* rendered, const float* gt, uint32_t width, uint32_t height, float lambda_dssim, int quantize) {
    const int W = (int)width, H = (int)height;
    auto rd = [&](const float* img, int x, int y, int c, bool q) {
        float v = img[((size_t)y * W + x) * 4 + c];
        if (q) v = floorf(fminf(fmaxf(v, 0.0f), 1.0f) * 255.0f + 0.5f) / 255.0f;
        return v;
    };
    const float two_sigma_sq = 2.0f * 1.5f * 1.5f;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    double total = 0.0;                                       /* the reference adds 1024 float partial sums atomically */
#pragma omp parallel for reduction(+ : total) schedule(static)
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            float mu_x = 0.f, mu_y = 0.f, wsum = 0.f;
            for (int dy = -5; dy <= 5; dy++)
                for (int dx = -5; dx <= 5; dx++) {
                    const int px = std::min(std::max(x + dx, 0), W - 1), py = std::min(std::max(y + dy, 0), H - 1);
                    const float w = expf(-(float)(dx * dx + dy * dy) / two_sigma_sq);
                    wsum += w;
                    const float xv = (rd(rendered, px, py, 0, quantize) + rd(rendered, px, py, 1, quantize) + rd(rendered, px, py, 2, quantize)) / 3.0f;
                    const float yv = (rd(gt, px, py, 0, false) + rd(gt, px, py, 1, false) + rd(gt, px, py, 2, false)) / 3.0f;
                    mu_x += w * xv; mu_y += w * yv;
                }
            mu_x /= wsum; mu_y /= wsum;
            float sxx = 0.f, syy = 0.f, sxy = 0.f;
            wsum = 0.f;
            for (int dy = -5; dy <= 5; dy++)
                for (int dx = -5; dx <= 5; dx++) {
                    const int px = std::min(std::max(x + dx, 0), W - 1), py = std::min(std::max(y + dy, 0), H - 1);
                    const float w = expf(-(float)(dx * dx + dy * dy) / two_sigma_sq);
                    wsum += w;
                    const float xv = (rd(rendered, px, py, 0, quantize) + rd(rendered, px, py, 1, quantize) + rd(rendered, px, py, 2, quantize)) / 3.0f;
                    const float yv = (rd(gt, px, py, 0, false) + rd(gt, px, py, 1, false) + rd(gt, px, py, 2, false)) / 3.0f;
                    const float a = xv - mu_x, b = yv - mu_y;
                    sxx += w * a * a; syy += w * b * b; sxy += w * a * b;
                }
            sxx /= wsum; syy /= wsum; sxy /= wsum;
            const float num = (2.0f * mu_x * mu_y + C1) * (2.0f * sxy + C2);
            const float den = (mu_x * mu_x + mu_y * mu_y + C1) * (sxx + syy + C2);
            const float dssim = fminf(fmaxf((1.0f - num / den) / 2.0f, 0.0f), 1.0f);
            const float l1 = (fabsf(rd(rendered, x, y, 0, quantize) - rd(gt, x, y, 0, false)) + fabsf(rd(rendered, x, y, 1, quantize) - rd(gt, x, y, 1, false)) +
                              fabsf(rd(rendered, x, y, 2, quantize) - rd(gt, x, y, 2, false))) / 3.0f;
            total += (double)((1.0f - lambda_dssim) * l1 + lambda_dssim * dssim);
        }
    return (float)(total / ((double)W * H));
}

} /* extern "C" */
