// micro-checks for packed fp32 (FFMA2/FMUL2/FADD2) on sm_100a: (1) is mul.rn.f32x2 ; add.rn.f32x2 kept unfused (bit-equal to
// __fmul_rn + __fadd_rn)?  (2) throughput of FFMA2 vs FFMA;  (3) rcp.approx(1.0f) == 1.0f ?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float2 up(u64 r) { float2 d; asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(r)); return d; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

__global__ void check(const float* x, const float* y, const float* z, int n, unsigned* bad) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float a = x[i], b = y[i], c = z[i];
    float2 r = up(add2(mul2(pk(a, b), pk(b, c)), pk(c, a)));
    float s0 = __fadd_rn(__fmul_rn(a, b), c), s1 = __fadd_rn(__fmul_rn(b, c), a);
    if (__float_as_uint(r.x) != __float_as_uint(s0) || __float_as_uint(r.y) != __float_as_uint(s1)) atomicAdd(bad, 1u);
    float2 f = up(fma2(pk(a, b), pk(b, c), pk(c, a)));
    if (__float_as_uint(f.x) != __float_as_uint(__fmaf_rn(a, b, c)) || __float_as_uint(f.y) != __float_as_uint(__fmaf_rn(b, c, a))) atomicAdd(bad + 1, 1u);
    float one = x[i] * 0.f + 1.0f, rc;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(one));
    if (rc != 1.0f) atomicAdd(bad + 2, 1u);
}
template <int MODE>
__global__ void __launch_bounds__(256) tput(float* out, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const float m = 0.999f + blockIdx.x * 1e-9f, c = 1e-3f;
    if (MODE == 0) {
        for (int i = 0; i < iters; i++) {
            a0 = __fmaf_rn(a0, m, c); a1 = __fmaf_rn(a1, m, c); a2 = __fmaf_rn(a2, m, c); a3 = __fmaf_rn(a3, m, c);
            a4 = __fmaf_rn(a4, m, c); a5 = __fmaf_rn(a5, m, c); a6 = __fmaf_rn(a6, m, c); a7 = __fmaf_rn(a7, m, c);
        }
    } else {
        u64 p0 = pk(a0, a1), p1 = pk(a2, a3), p2 = pk(a4, a5), p3 = pk(a6, a7), mm = pk(m, m), cc = pk(c, c);
        for (int i = 0; i < iters; i++) {
            p0 = fma2(p0, mm, cc); p1 = fma2(p1, mm, cc); p2 = fma2(p2, mm, cc); p3 = fma2(p3, mm, cc);
        }
        float2 r0 = up(p0), r1 = up(p1), r2 = up(p2), r3 = up(p3);
        a0 = r0.x; a1 = r0.y; a2 = r1.x; a3 = r1.y; a4 = r2.x; a5 = r2.y; a6 = r3.x; a7 = r3.y;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
// MODE 2/3: FP work mixed 1:1 with integer ALU work (does packing free issue slots for the other pipe?)
template <int MODE>
__global__ void __launch_bounds__(256) mixed(float* out, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
    unsigned q0 = threadIdx.x, q1 = q0 * 3, q2 = q0 * 5, q3 = q0 * 7;
    const float m = 0.999f + blockIdx.x * 1e-9f, c = 1e-3f;
    u64 p0 = pk(a0, a1), p1 = pk(a2, a3), mm = pk(m, m), cc = pk(c, c);
    for (int i = 0; i < iters; i++) {
        if (MODE == 2) { a0 = __fmaf_rn(a0, m, c); a1 = __fmaf_rn(a1, m, c); a2 = __fmaf_rn(a2, m, c); a3 = __fmaf_rn(a3, m, c); }
        else { p0 = fma2(p0, mm, cc); p1 = fma2(p1, mm, cc); }
        q0 = (q0 ^ (q0 >> 3)) + i; q1 = (q1 ^ (q1 >> 5)) + i; q2 = (q2 ^ (q2 >> 7)) + i; q3 = (q3 ^ (q3 >> 9)) + i;
    }
    if (MODE == 3) { float2 r0 = up(p0), r1 = up(p1); a0 = r0.x; a1 = r0.y; a2 = r1.x; a3 = r1.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + (float)(q0 + q1 + q2 + q3);
}
int main() {
    const int n = 1 << 22;
    float *x, *y, *z; unsigned* bad;
    cudaMallocManaged(&x, n * 4); cudaMallocManaged(&y, n * 4); cudaMallocManaged(&z, n * 4); cudaMallocManaged(&bad, 16);
    uint32_t s = 12345;
    for (int i = 0; i < n; i++) {
        s = s * 1664525u + 1013904223u; x[i] = (float)(s >> 8) / 16777216.0f * 4.f - 2.f;
        s = s * 1664525u + 1013904223u; y[i] = (float)(s >> 8) / 16777216.0f * 4.f - 2.f;
        s = s * 1664525u + 1013904223u; z[i] = (float)(s >> 8) / 16777216.0f * 1e-3f;
    }
    bad[0] = bad[1] = bad[2] = 0;
    check<<<n / 256, 256>>>(x, y, z, n, bad);
    cudaDeviceSynchronize();
    printf("mul2+add2 vs scalar rn mismatches: %u / %d ; fma2 vs fmaf_rn mismatches: %u ; rcp(1)!=1: %u\n", bad[0], n, bad[1], bad[2]);
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 1 << 16;
    for (int mode = 0; mode < 4; mode++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) tput<0><<<148 * 8, 256>>>(out, iters);
            else if (mode == 1) tput<1><<<148 * 8, 256>>>(out, iters);
            else if (mode == 2) mixed<2><<<148 * 8, 256>>>(out, iters);
            else mixed<3><<<148 * 8, 256>>>(out, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double fmas = (double)148 * 8 * 256 * iters * (mode < 2 ? 8 : 4);
            if (rep) printf("mode %d (%s): %.3f ms, %.1f TFLOP/s fp32\n", mode,
                            mode == 0 ? "FFMA x8" : mode == 1 ? "FFMA2 x4" : mode == 2 ? "FFMA x4 + 12 int ops" : "FFMA2 x2 + 12 int ops", ms, 2 * fmas / ms * 1e-9);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
