// peer_reduce.cu — the one exchange step of view-parallel training: the sum of the packed gradient buffers over
// the ranks of a node, done by this library's own kernels over NVLink peer memory instead of a library collective.
//
// Every rank maps every other rank's gradient buffer and flag words through CUDA IPC once (gs_dp_connect).  One call then is
//   barrier A   every rank's gradients are complete (its per-Gaussian chain kernel ran earlier on the same stream)
//   reduce      rank r sums slice r of the 16 planes over ALL ranks, in rank order, reading the peers' buffers directly
//               (16-byte system-scope loads that bypass L1) and writes the sum into its own buffer
//   barrier B   all slices are reduced
//   gather      every rank copies the other slices from their owners
//   barrier C   nobody still reads a buffer that the next step is about to overwrite
// Only the owner of a slice adds, and always in rank order, so every rank ends with bit-identical sums (the replicas of
// the model must not drift).  A barrier is one small kernel: thread t stores the call's epoch into rank t's flag word
// (release, system scope) and spins on its own word from rank t (acquire) — bounded: after 5 s it raises an error word and
// leaves, so a lost peer cannot hang the device.  Traffic per rank: 2 (g-1)/g of the payload over NVLink, both directions busy.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

__device__ __forceinline__ float4 ld_sys_f4(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void dp_barrier_kernel(const __grid_constant__ GsPeerSet P, int slot, uint32_t epoch, uint32_t* __restrict__ err) {
    const int t = (int)threadIdx.x;
    if (t >= P.world) return;
    __threadfence_system();
    st_release_sys(P.flags[t] + slot * GS_MAX_PEERS + P.rank, epoch);                  // tell rank t that this rank is here
    const uint32_t* mine = P.flags[P.rank] + slot * GS_MAX_PEERS + t;                  // wait for rank t
    const unsigned long long t0 = globaltimer_ns();
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
        if (globaltimer_ns() - t0 > 5000000000ull) { atomicExch(err, 1u); break; }     // 5 s: a peer is gone
        __nanosleep(64);
    }
}

// vector v of the flattened payload: plane v / n4, float4 index v % n4 inside the plane
__device__ __forceinline__ size_t vec_offset(uint32_t v, uint32_t n4, uint32_t stride4) { return (size_t)(v / n4) * stride4 + (v % n4); }

__global__ void __launch_bounds__(256) dp_reduce_kernel(const __grid_constant__ GsPeerSet P, uint32_t n4, uint32_t stride4, uint32_t v0, uint32_t v1) {
    float4* own = reinterpret_cast<float4*>(P.bufs[P.rank]);
    for (uint32_t v = v0 + blockIdx.x * blockDim.x + threadIdx.x; v < v1; v += gridDim.x * blockDim.x) {
        const size_t o = vec_offset(v, n4, stride4);
        float4 x[GS_MAX_PEERS];
#pragma unroll
        for (int q = 0; q < GS_MAX_PEERS; q++)
            if (q < P.world) x[q] = (q == P.rank) ? own[o] : ld_sys_f4(reinterpret_cast<const float4*>(P.bufs[q]) + o);
        float4 s = x[0];
#pragma unroll
        for (int q = 1; q < GS_MAX_PEERS; q++)
            if (q < P.world) { s.x = __fadd_rn(s.x, x[q].x); s.y = __fadd_rn(s.y, x[q].y); s.z = __fadd_rn(s.z, x[q].z); s.w = __fadd_rn(s.w, x[q].w); }
        own[o] = s;
    }
}

__global__ void __launch_bounds__(256) dp_gather_kernel(const __grid_constant__ GsPeerSet P, uint32_t n4, uint32_t stride4, uint32_t V) {
    float4* own = reinterpret_cast<float4*>(P.bufs[P.rank]);
    const uint32_t per = (V + (uint32_t)P.world - 1u) / (uint32_t)P.world;
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x) {
        const int owner = (int)(v / per);
        if (owner == P.rank) continue;
        const size_t o = vec_offset(v, n4, stride4);
        own[o] = ld_sys_f4(reinterpret_cast<const float4*>(P.bufs[owner]) + o);
    }
}

}  // namespace

void gs_launch_peer_allreduce(const GsPeerSet& P, uint32_t n, uint32_t stride, uint32_t epoch, uint32_t* err, cudaStream_t st) {
    const uint32_t n4 = (n + 3u) / 4u, stride4 = stride / 4u;
    const uint32_t V = 16u * n4;
    const uint32_t per = (V + (uint32_t)P.world - 1u) / (uint32_t)P.world;
    const uint32_t v0 = std::min(V, per * (uint32_t)P.rank), v1 = std::min(V, v0 + per);
    const int grid = 148 * 8;
    dp_barrier_kernel<<<1, 32, 0, st>>>(P, 0, epoch, err);
    if (v1 > v0) dp_reduce_kernel<<<grid, 256, 0, st>>>(P, n4, stride4, v0, v1);
    dp_barrier_kernel<<<1, 32, 0, st>>>(P, 1, epoch, err);
    dp_gather_kernel<<<grid, 256, 0, st>>>(P, n4, stride4, V);
    dp_barrier_kernel<<<1, 32, 0, st>>>(P, 2, epoch, err);
}
