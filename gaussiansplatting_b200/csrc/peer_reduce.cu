// peer_reduce.cu — the one exchange step of view-parallel training: the sum of the packed gradient buffers over
// the ranks of a node, done by this library's own kernels over NVLink peer memory instead of a library collective.
//
// Every rank maps every other rank's gradient buffer and flag words through CUDA IPC once (gs_dp_connect).  One call then is
//   barrier A   every rank's gradients are complete (its per-Gaussian chain kernel ran earlier on the same stream)
//   reduce +    rank r sums slice r of the 16 planes over ALL ranks, in rank order, reading the peers' buffers directly
//   broadcast   (16-byte system-scope loads that bypass L1), and stores the sum into slice r of EVERY rank's buffer (posted
//               writes: reads and writes keep both directions of the links busy).  Slice r of any buffer is touched by rank r
//               only — read, then overwritten by the same thread — so no rank ever sees a half-updated slice
//   barrier B   all sums have landed everywhere; nobody reads or writes a buffer that the next step is about to overwrite
// Only the owner of a slice adds, and always in rank order, so every rank ends with bit-identical sums (the replicas of
// the model must not drift).  A barrier is one small kernel: thread t stores the call's epoch into rank t's flag word
// (release, system scope) and spins on its own word from rank t (acquire) — bounded: after 5 s it raises an error word and
// leaves, so a lost peer cannot hang the device.  Traffic per rank: (g-1)/g of the payload in each direction of its links.
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

// err: device word the kernels of this stream read; host_err: the same flag in mapped pinned host memory, which the host polls
// without a stream synchronisation (gs_dp_allreduce_grads / gs_dp_check return GS_ERR_NCCL once it is set).  Both are sticky.
__global__ void dp_barrier_kernel(const __grid_constant__ GsPeerSet P, int slot, uint32_t epoch, uint32_t* __restrict__ err,
                                  volatile uint32_t* __restrict__ host_err) {
    const int t = (int)threadIdx.x;
    if (t >= P.world) return;
    __threadfence_system();
    st_release_sys(P.flags[t] + slot * GS_MAX_PEERS + P.rank, epoch);                  // tell rank t that this rank is here
    const uint32_t* mine = P.flags[P.rank] + slot * GS_MAX_PEERS + t;                  // wait for rank t
    if (*reinterpret_cast<volatile uint32_t*>(err)) return;   // already failed (sticky): the peers are told, nobody is waited for
    const unsigned long long t0 = globaltimer_ns();
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
        if (globaltimer_ns() - t0 > 5000000000ull) {                                   // 5 s: a peer is gone
            atomicExch(err, 1u);
            if (host_err) { *host_err = 1u; __threadfence_system(); }
            break;
        }
        __nanosleep(64);
    }
}

// vector v of the flattened payload: plane v / n4, float4 index v % n4 inside the plane
__device__ __forceinline__ size_t vec_offset(uint32_t v, uint32_t n4, uint32_t stride4) { return (size_t)(v / n4) * stride4 + (v % n4); }

template <int WORLD>
__global__ void __launch_bounds__(256) dp_reduce_kernel(const __grid_constant__ GsPeerSet P, uint32_t n4, uint32_t stride4, uint32_t v0, uint32_t v1,
                                                        const uint32_t* __restrict__ err) {
    constexpr int U = WORLD <= 2 ? 4 : 2;                       // vectors per thread per round: U * (WORLD - 1) remote loads in flight
    const float4* own = reinterpret_cast<const float4*>(P.bufs[P.rank]);
    const uint32_t T = gridDim.x * blockDim.x;
    if (*err) {
        // A barrier of this or an earlier call timed out: the peers' buffers may be incomplete, so nothing is summed and no
        // peer is written.  This rank's position-x gradients (plane 0) become NaN instead, which makes adamStep's guard
        // (shaders.metal:567-570) skip every Gaussian: the step is lost on this rank, never applied with a partial sum, and the
        // host sees GS_ERR_NCCL on its next gs_dp_allreduce_grads / gs_dp_check.
        float4* mine = reinterpret_cast<float4*>(P.bufs[P.rank]);
        const float qn = __int_as_float(0x7FC00000);
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += T) mine[i] = make_float4(qn, qn, qn, qn);
        return;
    }
    for (uint32_t vb = v0 + blockIdx.x * blockDim.x + threadIdx.x; vb < v1; vb += U * T) {
        float4 x[U][WORLD];
        size_t o[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t v = vb + (uint32_t)u * T;
            o[u] = vec_offset(v < v1 ? v : vb, n4, stride4);
#pragma unroll
            for (int q = 0; q < WORLD; q++)
                x[u][q] = (q == P.rank) ? own[o[u]] : ld_sys_f4(reinterpret_cast<const float4*>(P.bufs[q]) + o[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (vb + (uint32_t)u * T >= v1) break;
            float4 s = x[u][0];
#pragma unroll
            for (int q = 1; q < WORLD; q++) { s.x = __fadd_rn(s.x, x[u][q].x); s.y = __fadd_rn(s.y, x[u][q].y); s.z = __fadd_rn(s.z, x[u][q].z); s.w = __fadd_rn(s.w, x[u][q].w); }
#pragma unroll
            for (int q = 0; q < WORLD; q++) reinterpret_cast<float4*>(P.bufs[q])[o[u]] = s;
        }
    }
}

// NVLS variant: the NVSwitch adds.  Rank r owns slice r: multimem.ld_reduce on the multicast address returns the sum of the
// g copies of a 16-byte vector (one switch-side reduction, fp32 accumulate), multimem.st writes it to all g copies.  Every
// element is reduced exactly once and the one result is broadcast, so all ranks hold bit-identical sums; the order of the
// additions inside the switch is the hardware's (the rank-ordered kernel above stays the mode whose order is specified).
template <int U>
__global__ void __launch_bounds__(512) dp_reduce_mc_kernel(float4* __restrict__ mc, float4* __restrict__ own, uint32_t n4, uint32_t stride4,
                                                           uint32_t v0, uint32_t v1, const uint32_t* __restrict__ err) {
    const uint32_t T = gridDim.x * blockDim.x;
    if (*err) {                                                 // see dp_reduce_kernel: withhold the sum, poison the step
        const float qn = __int_as_float(0x7FC00000);
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += T) own[i] = make_float4(qn, qn, qn, qn);
        return;
    }
    for (uint32_t vb = v0 + blockIdx.x * blockDim.x + threadIdx.x; vb < v1; vb += U * T) {
        float4 x[U];
        float4* p[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t v = vb + (uint32_t)u * T;
            p[u] = mc + vec_offset(v < v1 ? v : vb, n4, stride4);
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(x[u].x), "=f"(x[u].y), "=f"(x[u].z), "=f"(x[u].w) : "l"(p[u]) : "memory");
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (vb + (uint32_t)u * T >= v1) break;
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                         :: "l"(p[u]), "f"(x[u].x), "f"(x[u].y), "f"(x[u].z), "f"(x[u].w) : "memory");
        }
    }
}

}  // namespace

// launch shape of the in-switch reduce (gs_tune "dp_mc_grid" / "dp_mc_threads" / "dp_mc_unroll").  Swept on 8 B200 by
// tools/allreduce_probe.py (profiles/r2_allreduce_probe_8gpu.txt): 64 MB in 158-161 us for every shape with 16-148 CTAs, but
// 490-770 us with 592-1184 CTAs x 8 vectors in flight -- the switch's multimem queues congest, so FEW requests in flight win:
// 64 CTAs x 512 threads x 1 vector (512 KiB in flight per GPU).
static int g_mc_grid = 64, g_mc_threads = 512, g_mc_unroll = 1;
void gs_tune_dp_mc(int which, int v) {
    if (which == 0 && v >= 1 && v <= 148 * 16) g_mc_grid = v;
    if (which == 1 && v >= 32 && v <= 512 && v % 32 == 0) g_mc_threads = v;
    if (which == 2 && (v == 1 || v == 2 || v == 4 || v == 8)) g_mc_unroll = v;
}

void gs_launch_peer_allreduce(const GsPeerSet& P, float* mc, uint32_t n, uint32_t stride, uint32_t epoch, uint32_t* err, uint32_t* host_err,
                              cudaStream_t st) {
    const uint32_t n4 = (n + 3u) / 4u, stride4 = stride / 4u;
    const uint32_t V = 16u * n4;
    const uint32_t per = (V + (uint32_t)P.world - 1u) / (uint32_t)P.world;
    const uint32_t v0 = std::min(V, per * (uint32_t)P.rank), v1 = std::min(V, v0 + per);
    const int grid = 148 * 8;
    dp_barrier_kernel<<<1, 32, 0, st>>>(P, 0, epoch, err, host_err);
    if (v1 > v0 && mc) {
        float4* m4 = reinterpret_cast<float4*>(mc); float4* o4 = reinterpret_cast<float4*>(P.bufs[P.rank]);
        const int g = g_mc_grid, t = g_mc_threads;
        if (g_mc_unroll == 1) dp_reduce_mc_kernel<1><<<g, t, 0, st>>>(m4, o4, n4, stride4, v0, v1, err);
        else if (g_mc_unroll == 2) dp_reduce_mc_kernel<2><<<g, t, 0, st>>>(m4, o4, n4, stride4, v0, v1, err);
        else if (g_mc_unroll == 8) dp_reduce_mc_kernel<8><<<g, t, 0, st>>>(m4, o4, n4, stride4, v0, v1, err);
        else dp_reduce_mc_kernel<4><<<g, t, 0, st>>>(m4, o4, n4, stride4, v0, v1, err);
    } else if (v1 > v0) {
        switch (P.world) {
#define GS_DP_CASE(W) case W: dp_reduce_kernel<W><<<grid, 256, 0, st>>>(P, n4, stride4, v0, v1, err); break;
            GS_DP_CASE(2) GS_DP_CASE(3) GS_DP_CASE(4) GS_DP_CASE(5) GS_DP_CASE(6) GS_DP_CASE(7) GS_DP_CASE(8)
#undef GS_DP_CASE
        default: break;
        }
    }
    dp_barrier_kernel<<<1, 32, 0, st>>>(P, 1, epoch, err, host_err);
}
