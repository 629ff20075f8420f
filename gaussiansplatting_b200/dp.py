"""View-parallel data parallelism: host-side logic of the N > 1 path.

The reference trains one view per step on one device (mtl_engine.mm:1085-1093).  The multi-GPU step is defined as:
every rank holds a replica of the Gaussians, the Adam moments and the densify statistics; the views of a step are
dealt round-robin over the ranks; each rank accumulates its views' gradients into its packed gradient buffer
(GS_BWD_ACCUMULATE); ONE sum all-reduce of that buffer (16 fp32 planes: 14 parameter gradients + the two densify
statistics, which are per-view quantities and therefore ride as their own columns); then the identical Adam step and
identical densify decisions on every rank (the split RNG is counter-based).  Render-only work shards views with no
communication.

The collective itself is torch.distributed (NCCL over NVLink on the GPUs; gloo in the CPU tests) on a zero-copy view
of the device buffer — the only exchange step on this path.
"""
import numpy as np

PLANES = 16          # packed gradient planes, include/gs_abi.h gs_grads_packed
PLANE_OF = {"position": (0, 3), "opacity": (3, 1), "scale": (4, 3), "rotation": (7, 4), "sh_dc": (11, 3), "stat": (14, 2)}


def views_for_rank(step, views_per_step, world, rank, num_views):
    """indices (into the view pool) this rank renders at `step`: view slot s of the step goes to rank s % world"""
    first = step * views_per_step
    return [(first + s) % num_views for s in range(views_per_step) if s % world == rank]


def render_views_for_rank(num_views, world, rank):
    """render-only sharding (BASELINE config 4): view v -> rank v % world, no communication"""
    return list(range(rank, num_views, world))


def pack_gradients(grads, per_view_stats):
    """numpy mirror of the device's packed layout for one view's GaussianGradients (tests and host tooling only).
    per_view_stats: columns 14/15 = min(|viewspace|, 1) and 1 where that is finite and > 0 (density_control.mm:157-168)
    instead of the raw view-space x, y."""
    n = len(grads)
    out = np.zeros((PLANES, n), np.float32)
    out[0:3] = grads["position"].T
    out[3] = grads["opacity"]
    out[4:7] = grads["scale"].T
    out[7:11] = grads["rotation"].T
    out[11:14] = grads["sh"][:, [0, 4, 8]].T
    vx, vy = grads["viewspace"][:, 0], grads["viewspace"][:, 1]
    if per_view_stats:
        mag = np.minimum(np.sqrt(vx * vx + vy * vy, dtype=np.float32), np.float32(1.0))
        ok = np.isfinite(mag) & (mag > 0)
        out[14] = np.where(ok, mag, 0)
        out[15] = ok.astype(np.float32)
    else:
        out[14], out[15] = vx, vy
    return out


def unpack_gradients(packed, dtype):
    """inverse of pack_gradients (columns 14/15 are returned in `viewspace` as they are)"""
    n = packed.shape[1]
    g = np.zeros(n, dtype)
    g["position"] = packed[0:3].T
    g["opacity"] = packed[3]
    g["scale"] = packed[4:7].T
    g["rotation"] = packed[7:11].T
    g["sh"][:, [0, 4, 8]] = packed[11:14].T
    g["viewspace"] = packed[14:16].T
    return g


class CudaArrayView:
    """zero-copy __cuda_array_interface__ view of a raw device pointer"""
    def __init__(self, ptr, nfloats):
        self.__cuda_array_interface__ = {"shape": (nfloats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


class ViewParallel:
    """one per process/GPU; wraps the per-step gradient all-reduce"""
    def __init__(self, dist=None):
        self.dist = dist
        self.world = dist.get_world_size() if dist is not None else 1
        self.rank = dist.get_rank() if dist is not None else 0

    def views(self, step, views_per_step, num_views):
        return views_for_rank(step, views_per_step, self.world, self.rank, num_views)

    def allreduce_(self, tensor):
        """in-place sum over ranks of a torch tensor (a CPU tensor under gloo, a device view under nccl)"""
        if self.dist is not None and self.world > 1:
            self.dist.all_reduce(tensor, op=self.dist.ReduceOp.SUM)
        return tensor

    def device_tensor(self, grads_handle):
        """torch view of a gs.Grads packed buffer (device memory is not copied)"""
        import torch
        ptr, nfl, _ = grads_handle.packed()
        return torch.as_tensor(CudaArrayView(ptr, nfl), device="cuda")


class SymmetricGrads:
    """The packed gradient buffer of view-parallel training in SYMMETRIC memory: every rank of the node maps every other rank's
    buffer and the NVSwitch maps all of them at one multicast address, so that gs_dp_allreduce_grads can reduce in the
    switch (multimem.ld_reduce / multimem.st, csrc/peer_reduce.cu).  torch.distributed._symmetric_memory does the plumbing
    (allocation + handle exchange); the kernels are this library's.  use_multicast=False keeps the rank-ordered peer kernel
    on the same pointers."""
    FLAG_FLOATS = 64                      # 24 u32 flag words per rank live behind the 16 planes

    def __init__(self, gs, ctx, dist, capacity, device, use_multicast=True):
        import torch
        import torch.distributed._symmetric_memory as symm_mem
        self.capacity = (int(capacity) + 63) // 64 * 64
        self.tensor = symm_mem.empty(16 * self.capacity + self.FLAG_FLOATS, dtype=torch.float32, device=torch.device("cuda", device))
        self.tensor.zero_()
        torch.cuda.synchronize(device)
        self.handle = symm_mem.rendezvous(self.tensor, dist.group.WORLD.group_name)
        rank, world = dist.get_rank(), dist.get_world_size()
        bufs = [int(p) for p in self.handle.buffer_ptrs]
        assert bufs[rank] == self.tensor.data_ptr()
        flags = [p + 16 * self.capacity * 4 for p in bufs]
        mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        self.multicast = bool(mc) and use_multicast
        self.grads = gs.Grads(ctx, self.capacity, external_ptr=bufs[rank])
        self.reduce = gs.PeerReduce(ctx, rank, world)
        self.reduce.connect_ptrs(self.grads, bufs, flags, mc if self.multicast else None)
        dist.barrier()                    # every rank's flag words are zero before anybody signals

    def close(self):
        self.reduce.close(); self.grads.close()
