"""Generates tests/golden/*.npz — every intermediate of the hot path for small fixed-seed scenes.

The reference (ctaylo41/GaussianSplatting) ships no golden vectors, no tests and cannot be built on Linux (Metal +
Objective-C++), so these fixtures are dumped from oracle/ (the scalar C++ restatement of its kernels) with ONE thread
and pin (a) the oracle against accidental change and (b) the CUDA path on the GPU box without needing the oracle's
arithmetic to be re-derived there.  PARITY UNPINNED against the reference itself.

usage: python tests/golden/make_golden.py        (writes next to this file; commit the result)
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from gaussiansplatting_b200 import synth  # noqa: E402
from oracle import orc  # noqa: E402

LRS = [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3]      # mtl_engine.mm:1063-1068

# name: (n, W, H, s0, seed, view, num_views, densify iteration)
CASES = {
    "n64_64x64": (64, 64, 64, 0.05, 101, 1, 8, 700),
    "n1000_100x70": (1000, 100, 70, 0.03, 102, 3, 8, 700),          # ragged image (not a multiple of 16)
    "n2000_160x128": (2000, 160, 128, 0.02, 103, 5, 8, 3500),
}


def make_case(n, W, H, s0, seed, view, nviews, dens_iter):
    g = synth.make_scene(n, seed, s0)
    gt_g = synth.make_scene(n, seed ^ 0xA5A5, s0)
    u = synth.make_camera(view, nviews, W, H, n)
    r = orc.render(g, u, threads=1)
    gt = orc.render(gt_g, u, threads=1)["image"]
    grads, tf, gabs, gterms = orc.backward(g, r["proj"], r["svals"], r["ranges"], u, r["last"], r["image"], gt, threads=1,
                                           want_tfinal=True, want_abs=True)
    st = orc.AdamState(n)
    g1 = g.copy()
    orc.adam_step(g1, grads, st, LRS)
    accum = np.zeros(n, np.float32)
    counts = np.zeros(n, np.uint32)
    orc.accumulate_gradients(grads, accum, counts)
    # densify on amplified statistics so that clone / split / prune all occur
    accum_d = accum.copy()
    accum_d[::2] *= 50.0
    g1d = g1.copy()
    g1d["opacity"][::7] = -7.0                   # sigma < 0.005 -> pruned (density_control.mm:295-301)
    g1d["scale"][5::11] = np.log(0.5)            # large: split, and world-size prune once iteration > 3000
    dens, markers, dstats = orc.densify_apply(g1d, accum_d, counts, dens_iter, synth.scene_extent(3.0), 0.9 * W, float(W),
                                              2.0 * synth.scene_extent(3.0), 0x5EED)
    return dict(gaussians=g, gt_gaussians=gt_g, uniforms=u, proj=r["proj"], keys=r["keys"], vals=r["vals"],
                skeys=r["skeys"], svals=r["svals"], ranges=r["ranges"], image=r["image"], last=r["last"],
                evals=np.array([r["E"], r["C"]], np.uint64), gt=gt, grads=grads, tfinal=tf,
                grads_abs=gabs, grads_terms=gterms,      # per-component sum|term| and term count: the gradient tolerance band (oracle/parity.py)
                adam_gaussians=g1, dens_in=g1d, dens_iter=np.uint32(dens_iter), adam_m=np.concatenate(st.m), adam_v=np.concatenate(st.v),
                dens_accum=accum, dens_counts=counts, dens_accum_in=accum_d, dens_out=dens, dens_markers=markers,
                dens_stats=dstats)


def main():
    orc.build()
    for name, args in CASES.items():
        d = make_case(*args)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **d)
        print(name, {k: (v.shape, str(v.dtype)[:12]) for k, v in d.items() if k in ("keys", "image", "dens_out")},
              os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
