"""Parity bookkeeping shared by tests/, __graft_entry__.smoke() and bench.py's parity leg (TEST INFRASTRUCTURE ONLY).

Gradient bar (BASELINE.json north_star: "gradients must match to 1e-3 relative, since atomic ordering differs"):

    |a - b| <= 1e-3 * max(|a|, |b|)  +  k * 2^-23 * sum|term|

a = this library's component, b = the oracle's (accumulated in double), and the second summand is the component's own
accumulation band: the reference adds k per-(pixel, Gaussian) terms with relaxed float atomics in arbitrary order
(tiled_shaders.metal:699-736); any such order stays within k * 2^-24 * sum|term| of the exact sum, and two different
orders within twice that.  k and sum|term| come from the oracle itself (orc.backward(want_abs=True)); "|term|" is the
term's magnitude chain (gs_oracle.cpp: the same chain on magnitudes with all-plus sums), which equals |term| for the
plain products (opacity, colour) and exceeds it where the chain itself cancels (conic -> cov2D -> cov3D -> scale /
quaternion: median 20-50 x), because an fp32 evaluation of such a term -- the reference's own included -- is only
accurate relative to that magnitude.  A component without cancellation has sum|term| = |b| and k * 2^-23 <= 1e-3 for
k < 8388, so the band only matters where terms cancel.  There is no group-wide floor.
grad_report() also states, per parameter group, the worst PURE relative error |a-b| / max(|a|,|b|), the fraction of
components above 1e-3 pure relative, and the fraction that pass only thanks to the band.
"""
import numpy as np

GROUPS = ("position", "opacity", "scale", "rotation", "sh", "viewspace")
RTOL = 1e-3
ULP = 2.0 ** -23


def grad_report(got, ref, absref, count, groups=GROUPS, rtol=RTOL):
    """per group: n (finite non-zero components), worst_pure_rel, frac_over_rtol_pure, frac_need_band, n_fail, worst_excess
    (largest |a-b| - tolerance, <= 0 when everything passes), nonfinite_mismatch"""
    rep = {}
    k = count.astype(np.float64)
    for f in groups:
        a, b, s = got[f].astype(np.float64), ref[f].astype(np.float64), absref[f].astype(np.float64)
        kk = k.reshape((-1,) + (1,) * (a.ndim - 1))
        fin = np.isfinite(b) & np.isfinite(s)
        # components the oracle itself overflows on (giant splats, denom^2 > FLT_MAX) must be non-finite here too
        bad_nf = int(np.count_nonzero(np.isfinite(a[~fin]))) if f != "viewspace" else 0
        bad_nf += int(np.count_nonzero(~np.isfinite(a[fin])))
        ok = fin & np.isfinite(a)
        a, b, s, kk = a[ok], b[ok], np.broadcast_to(s, got[f].shape)[ok], np.broadcast_to(kk, got[f].shape)[ok]
        mag = np.maximum(np.abs(a), np.abs(b))
        nz = mag > 0
        diff = np.abs(a - b)
        pure = np.zeros_like(diff)
        pure[nz] = diff[nz] / mag[nz]
        band = kk * ULP * s
        excess = diff - (rtol * mag + band)
        over = pure > rtol
        n = int(np.count_nonzero(nz))
        rep[f] = {"n": n, "worst_pure_rel": float(pure.max()) if pure.size else 0.0,
                  "frac_over_rtol_pure": float(np.count_nonzero(over)) / max(n, 1),
                  "frac_need_band": float(np.count_nonzero(over & (excess <= 0))) / max(n, 1),
                  "n_fail": int(np.count_nonzero(excess > 0)), "worst_excess": float(excess.max()) if excess.size else 0.0,
                  "nonfinite_mismatch": bad_nf}
    return rep


def grads_ok(rep):
    return all(r["n_fail"] == 0 and r["nonfinite_mismatch"] == 0 for r in rep.values())


def assert_grads(got, ref, absref, count, groups=GROUPS, rtol=RTOL):
    rep = grad_report(got, ref, absref, count, groups, rtol)
    assert grads_ok(rep), {f: r for f, r in rep.items() if r["n_fail"] or r["nonfinite_mismatch"]}
    return rep


def frame_parity(gs, R, image, ref, grads=None, ref_grads=None, absref=None, count=None, what=""):
    """compare the state a gs.Rasterizer forward (+ backward) left behind with an oracle frame (orc.render dict).
    keys / lists / ranges / lastContribIdx bit-exact, image max-abs, gradients by grad_report()"""
    out = {"config": what}
    out["pairs"] = int(R.pair_count)
    out["pairs_equal"] = bool(R.pair_count == len(ref["skeys"]))
    out["keys_equal"] = bool(out["pairs_equal"] and np.array_equal(R.get(gs.RAST_SORTED_KEYS), ref["skeys"]))
    out["lists_equal"] = bool(out["pairs_equal"] and np.array_equal(R.get(gs.RAST_SORTED_VALUES), ref["svals"]))
    rg = R.get(gs.RAST_TILE_RANGES)
    out["ranges_equal"] = bool(np.array_equal(rg["start"], ref["ranges"]["start"]) and np.array_equal(rg["count"], ref["ranges"]["count"]))
    out["last_equal"] = bool(np.array_equal(R.get(gs.RAST_LAST_IDX), ref["last"]))
    out["image_max_abs"] = float(np.abs(image - ref["image"]).max())
    out["image_bit_equal"] = bool(np.array_equal(image, ref["image"]))
    if grads is not None:
        rep = grad_report(grads, ref_grads, absref, count)
        main = [f for f in GROUPS if f != "viewspace"]
        out["grad_worst_pure_rel"] = max(rep[f]["worst_pure_rel"] for f in main)
        n = sum(rep[f]["n"] for f in main)
        out["grad_frac_over_1e-3_pure"] = sum(rep[f]["frac_over_rtol_pure"] * rep[f]["n"] for f in main) / max(n, 1)
        out["grad_frac_need_band"] = sum(rep[f]["frac_need_band"] * rep[f]["n"] for f in main) / max(n, 1)
        out["grad_n_fail"] = sum(rep[f]["n_fail"] + rep[f]["nonfinite_mismatch"] for f in GROUPS)
        out["grad_components"] = n
        out["grad_groups"] = rep
        out["grad_rule"] = "|a-b| <= 1e-3*max(|a|,|b|) + k*2^-23*sum|term| (k, sum|term| per component from the oracle)"
    out["ok"] = bool(out["keys_equal"] and out["lists_equal"] and out["ranges_equal"] and out["last_equal"] and
                     out["image_max_abs"] <= 1e-4 and (grads is None or out["grad_n_fail"] == 0))
    return out
