"""GPU parity of the radix sorts (K4), Adam (K9) and density control (A13, K10/K11) through the C ABI."""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth

pytestmark = pytest.mark.gpu


def rand_u64(seed, n):
    return synth.splitmix64(seed, n, 77)


# ---- GPURadixSort64 / GPURadixSort32 ------------------------------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 31, 1000, 4096, 4097, 100003, 1 << 20])
def test_sort64_matches_reference_cpu_sort(gs, ctx, orc, n):
    keys = rand_u64(n + 1, n)
    vals = np.arange(n, dtype=np.uint32)
    S = gs.Sort64(ctx, max(n, 1))
    try:
        k, v = S.sort_host(keys, vals)
    finally:
        S.close()
    rk, rv = orc.sort_pairs(keys, vals, threads=4)
    assert np.array_equal(k, rk) and np.array_equal(v, rv)


def test_sort64_is_stable_and_respects_bit_range(gs, ctx, orc):
    n = 300000
    keys = rand_u64(5, n) & np.uint64(0x0000000F000000FF)      # many ties -> stability matters
    vals = np.arange(n, dtype=np.uint32)
    S = gs.Sort64(ctx, n)
    try:
        k, v = S.sort_host(keys, vals)
        rk, rv = orc.sort_pairs(keys, vals)
        assert np.array_equal(k, rk) and np.array_equal(v, rv)
        # only bits [32,40): order by that digit, ties in input order
        k2, v2 = S.sort_host(keys, vals, 32, 40)
        order = np.argsort((keys >> np.uint64(32)) & np.uint64(0xFF), kind="stable")
        assert np.array_equal(v2, vals[order]) and np.array_equal(k2, keys[order])
    finally:
        S.close()


def test_sort64_checksum_at_full_size(gs, ctx):
    """BASELINE-size property test (no oracle needed): sortedness, permutation validity, key/value pairing"""
    n = 1 << 23
    keys = rand_u64(99, n)
    vals = np.arange(n, dtype=np.uint32)
    S = gs.Sort64(ctx, n)
    try:
        k, v = S.sort_host(keys, vals)
    finally:
        S.close()
    assert np.all(k[:-1] <= k[1:])
    assert np.array_equal(keys[v], k)
    assert np.array_equal(np.sort(v), vals)


def test_sort32_depth_back_to_front(gs, ctx, orc):
    n = 50000
    pos = np.zeros((n, 4), np.float32)
    pos[:, :3] = synth.make_scene(n, 3, 0.01)["position"]
    pos[7, 0] = np.nan; pos[9, 2] = np.inf
    cam = np.array([0.5, -3.0, 1.0], np.float32)
    rk, rv = orc.depth_keys32(pos, cam)
    rk, rv = orc.sort_pairs32(rk, rv)
    d = ctx.alloc(n * 16)
    S = gs.Sort32(ctx, n)
    try:
        ctx.upload(d, pos)
        S.sort_depth(d, cam, n)
        ctx.sync()
        assert np.array_equal(ctx.download(S.sorted_keys, n, np.uint32), rk)
        assert np.array_equal(ctx.download(S.sorted_indices, n, np.uint32), rv)
    finally:
        S.close(); ctx.free(d)


# ---- AdamOptimizer ----------------------------------------------------------------------------------------
def make_grads(orc_mod, n, seed):
    from gaussiansplatting_b200.layouts import GRADIENTS
    g = np.zeros(n, GRADIENTS)
    u = lambda s, k=1: (synth.uniform01(seed, n * k, s).reshape(n, k) * 2 - 1).astype(np.float32)
    g["position"] = u(1, 3) * 0.8
    g["opacity"] = u(2)[:, 0] * 0.8
    g["scale"] = u(3, 3) * 0.3
    g["rotation"] = u(4, 4) * 1e-3
    for c in range(3):
        g["sh"][:, 4 * c] = u(5 + c)[:, 0] * 1e-2
    g["viewspace"] = u(9, 2) * 1e-3
    g["position"][3, 0] = np.nan          # skipped Gaussian (shaders.metal:567-570)
    g["opacity"][5] = np.inf
    return g


def test_adam_parity_three_steps(gs, ctx, orc):
    n = 5000
    g = synth.make_scene(n, 61, 0.02)
    g["position"][11, 0] = 2e6            # corrupted Gaussian is skipped (:573-576)
    lrs = [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3]
    ref = g.copy()
    st = orc.AdamState(n)
    G = gs.Gaussians(ctx, n); G.upload(g)
    A = gs.Adam(ctx, n)
    D = gs.Grads(ctx, n)
    try:
        for step in range(3):
            gr = make_grads(orc, n, 100 + step)
            orc.adam_step(ref, gr, st, lrs)
            D.upload(gr)
            A.step(G, D, lrs)
        assert A.timestep == 3
        got = G.download()
        for f in ("position", "scale", "rotation", "opacity", "sh"):
            assert np.array_equal(got[f], ref[f], equal_nan=True), f
        for grp in range(5):
            assert np.array_equal(A.download(grp, 0, n), st.m[grp], equal_nan=True), grp
            assert np.array_equal(A.download(grp, 1, n), st.v[grp], equal_nan=True), grp
        # state helpers (optimizer.mm:137-188)
        A.reset_opacity_momentum(); A.reset_scale_momentum()
        assert not A.download(3, 0, n).any() and not A.download(1, 1, n).any()
        assert A.download(0, 0, n).any()
        A.reset_state_for_new(n - 100)
        assert not A.download(0, 0, n)[3 * (n - 100):].any() and A.download(0, 0, n)[:3 * (n - 100)].any()
        A.resize_if_needed(3 * n)
        assert np.array_equal(A.download(0, 0, n), np.concatenate([st.m[0][:3 * (n - 100)], np.zeros(300, np.float32)]))
        assert not A.download(4, 1, 3 * n)[12 * n:].any()
    finally:
        D.close(); A.close(); G.close()


def test_adam_idle_plane_skipping_stays_exact(gs, ctx, orc):
    """The library stops streaming the nine non-DC SH planes once they are known to sit inside Adam's clamp with zero
    moments, and writes nothing for a parameter whose gradient and moments are zero.  Every path that can invalidate
    that (parameter upload with out-of-range values, growth of the covered count, caller-supplied SH moments, zero
    gradients turning non-zero) must still reproduce the oracle bit for bit."""
    n = 3000
    lrs = [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3]
    g = synth.make_scene(n, 71, 0.02)
    g["sh"][::7, 1] = 3.5; g["sh"][::5, 6] = -9.0            # non-DC coefficients outside [-2, 2]: clamped by the first step
    ref = g.copy()
    st = orc.AdamState(n)
    G = gs.Gaussians(ctx, n); G.upload(g)
    A = gs.Adam(ctx, n)
    D = gs.Grads(ctx, n)

    def step(seed, zero_from=None):
        gr = make_grads(orc, n, seed)
        if zero_from is not None:                              # Gaussians no pixel touched: all-zero gradient rows
            for f in ("position", "opacity", "scale", "rotation", "sh", "viewspace"):
                gr[f][zero_from:] = 0
        orc.adam_step(ref, gr, st, lrs)
        D.upload(gr); A.step(G, D, lrs)

    def check(tag):
        got = G.download()
        for f in ("position", "scale", "rotation", "opacity", "sh"):
            assert np.array_equal(got[f], ref[f], equal_nan=True), (tag, f)
        for grp in range(5):
            assert np.array_equal(A.download(grp, 0, n), st.m[grp], equal_nan=True), (tag, grp)
            assert np.array_equal(A.download(grp, 1, n), st.v[grp], equal_nan=True), (tag, grp)
    try:
        step(1, zero_from=n // 2); check("first step clamps")
        assert np.abs(ref["sh"][20:]).max() <= 2.0 and ref["sh"][5, 6] == -9.0   # Gaussian 5 is skipped (inf gradient): not clamped
        step(2, zero_from=n // 2); step(3, zero_from=n // 3); check("idle planes skipped")
        step(4); check("idle rows wake up")
        # a skipped Gaussian keeps an out-of-range coefficient until its gradient is finite again, then it is clamped
        gr = make_grads(orc, n, 40); gr["opacity"][5] = 0.25
        orc.adam_step(ref, gr, st, lrs); D.upload(gr); A.step(G, D, lrs)
        check("skipped Gaussian comes back"); assert ref["sh"][5, 6] == -2.0
        g2 = G.download(); g2["sh"][::3, 9] = 7.0; ref["sh"][::3, 9] = 7.0
        G.upload(g2)
        step(5); check("re-upload outside the clamp")
        m = st.m[4].copy(); m[5::12] = 1e-3; st.m[4][:] = m      # caller-supplied moments on a non-DC plane
        A.upload(4, 0, m, n)
        step(6); step(7); check("uploaded SH moments")
    finally:
        D.close(); A.close(); G.close()


# ---- DensityController ------------------------------------------------------------------------------------
def test_accumulate_gradients_parity(gs, ctx, orc):
    n = 4000
    gr = make_grads(orc, n, 7)
    gr["position"][3, 0] = 0; gr["opacity"][5] = 0
    gr["viewspace"][10] = [3.0, 4.0]      # clamped to 1
    gr["viewspace"][11] = 0               # not counted
    gr["viewspace"][12, 0] = np.nan       # not counted
    accum, count = np.zeros(n, np.float32), np.zeros(n, np.uint32)
    D = gs.Grads(ctx, n); Dn = gs.Density(ctx)
    try:
        D.upload(gr)
        for _ in range(2):
            orc.accumulate_gradients(gr, accum, count)
            Dn.accumulate(D, n)
        a, c = Dn.download(n)
        assert np.array_equal(c, count) and np.array_equal(a, accum)
    finally:
        Dn.close(); D.close()


@pytest.mark.parametrize("iteration,max_g", [(700, 1500000), (3500, 1500000), (700, 2600), (700, 2000), (200, 1500000),
                                             (15000, 1500000)])
def test_densify_apply_parity(gs, ctx, orc, iteration, max_g):
    n = 2000
    g = synth.make_scene(n, 71, 0.02)
    g["scale"][:400] = np.log(0.2)         # big: split candidates
    g["scale"][:100] = np.log(0.5)         # world-space / screen-space prune once iteration > 3000
    g["opacity"][400:500] = -7.0           # pruned
    accum = (synth.uniform01(5, n, 1) * 1e-3).astype(np.float32)
    count = (synth.uniform01(5, n, 2) * 3).astype(np.uint32)
    extent, focal, width, depth, seed = 3.3, 230.0, 256.0, 6.6, 1234
    ref, rmark, rstats = orc.densify_apply(g, accum, count, iteration, extent, focal, width, depth, seed, max_g)
    G = gs.Gaussians(ctx, n); G.upload(g)
    Dn = gs.Density(ctx, max_g)
    try:
        Dn.set_scene_extent(extent); Dn.set_seed(seed)
        Dn.upload(accum, count)
        stats = Dn.apply(G, iteration, focal, width, depth)
        assert stats == tuple(int(x) for x in rstats)
        assert G.count == len(ref)
        got = G.download()
        if iteration < 15000:                               # past DENSIFY_UNTIL_ITER apply() returns before marking
            _, _, m = Dn.download(n, markers=True)
            assert np.array_equal(m, rmark)
        for f in ("position", "scale", "rotation", "opacity", "sh"):
            assert np.array_equal(got[f], ref[f]), f
        a, c = Dn.download(G.count)
        assert not a.any() and not c.any()                 # resetAccumulator(newCount), density_control.mm:493
    finally:
        Dn.close(); G.close()
    if max_g < 3000:
        assert len(ref) <= max_g and rstats[2] < 190 + (max_g > 2000) * 1000


def test_adam_moments_follow_the_gaussians_in_the_corrected_mode(gs, ctx, orc):
    """gs_adam_permute_after_densify (opt-in correction of SURVEY note D): after DensityController::apply re-orders the model,
    a kept Gaussian and the first copy of a clone keep their moments at the new index; everything new starts from zero.
    Expected layout stated here in numpy from the post-cap markers (prune 1 -> 0 outputs, keep 0 -> 1, clone 2 / split 3 -> 2)."""
    n = 3000
    g = synth.make_scene(n, 72, 0.02)
    g["scale"][:500] = np.log(0.2)
    g["opacity"][500:700] = -7.0
    accum = (synth.uniform01(6, n, 1) * 1e-3).astype(np.float32)
    count = (synth.uniform01(6, n, 2) * 3).astype(np.uint32)
    rng = np.random.default_rng(4)
    widths = (3, 3, 4, 1, 12)
    G = gs.Gaussians(ctx, n); G.upload(g)
    Dn = gs.Density(ctx, 1500000)
    A = gs.Adam(ctx, n)
    try:
        mom = [[rng.standard_normal(n * w).astype(np.float32) for w in widths] for _ in range(2)]
        for mv in range(2):
            for grp, w in enumerate(widths):
                A.upload(grp, mv, mom[mv][grp], n)
        Dn.set_scene_extent(3.3); Dn.set_seed(99)
        Dn.upload(accum, count)
        stats = Dn.apply(G, 1000, 230.0, 256.0, 6.6)
        assert min(stats) > 0, "the case must prune, clone and split"
        _, _, m = Dn.download(n, markers=True)
        outs = np.array([1, 0, 2, 2], np.int64)[m]
        off = np.concatenate([[0], np.cumsum(outs)[:-1]])
        new_n = int(outs.sum())
        assert new_n == G.count
        A.permute_after_densify(Dn, G, n)
        carry = (m == 0) | (m == 2)
        for mv in range(2):
            for grp, w in enumerate(widths):
                want = np.zeros((new_n, w), np.float32)
                want[off[carry]] = mom[mv][grp].reshape(n, w)[carry]
                got = A.download(grp, mv, new_n).reshape(new_n, w)
                assert np.array_equal(got, want), (mv, grp)
        # the model rows the moments now sit beside are the survivors themselves
        after = G.download()
        assert np.array_equal(after["opacity"][off[carry]], g["opacity"][carry])
    finally:
        A.close(); Dn.close(); G.close()


def _sorted_ok(S, keys):
    vals = np.arange(len(keys), dtype=np.uint32)
    k, v = S.sort_host(keys, vals)
    order = np.argsort(keys, kind="stable")
    return np.array_equal(k, keys[order]) and np.array_equal(v, vals[order])


def test_sort64_hybrid_plan_and_its_device_side_fallback(gs, ctx):
    """GPURadixSort64 on full 64-bit keys: spread keys finish on the prefix-passes + local-pass plan; keys the plan cannot
    take (one hot value, no entropy in the top bits, crowded bins) are caught ON THE DEVICE and sorted by the 8-pass LSD plan.
    Either way the result is numpy's stable sort, bit for bit."""
    n = (1 << 20) + 4321
    rng = np.random.default_rng(3)
    S = gs.Sort64(ctx, n)
    try:
        cases = [("random", synth.splitmix64(41, n, 1), True),
                 ("random, every key twice", np.repeat(synth.splitmix64(42, (n + 1) // 2, 1), 2)[:n], True),
                 ("already sorted", np.sort(synth.splitmix64(43, n, 1)), True),
                 ("all equal", np.full(n, 0xDEADBEEF12345678, np.uint64), False),
                 ("only the low 32 bits vary", rng.integers(0, 1 << 32, n, dtype=np.uint64), False),
                 ("top 16 bits random, rest zero", rng.integers(0, 1 << 16, n, dtype=np.uint64) << np.uint64(48), False),
                 ("30 % one hot key", np.where(rng.random(n) < 0.3, np.uint64(1) << np.uint64(63), synth.splitmix64(44, n, 1)), False)]
        for name, keys, want_hybrid in cases:
            assert _sorted_ok(S, keys), name
            assert S.last_plan_hybrid == want_hybrid, name
        gs.tune("sort_hybrid", 0)
        assert _sorted_ok(S, cases[0][1]) and not S.last_plan_hybrid
    finally:
        gs.tune("sort_hybrid", 1)
        S.close()


@pytest.mark.parametrize("logn", [22, 24])
def test_sort64_hybrid_full_size(gs, ctx, logn):
    """the bench's standalone sort sizes: sortedness, stability and permutation checked without a host sort of 16 M keys
    taking minutes: numpy's stable argsort is the checker"""
    n = 1 << logn
    S = gs.Sort64(ctx, n)
    try:
        assert _sorted_ok(S, synth.splitmix64(1234 + logn, n, 5)) and S.last_plan_hybrid
        # a sub-range sort never takes the hybrid plan: bits [8, 40) only
        keys = synth.splitmix64(7, 1 << 20, 2)
        vals = np.arange(len(keys), dtype=np.uint32)
        k, v = S.sort_host(keys, vals, 8, 40)
        order = np.argsort((keys >> np.uint64(8)) & np.uint64((1 << 32) - 1), kind="stable")
        assert np.array_equal(v, vals[order]) and not S.last_plan_hybrid
    finally:
        S.close()


@pytest.mark.parametrize("n", [440_000, 700_000, 1_000_000, 1 << 20, 1_200_000, 1_360_000])
def test_sort64_one_wave_tile_sizes_agree(gs, ctx, n):
    """one-wave inputs pick their tile size per n (rs_ipt: 8 / 9 / 10 / 12 items x 384 threads, radix_sort.cu); every choice,
    and the fixed 3072-pair tiles ("sort_even" 0), must give numpy's stable sort on keys with many ties"""
    keys = rand_u64(n, n) & np.uint64(0x000000FF00FF0FFF)
    vals = np.arange(n, dtype=np.uint32)
    order = np.argsort(keys, kind="stable")
    S = gs.Sort64(ctx, n)
    try:
        for even in (1, 0):
            gs.tune("sort_even", even)
            k, v = S.sort_host(keys, vals)
            assert np.array_equal(v, vals[order]) and np.array_equal(k, keys[order]), even
    finally:
        gs.tune("sort_even", 1)
        S.close()


def test_sort_rejects_its_own_result_buffers_as_input(gs, ctx):
    """getSortedKeys / getSortedValues are borrowed pointers into the sorter's ping-pong buffers; sorting them in
    place would read what pass 0 overwrites, so it is an argument error (copy the result out first)"""
    n = 5000
    keys = synth.splitmix64(77, n, 1)
    vals = np.arange(n, dtype=np.uint32)
    S = gs.Sort64(ctx, n)
    dk, dv = ctx.alloc(n * 8), ctx.alloc(n * 4)
    try:
        ctx.upload(dk, keys); ctx.upload(dv, vals)
        S.sort(dk, dv, n); ctx.sync()
        with pytest.raises(gs.GsError, match="aliases"):
            S.sort(S.sorted_keys, dv, n)
        with pytest.raises(gs.GsError, match="aliases"):
            S.sort(dk, S.sorted_values, n)
        S.sort(dk, dv, n); ctx.sync()                      # the sorter is still usable
        o = np.argsort(keys, kind="stable")
        assert np.array_equal(ctx.download(S.sorted_keys, n, np.uint64), keys[o])
    finally:
        ctx.free(dk); ctx.free(dv); S.close()
