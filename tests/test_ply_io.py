"""PLY export / load of the host layer (gaussiansplatting_b200/host/ply_io.hpp restating PLYExporter::exportPLY,
ply_exporter.hpp:17-162, and load_ply / detectLinearScales, ply_loader.cpp:17-290).  Host-only: runs without a GPU.  The
expected values are stated here in numpy from the reference's rules, independently of the C++."""
import numpy as np
import pytest
from gaussiansplatting_b200 import host, synth
from gaussiansplatting_b200.layouts import GAUSSIAN

NAMES = (["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(9)] +
         ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"])


def _read_ply_raw(path):
    raw = open(path, "rb").read()
    end = raw.index(b"end_header\n") + len(b"end_header\n")
    header = raw[:end].decode().splitlines()
    n = int([l for l in header if l.startswith("element vertex")][0].split()[2])
    props = [l.split()[2] for l in header if l.startswith("property")]
    body = np.frombuffer(raw, "<f4", n * len(props), end).reshape(n, len(props))
    assert end + body.nbytes == len(raw)
    return header, props, body


def _scene(n=500, seed=3):
    g = synth.make_scene(n, seed, 0.05)
    rng = np.random.default_rng(seed)
    g["sh"] = rng.standard_normal((n, 12)).astype(np.float32)          # exercise the f_rest interleave
    return g


def test_export_layout_and_values(tmp_path):
    g = _scene()
    g["position"][5, 0] = np.nan; g["position"][6, 0] = np.inf; g["position"][7, 0] = 2e6      # skipped: x only (ply_exporter.hpp:74-77)
    g["position"][8, 1] = np.nan                                                            # kept: only x is tested
    p = tmp_path / "out.ply"
    host.ply_export(p, g)
    header, props, body = _read_ply_raw(p)
    assert header[:2] == ["ply", "format binary_little_endian 1.0"] and props == NAMES
    keep = np.ones(len(g), bool); keep[[5, 6, 7]] = False
    k = g[keep]
    assert body.shape == (len(k), 26)
    want = np.concatenate([k["position"], np.zeros((len(k), 3), np.float32), k["sh"][:, [0, 4, 8]],
                           k["sh"][:, [1, 5, 9, 2, 6, 10, 3, 7, 11]], k["opacity"][:, None], k["scale"], k["rotation"]], axis=1)
    assert np.array_equal(body.view(np.uint32), want.astype(np.float32).view(np.uint32))


def _expected_load(pos, scale, rot, op, dc, rest):
    """load_ply's rules (ply_loader.cpp:139-235) in numpy float32"""
    f32 = np.float32
    samp = scale[:1000]
    linear = bool((samp < 0).sum() == 0 and samp.max() <= 1.0 and samp.min() > 0.0)
    ok = ~(np.isnan(pos).any(1) | np.isinf(pos).any(1) | (np.abs(pos) > 1e6).any(1))
    s = scale.astype(f32)
    if linear:
        s = np.log(np.maximum(s, f32(1e-8))).astype(f32)
    s = np.clip(s, -8, 8).astype(f32)
    ql = np.sqrt((rot[:, 0] * rot[:, 0] + rot[:, 1] * rot[:, 1] + rot[:, 2] * rot[:, 2] + rot[:, 3] * rot[:, 3]).astype(f32)).astype(f32)
    q = np.where((ql > 1e-4)[:, None], rot / np.where(ql > 0, ql, 1)[:, None], np.array([1, 0, 0, 0], f32)).astype(f32)
    out = np.zeros(int(ok.sum()), GAUSSIAN)
    out["position"] = pos[ok]; out["scale"] = s[ok]; out["rotation"] = q[ok]; out["opacity"] = op[ok]
    sh = np.zeros((len(pos), 12), f32)
    sh[:, [0, 4, 8]] = dc
    if rest is not None:
        sh[:, [1, 5, 9, 2, 6, 10, 3, 7, 11]] = rest
    out["sh"] = sh[ok]
    return out, linear, int((~ok).sum())


def _same(a, b, log_tol=0.0):
    for f in ("position", "rotation", "opacity", "sh"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f
    if log_tol:
        assert np.abs(a["scale"] - b["scale"]).max() <= log_tol
    else:
        assert np.array_equal(a["scale"].view(np.uint32), b["scale"].view(np.uint32))
    assert not a["_pad0"].any() and not a["_pad1"].any() and not a["_pad2"].any()


def test_export_then_load_round_trip(tmp_path):
    g = _scene(2000, 9)
    g["rotation"][:100] *= 2.5                       # re-normalised on load
    g["rotation"][100:110] = 0                       # -> identity
    g["scale"][110:130] = 9.0; g["scale"][130:150] = -11.0      # clamped to +-8
    g["position"][150:160] = np.nan                  # dropped by the exporter (x) and the loader
    g["position"][160:165, 2] = np.inf               # written by the exporter (x is fine), skipped by the loader
    p = tmp_path / "rt.ply"
    host.ply_export(p, g)
    got, info = host.ply_load(p)
    kept = g[~np.isnan(g["position"][:, 0])]
    want, linear, skipped = _expected_load(kept["position"], kept["scale"], kept["rotation"], kept["opacity"],
                                           kept["sh"][:, [0, 4, 8]], kept["sh"][:, [1, 5, 9, 2, 6, 10, 3, 7, 11]])
    assert info == dict(vertices=len(kept), skipped=5, linear_scales=False, has_sh_rest=True) and skipped == 5 and not linear
    _same(got, want)
    # a second trip is the identity: the loader's output is already normalised and clamped
    p2 = tmp_path / "rt2.ply"
    host.ply_export(p2, got)
    again, _ = host.ply_load(p2)
    assert np.array_equal(again["position"], got["position"]) and np.array_equal(again["scale"], got["scale"])
    assert np.abs(again["rotation"] - got["rotation"]).max() <= 1e-6          # re-normalising a unit quaternion moves it by <= 1 ulp


def _write_ply(path, cols, order, fmt="binary_little_endian", types=None, pre_element=False, post_element=False):
    n = len(next(iter(cols.values())))
    types = types or {}
    lines = ["ply", f"format {fmt} 1.0", "comment written by tests/test_ply_io.py"]
    if pre_element:
        lines += ["element camera 2", "property float cx", "property list uchar int idx"]
    lines.append(f"element vertex {n}")
    lines += [f"property {types.get(k, 'float')} {k}" for k in order]
    if post_element:
        lines += ["element face 1", "property list uchar int vertex_indices"]
    lines.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(lines) + "\n").encode())
        if fmt == "ascii":
            if pre_element:
                f.write(b"0.5 2 7 8\n1.5 0\n")
            for i in range(n):
                f.write((" ".join(repr(float(cols[k][i])) for k in order) + "\n").encode())
            if post_element:
                f.write(b"3 0 1 2\n")
            return
        e = "<" if fmt == "binary_little_endian" else ">"
        if pre_element:
            f.write(np.array([0.5], e + "f4").tobytes() + bytes([2]) + np.array([7, 8], e + "i4").tobytes())
            f.write(np.array([1.5], e + "f4").tobytes() + bytes([0]))
        dt = np.dtype([(k, e + {"float": "f4", "double": "f8", "uchar": "u1", "int": "i4"}[types.get(k, "float")]) for k in order])
        rows = np.zeros(n, dt)
        for k in order:
            rows[k] = cols[k]
        f.write(rows.tobytes())
        if post_element:
            f.write(bytes([3]) + np.array([0, 1, 2], e + "i4").tobytes())


def _cols(n, seed, linear=False, rest=True):
    rng = np.random.default_rng(seed)
    c = {k: rng.standard_normal(n).astype(np.float32) for k in ["x", "y", "z", "opacity", "f_dc_0", "f_dc_1", "f_dc_2", "rot_0", "rot_1", "rot_2", "rot_3"]}
    for k in ("scale_0", "scale_1", "scale_2"):
        c[k] = (rng.uniform(0.001, 1.0, n) if linear else rng.uniform(-6, 1, n)).astype(np.float32)
    if rest:
        for i in range(9):
            c[f"f_rest_{i}"] = rng.standard_normal(n).astype(np.float32)
    c["nx"] = np.zeros(n, np.float32); c["red"] = rng.integers(0, 255, n).astype(np.float32)
    return c


def _want(c, rest=True):
    pos = np.stack([c["x"], c["y"], c["z"]], 1); sc = np.stack([c["scale_0"], c["scale_1"], c["scale_2"]], 1)
    rot = np.stack([c[f"rot_{i}"] for i in range(4)], 1); dc = np.stack([c[f"f_dc_{i}"] for i in range(3)], 1)
    rs = np.stack([c[f"f_rest_{i}"] for i in range(9)], 1) if rest else None
    return _expected_load(pos, sc, rot, c["opacity"], dc, rs)


@pytest.mark.parametrize("fmt", ["binary_little_endian", "binary_big_endian", "ascii"])
def test_loader_finds_properties_by_name_in_any_order(tmp_path, fmt):
    c = _cols(300, 5)
    order = list(c.keys())
    np.random.default_rng(1).shuffle(order)
    p = tmp_path / "shuffled.ply"
    _write_ply(p, c, order, fmt, types={"red": "uchar"}, pre_element=True, post_element=True)
    got, info = host.ply_load(p)
    want, linear, skipped = _want(c)
    assert info == dict(vertices=300, skipped=0, linear_scales=False, has_sh_rest=True)
    _same(got, want)


def test_loader_converts_linear_scales_and_tolerates_missing_sh_rest(tmp_path):
    c = _cols(1500, 6, linear=True, rest=False)
    c["x"][10] = np.nan; c["z"][11] = 1e7
    c["scale_0"][1200] = 0.0                                    # beyond the 1000-vertex sample: log(max(0, 1e-8)) clamps to -8
    p = tmp_path / "linear.ply"
    _write_ply(p, c, list(c.keys()))
    got, info = host.ply_load(p)
    want, linear, skipped = _want(c, rest=False)
    assert linear and info == dict(vertices=1500, skipped=2, linear_scales=True, has_sh_rest=False)
    _same(got, want, log_tol=2e-6)                              # libm logf vs numpy's: last-ulp differences allowed
    assert not got["sh"][:, [1, 2, 3, 5, 6, 7, 9, 10, 11]].any()
    assert got["scale"].min() == -8.0


def test_loader_converts_a_double_property_instead_of_misreading_it(tmp_path):
    c = _cols(50, 8)
    p = tmp_path / "double.ply"
    _write_ply(p, c, list(c.keys()), types={"opacity": "double", "x": "double"})
    got, _ = host.ply_load(p)
    _same(got, _want(c)[0])


def test_loader_failures_are_loud_and_empty(tmp_path):
    c = _cols(20, 2)
    del c["rot_3"]
    p = tmp_path / "norot.ply"
    _write_ply(p, c, list(c.keys()))
    with pytest.raises(host.GsError, match="rotations"):
        host.ply_load(p)
    with pytest.raises(host.GsError):
        host.ply_load(tmp_path / "missing.ply")
    trunc = tmp_path / "trunc.ply"
    c = _cols(20, 2)
    _write_ply(trunc, c, list(c.keys()))
    data = open(trunc, "rb").read()
    open(trunc, "wb").write(data[:-40])
    with pytest.raises(host.GsError, match="truncated"):
        host.ply_load(trunc)
