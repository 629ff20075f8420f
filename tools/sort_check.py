"""Correctness sweep of GPURadixSort64 (gs_sort64_sort) against numpy's stable argsort on inputs that exercise the hybrid plan
(prefix passes + local pass) and every way it has to give up on the device and fall back to the 8-pass LSD plan.
usage: python tools/sort_check.py"""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs, synth

ctx = gs.Context(0)


def check(name, keys):
    n = len(keys)
    vals = np.arange(n, dtype=np.uint32)
    S = gs.Sort64(ctx, n)
    k, v = S.sort_host(keys, vals)
    S.close()
    order = np.argsort(keys, kind="stable")
    ok = np.array_equal(k, keys[order]) and np.array_equal(v, vals[order])
    print(f"{'ok  ' if ok else 'FAIL'} {name} n={n}")
    return ok


rng = np.random.default_rng(1)
allok = True
for n in (1 << 20, (1 << 20) + 12345, 3000000, 1 << 22, (1 << 24) + 7, 1 << 24):
    allok &= check("random u64", synth.splitmix64(99, n, 3))
n = (1 << 21) + 999
allok &= check("all equal", np.full(n, 0x123456789ABCDEF0, np.uint64))
allok &= check("low 48 bits zero", (rng.integers(0, 1 << 16, n, dtype=np.uint64) << np.uint64(48)))
allok &= check("only low 32 bits random", rng.integers(0, 1 << 32, n, dtype=np.uint64))
allok &= check("256 distinct keys", synth.splitmix64(5, 256, 1)[rng.integers(0, 256, n)])
allok &= check("already sorted", np.sort(synth.splitmix64(7, n, 1)))
allok &= check("reverse sorted", np.sort(synth.splitmix64(8, n, 1))[::-1].copy())
allok &= check("top 24 bits random, rest zero", (rng.integers(0, 1 << 24, n, dtype=np.uint64) << np.uint64(40)))
allok &= check("random with 30% duplicates of one key", np.where(rng.random(n) < 0.3, np.uint64(0x8000000000000000), synth.splitmix64(9, n, 1)))
allok &= check("half the keys in one top byte", np.where(rng.random(n) < 0.5, synth.splitmix64(10, n, 1) >> np.uint64(8), synth.splitmix64(11, n, 1)))
allok &= check("random, every key twice", np.repeat(synth.splitmix64(12, n // 2, 1), 2))
print("ALL OK" if allok else "FAILURES")
sys.exit(0 if allok else 1)
