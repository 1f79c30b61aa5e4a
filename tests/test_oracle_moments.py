"""CPU tests: the oracle's STOCHASTIC pieces (prior tree, SMC' move, next recombination position; Philox stream) against
sample moments of the unmodified reference's own code and mt19937_64 stream (tests/golden/moments.npz).  RNG streams
differ by design, so agreement is distributional: z-tests at 5 standard errors (both sides sampled)."""
import ctypes
import os

import numpy as np
import pytest

import _libs
from _libs import OrcParticle, OrcRng

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = ["const4", "bottleneck8", "zigzag8"]
ZMAX = 5.0


def _z(mean, se, ref):
    return abs(mean - ref[0]) / np.sqrt(se * se + ref[1] * ref[1] + 1e-300)


def _model(g, name, seqlen=1e8):
    n = int(g[f"{name}_n"])
    return n, _libs.make_model(n, list(g[f"{name}_t_start"]), list(g[f"{name}_inv2N"]), float(g[f"{name}_mu"]),
                               float(g[f"{name}_rho"]), seqlen)


@pytest.mark.parametrize("name", NAMES)
def test_prior_tree_moments(name):
    L = _libs.oracle()
    g = np.load(os.path.join(G, "moments.npz"))
    n, m = _model(g, name)
    nn = 2 * n - 1
    count = 40000
    tm, tl, first = np.zeros(count), np.zeros(count), np.zeros(count)
    p, r = OrcParticle(), OrcRng()
    for t in range(count):
        L.orc_rng_init(ctypes.byref(r), 12345, t, 0, 0, 1)
        L.orc_init_particle(ctypes.byref(m), ctypes.byref(p), ctypes.byref(r), 1.0)
        h = np.ctypeslib.as_array(p.height)[:nn]
        tm[t], first[t] = h[nn - 1], h[n]
        tl[t] = L.orc_tree_length(n, p.parent, p.height)
    for x, ref in zip((tm, tl, first, tm ** 2), g[f"{name}_prior"]):
        assert _z(x.mean(), x.std(ddof=1) / np.sqrt(count), ref) < ZMAX


@pytest.mark.parametrize("name", NAMES)
def test_smc_move_moments(name):
    L = _libs.oracle()
    g = np.load(os.path.join(G, "moments.npz"))
    n, m = _model(g, name)
    nn = 2 * n - 1
    p0, h0 = g[f"{name}_move_start_parent"], g[f"{name}_move_start_height"]
    reps = 40000
    rh, mtl, mtm, unch, newroot = (np.zeros(reps) for _ in range(5))
    p, r = OrcParticle(), OrcRng()
    rec_h = ctypes.c_double()
    for i in range(reps):
        for v in range(nn):
            p.parent[v] = int(p0[v]); p.height[v] = float(h0[v])
        p.cur_base = 100.0
        L.orc_rng_init(ctypes.byref(r), 777, i, 0, 0, 2)
        L.orc_smc_move(ctypes.byref(m), ctypes.byref(p), ctypes.byref(r), 1 << 30, None, ctypes.byref(rec_h))
        h = np.ctypeslib.as_array(p.height)[:nn]
        rh[i] = rec_h.value
        mtl[i] = L.orc_tree_length(n, p.parent, p.height)
        mtm[i] = h.max()
        unch[i] = float(np.abs(h - h0).max() < 1e-9 * h0.max())
        newroot[i] = float(mtm[i] > h0.max() * (1 + 1e-12))
        # structural invariants: internal slots sorted by height, exactly one root, every internal node has 2 children
        assert np.all(np.diff(h[n:]) >= 0)
        par = np.ctypeslib.as_array(p.parent)[:nn]
        assert (par < 0).sum() == 1 and par[nn - 1] == -1
        assert np.array_equal(np.bincount(par[par >= 0], minlength=nn)[n:], np.full(n - 1, 2))
    for x, ref in zip((rh, mtl, mtm, unch, newroot), g[f"{name}_move"]):
        assert _z(x.mean(), x.std(ddof=1) / np.sqrt(reps), ref) < ZMAX


@pytest.mark.parametrize("name", NAMES)
def test_arg_walk_moments(name):
    """Walk the prior ARG with the oracle's row extension (no data: all alleles missing => weight untouched)."""
    L = _libs.oracle()
    g = np.load(os.path.join(G, "moments.npz"))
    n, m = _model(g, name)
    nn = 2 * n - 1
    p, r = OrcParticle(), OrcRng()
    L.orc_rng_init(ctypes.byref(r), 99, 0, 0, 0, 3)
    L.orc_init_particle(ctypes.byref(m), ctypes.byref(p), ctypes.byref(r), 0.0)
    cnt = 20000
    pos, tmr, tln = [], [], []
    rec_h = ctypes.c_double()
    rhs = []
    for i in range(cnt):
        if p.next_base >= m.seqlen:
            break
        p.cur_base = p.next_base
        L.orc_smc_move(ctypes.byref(m), ctypes.byref(p), ctypes.byref(r), 1 << 30, None, ctypes.byref(rec_h))
        L.orc_redraw_next_base(ctypes.byref(m), ctypes.byref(p), ctypes.byref(r))
        pos.append(p.cur_base); rhs.append(rec_h.value)
        tmr.append(p.height[nn - 1]); tln.append(L.orc_tree_length(n, p.parent, p.height))
    pos, rhs, tmr, tln = map(np.array, (pos, rhs, tmr, tln))
    assert len(pos) > 1000

    def bm(x):
        b = np.array_split(x, 50)
        mm = np.array([c.mean() for c in b])
        return x.mean(), mm.std(ddof=1) / np.sqrt(len(mm))
    for x, ref in zip((np.diff(pos), rhs, tmr[1:], tln[1:]), g[f"{name}_walk"]):
        mean, se = bm(x)
        assert _z(mean, se, ref) < ZMAX
