"""Structured populations (several demes, migration, joins) -- the CPU restatement oracle/smc_struct_oracle.c against
the UNMODIFIED reference, distributionally: prior trees (Forest::buildInitialTree) and prior ARG walks
(sampleNextGenealogy + sampleRecSeqPosition) under five structured models, reference side committed as
tests/golden/struct_moments.npz (make_golden.py::golden_struct_moments, own mt19937_64 stream).  This pins the
algorithm the device code for SURVEY.md section 8 rows a2/a4 with P > 1 will be compared with draw for draw."""
import ctypes
import os

import numpy as np
import pytest

import _libs

HERE = os.path.dirname(os.path.abspath(__file__))
G = np.load(os.path.join(HERE, "golden", "struct_moments.npz"))
MODELS = sorted({k.rsplit("_prior", 1)[0] for k in G.files if k.endswith("_prior")})


def P(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def model(name):
    n = int(G[name + "_n"])
    return n, _libs.make_smodel(n, G[name + "_t_start"], G[name + "_inv2N"], G[name + "_mig"], G[name + "_single"],
                                G[name + "_leaf_pop"], float(G[name + "_mu"]), float(G[name + "_rho"]), 1e8)


def batch_means(x, nb=50):
    m = np.array([c.mean() for c in np.array_split(x, nb)])
    return x.mean(), m.std(ddof=1) / np.sqrt(nb)


@pytest.mark.parametrize("name", MODELS)
def test_prior_trees_match_reference(name):
    L = _libs.oracle()
    n, m = model(name)
    count = 40000
    tm, tl, mk = np.zeros(count), np.zeros(count), np.zeros(count)
    assert L.orc_sprior_stats(ctypes.byref(m), ctypes.c_uint64(11), count, P(tm), P(tl), P(mk)) == 0
    ref = G[name + "_prior"]
    for k, x in enumerate((tm, tl, tm ** 2, tl ** 2)):
        z = (x.mean() - ref[k, 0]) / np.hypot(ref[k, 1], x.std(ddof=1) / np.sqrt(count))
        assert abs(z) < 4.5, (name, k, x.mean(), ref[k].tolist(), z)
    if G[name + "_mig"].sum() > 0 or G[name + "_single"].sum() > 0:
        assert mk.mean() > 0.05                      # migration marks do occur


@pytest.mark.parametrize("name", MODELS)
def test_arg_walk_matches_reference(name):
    L = _libs.oracle()
    n, m = model(name)
    count = 20000
    stats = []
    for seed in (3, 4):                              # two walks: more batches, smaller error
        pos, rh, tm, tl, mk = (np.zeros(count) for _ in range(5))
        got = L.orc_sarg_walk(ctypes.byref(m), ctypes.c_uint64(seed), count, P(pos), P(rh), P(tm), P(tl), P(mk))
        assert got > 1000, got                       # negative: orc_scheck found an inconsistent tree
        stats.append((np.diff(pos[:got]), rh[:got], tm[1:got], tl[1:got]))
    ref = G[name + "_walk"]
    for k in range(4):
        x = np.concatenate([s[k] for s in stats])
        mean, se = batch_means(x, 100)
        z = (mean - ref[k, 0]) / np.hypot(ref[k, 1], se)
        assert abs(z) < 4.5, (name, k, mean, ref[k].tolist(), z)


def test_single_population_limit_is_the_panmictic_oracle_in_law():
    """P = 1: the structured restatement and the panmictic one (smc_oracle.c, pinned separately) give the same moments"""
    L = _libs.oracle()
    n = 6
    t_start = [0.0, 4000.0, 30000.0]
    inv = [[1 / 2e4], [1 / 6e3], [1 / 4e4]]
    m = _libs.make_smodel(n, t_start, inv, np.zeros((3, 1, 1)), np.zeros((3, 1, 1)), [0] * n, 2.5e-8, 1e-8, 1e8)
    count = 40000
    tm, tl = np.zeros(count), np.zeros(count)
    assert L.orc_sprior_stats(ctypes.byref(m), ctypes.c_uint64(5), count, P(tm), P(tl), None) == 0
    pm = _libs.make_model(n, t_start, [x[0] for x in inv], 2.5e-8, 1e-8, 1e8)
    p, r = _libs.OrcParticle(), _libs.OrcRng()
    tm1, tl1 = np.zeros(count), np.zeros(count)
    for t in range(count):
        L.orc_rng_init(ctypes.byref(r), 9, t, 0, 0, 7)
        L.orc_init_particle(ctypes.byref(pm), ctypes.byref(p), ctypes.byref(r), ctypes.c_double(0.0))
        h = np.ctypeslib.as_array(p.height)[:2 * n - 1]
        par = np.ctypeslib.as_array(p.parent)[:2 * n - 1]
        tm1[t] = h.max()
        tl1[t] = (h[par[par >= 0]] - h[par >= 0]).sum()
    for a, b in ((tm, tm1), (tl, tl1)):
        z = (a.mean() - b.mean()) / np.hypot(a.std(ddof=1), b.std(ddof=1)) * np.sqrt(count)
        assert abs(z) < 4.5, (a.mean(), b.mean(), z)
