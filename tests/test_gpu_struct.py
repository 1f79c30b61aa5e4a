"""Structured populations on the GPU (smc_set_populations: several demes, migration, joins) against the CPU restatement
oracle/smc_struct_oracle.c on the SAME Philox streams: prior trees by sequential insertion, the structured SMC' move
(two floating lineages above the root, migration marks, fixed-time moves), weights, resampling and ancestry, and the
per-population coalescence / migration tallies.  The restatement itself is pinned to the unmodified reference in law
(tests/test_struct_oracle.py).  Models: the committed fixtures of tests/golden/struct_moments.npz, among them
BASELINE.json configs[2] (`-I 2 4 4 -em 0 2 1 1 -ej 0.5 2 1`)."""
import os

import numpy as np
import pytest

import _libs
from test_gpu_parity import _rows

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
G = np.load(os.path.join(HERE, "golden", "struct_moments.npz"))
MODELS = sorted({k.rsplit("_prior", 1)[0] for k in G.files if k.endswith("_prior")})


@pytest.fixture(scope="module")
def S():
    import smcsmc_b200
    smcsmc_b200.lib()
    return smcsmc_b200


def _model(name, seqlen, mu, rho):
    n = int(G[name + "_n"])
    t_start, inv2N, mig, single, leaf = (G[name + k] for k in ("_t_start", "_inv2N", "_mig", "_single", "_leaf_pop"))
    return n, t_start, inv2N, mig, single, leaf, _libs.make_smodel(n, t_start, inv2N, mig, single, leaf, mu, rho, seqlen)


def _run(S, name, N, R, seqlen, mu, rho, seed, ess=0.5, max_rec=None, **kw):
    n, t_start, inv2N, mig, single, leaf, sm = _model(name, seqlen, mu, rho)
    E = len(t_start)
    lags = np.linspace(0.25 * seqlen, 0.02 * seqlen, E)
    x_end, al, is_site = _rows(n, R, seqlen, seed=len(name) + seed, **kw)
    pf = S.ParticleFilter(n, N, t_start, inv2N[:, 0], mu, rho, seqlen, lags, seed=seed, chunk=2, parity_scan=True,
                          ess_fraction=ess,
                          populations=dict(inv2N=inv2N, mig=mig, single_mig=single, leaf_pop=leaf,
                                           max_marks=120 if name == "island2_epochs" else 40))
    pf.upload_rows(x_end, al, is_site, max_rec)
    pf.init_particles()
    ref = _libs.oracle_sfilter(sm, N, seed, 2, 0.0, ess, lags, x_end, al.astype(np.int32), is_site.astype(np.int32),
                               max_rec)
    return pf, ref, (x_end, al, is_site)


@pytest.mark.parametrize("name", MODELS)
def test_prior_trees_draw_for_draw(S, name):
    """Forest::buildInitialTree with populations: trees, populations and migration marks of every particle"""
    n, t_start, inv2N, mig, single, leaf, sm = _model(name, 1e5, 2.5e-8, 1e-8)
    N = 300
    lags = np.full(len(t_start), 1000.0)
    pf = S.ParticleFilter(n, N, t_start, inv2N[:, 0], 2.5e-8, 1e-8, 1e5, lags, seed=17, chunk=1,
                          populations=dict(inv2N=inv2N, mig=mig, single_mig=single, leaf_pop=leaf,
                                           max_marks=120 if name == "island2_epochs" else 40))
    pf.upload_rows([1e5], np.full((1, n), -1, dtype=np.int8), [0])
    pf.init_particles()
    parent, height, logw, nb = pf.particles()
    pop, marks = pf.population_state()
    L = _libs.oracle()
    import ctypes
    for i in range(N):
        p, r = _libs.OrcSParticle(), _libs.OrcRng()
        L.orc_rng_init(ctypes.byref(r), ctypes.c_uint64(17), i, 0, 1, 1)
        L.orc_sinit_particle(ctypes.byref(sm), ctypes.byref(p), ctypes.byref(r), ctypes.c_double(0.0))
        nn = 2 * n - 1
        assert np.array_equal(parent[i], np.ctypeslib.as_array(p.tree.parent)[:nn]), i
        np.testing.assert_allclose(height[i], np.ctypeslib.as_array(p.tree.height)[:nn], rtol=1e-11)
        assert np.array_equal(pop[i], np.ctypeslib.as_array(p.pop)[:nn]), i
        want = sorted((p.mark_branch[q], p.mark_time[q], p.mark_pop[q]) for q in range(p.nmarks))
        got = sorted(marks[i])
        assert [(a, c) for a, _, c in got] == [(a, c) for a, _, c in want], i
        np.testing.assert_allclose([b for _, b, _ in got], [b for _, b, _ in want], rtol=1e-11)
        assert nb[i] == pytest.approx(p.tree.next_base, rel=1e-11)
    assert pf.ln_normalization_factor() == 0.0          # no flag was raised
    pf.close()


@pytest.mark.parametrize("name", MODELS)
def test_structured_filter_vs_oracle(S, name):
    """the whole sweep under a structured model, including the tallies per population and the migration tallies"""
    mu, rho = (1e-7, 4e-8) if name in ("config3", "three") else (2.5e-8, 1e-8)
    N, R, seqlen = 160, 60, 50000.0
    rng = np.random.default_rng(5)
    max_rec = None
    if name == "island2_epochs":
        max_rec = rng.integers(0, 4, size=R).astype(np.int32)
    pf, ref, _ = _run(S, name, N, R, seqlen, mu, rho, seed=31, max_rec=max_rec,
                      p_missing=0.1 if name == "split2" else 0.0, p_unphased=0.1 if name == "island2" else 0.0)
    pf.sweep(0, R // 3)
    pf.sweep(R // 3, R)
    pf.finish()
    parent, height, logw, nb = pf.particles()
    pop, marks = pf.population_state()
    c = pf.counters()
    assert c["err"] == 0 and c["rows"] == R
    assert c["resamples"] == ref["n_resample"] and ref["n_resample"] > 0
    assert c["recombinations"] == ref["n_recomb"] and ref["n_recomb"] > 0
    assert np.array_equal(parent, ref["parent"])
    assert np.array_equal(pop, ref["pop"])
    np.testing.assert_allclose(height, ref["height"], rtol=1e-11)
    np.testing.assert_allclose(nb, ref["next_base"], rtol=1e-11)
    np.testing.assert_allclose(logw, ref["logw"], rtol=0, atol=1e-9)
    assert pf.ln_normalization_factor() == pytest.approx(ref["lnZ"], rel=1e-11)
    assert np.array_equal(pf.last_ancestry(), ref["last_anc"])
    for i in range(N):
        got, want = sorted(marks[i]), ref["marks"][i]
        assert [(a, c_) for a, _, c_ in got] == [(a, c_) for a, _, c_ in want], i
        np.testing.assert_allclose([b for _, b, _ in got], [b for _, b, _ in want], rtol=1e-11)
    t = pf.tallies()
    P = G[name + "_inv2N"].shape[1]
    assert t.shape == ref["tallies"].shape == (2, 2 + 3 * P + P * P, len(G[name + "_t_start"]))
    np.testing.assert_allclose(t[0], ref["tallies"][0], rtol=1e-10, atol=1e-13)
    # w^2 sums: copies do not share records, so the device value is a lower bound
    assert np.all(t[1] <= ref["tallies"][1] * (1 + 1e-9) + 1e-300)
    # ... and exactly the engine's definition (pooled at the row in which the record's update window was passed)
    np.testing.assert_allclose(t[1], ref["w2_pooled_at_close"], rtol=1e-9, atol=1e-300)
    # the tallies carry what the model implies: coalescences in every population that has samples or receives lineages,
    # migration counts only where the model has a rate
    mig = G[name + "_mig"]
    for p in range(P):
        assert t[0][3 + 3 * p].sum() >= 0 and t[0][4 + 3 * p].sum() > 0 or mig[:, :, p].sum() == 0
        for q in range(P):
            cnt = t[0][2 + 3 * P + p * P + q]
            assert np.all(cnt[mig[:, p, q] == 0] == 0)
    if mig.sum() > 0:
        assert t[0][2 + 3 * P:].sum() > 0
    pf.close()


@pytest.mark.parametrize("name", ["config3", "island2_epochs"])
def test_structured_vb_factors(S, name):
    """-vb with populations (smc_set_population_vb): a factor per coalescence (epoch, population) and per migration
    (epoch, from, to) enters the particle weights exactly as in the restatement -- weights, resampling, ancestry, ln
    normaliser -- and the run differs from the one without factors"""
    mu, rho = (1e-7, 4e-8) if name == "config3" else (2.5e-8, 1e-8)
    N, R, seqlen = 128, 50, 40000.0
    n, t_start, inv2N, mig, single, leaf, _ = _model(name, seqlen, mu, rho)
    E, P = inv2N.shape
    rng = np.random.default_rng(9)
    vb_c = np.log(rng.uniform(0.7, 1.0, size=(E, P)))
    vb_m = np.log(rng.uniform(0.5, 1.0, size=(E, P, P)))
    sm = _libs.make_smodel(n, t_start, inv2N, mig, single, leaf, mu, rho, seqlen, vb_coal=vb_c, vb_mig=vb_m)
    lags = np.linspace(0.25 * seqlen, 0.02 * seqlen, E)
    x_end, al, is_site = _rows(n, R, seqlen, seed=77)
    out = []
    for with_vb in (True, False):
        pf = S.ParticleFilter(n, N, t_start, inv2N[:, 0], mu, rho, seqlen, lags, seed=13, chunk=2, parity_scan=True,
                              populations=dict(inv2N=inv2N, mig=mig, single_mig=single, leaf_pop=leaf,
                                               max_marks=120 if name == "island2_epochs" else 40))
        if with_vb:
            pf.set_population_vb(vb_c, vb_m)
        pf.upload_rows(x_end, al, is_site)
        pf.init_particles()
        pf.sweep()
        pf.finish()
        out.append((pf.particles(), pf.ln_normalization_factor(), pf.last_ancestry(), pf.tallies(), pf.counters()))
        pf.close()
    ref = _libs.oracle_sfilter(sm, N, 13, 2, 0.0, 0.5, lags, x_end, al.astype(np.int32), is_site.astype(np.int32))
    (parent, height, logw, nb), lnZ, anc, t, c = out[0]
    assert c["resamples"] == ref["n_resample"] and ref["n_resample"] > 0
    assert np.array_equal(parent, ref["parent"]) and np.array_equal(anc, ref["last_anc"])
    np.testing.assert_allclose(height, ref["height"], rtol=1e-11)
    np.testing.assert_allclose(logw, ref["logw"], rtol=0, atol=1e-9)
    assert lnZ == pytest.approx(ref["lnZ"], rel=1e-11)
    np.testing.assert_allclose(t[0], ref["tallies"][0], rtol=1e-10, atol=1e-13)
    assert abs(out[1][1] - lnZ) > 1e-3            # the factors do something


def test_no_coalescence_is_reported(S):
    """two isolated populations that never join: the reference throws 'Lines did not coalescence'; so does the engine"""
    n = 4
    t_start = [0.0]
    inv2N = np.array([[1 / 2e4, 1 / 2e4]])
    pf = S.ParticleFilter(n, 64, t_start, inv2N[:, 0], 2.5e-8, 1e-8, 1e4, [1000.0], seed=1,
                          populations=dict(inv2N=inv2N, leaf_pop=[0, 0, 1, 1]))
    pf.upload_rows([1e4], np.full((1, n), -1, dtype=np.int8), [0])
    pf.init_particles()
    with pytest.raises(S.SmcError, match="did not coalesce"):
        pf.ln_normalization_factor()
    pf.close()


def test_population_argument_checks(S):
    n = 4
    mk = lambda: S.ParticleFilter(n, 32, [0.0], [1 / 2e4], 2.5e-8, 1e-8, 1e4, [1000.0])   # noqa: E731
    inv = np.array([[1 / 2e4, 1 / 2e4]])
    pf = mk()
    with pytest.raises(S.SmcError):
        pf.set_populations(inv2N=np.array([[1 / 2e4]]), leaf_pop=[0] * n)                   # one population
    with pytest.raises(S.SmcError):
        pf.set_populations(inv2N=inv, leaf_pop=[0, 0, 2, 1])                                # population out of range
    with pytest.raises(S.SmcError):
        pf.set_populations(inv2N=-inv, leaf_pop=[0, 0, 1, 1])
    with pytest.raises(S.SmcError):
        pf.set_populations(inv2N=inv, leaf_pop=[0, 0, 1, 1], mig=np.array([[[0, -1.0], [0, 0]]]))
    with pytest.raises(S.SmcError):
        pf.set_populations(inv2N=inv, leaf_pop=[0, 0, 1, 1], max_marks=500)
    pf.upload_rows([1e4], np.full((1, n), -1, dtype=np.int8), [0])
    with pytest.raises(S.SmcError):
        pf.set_populations(inv2N=inv, leaf_pop=[0, 0, 1, 1])                                # after the rows
    pf.close()
