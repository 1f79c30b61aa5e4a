"""GPU parity tests (run on the B200 box: pytest -m gpu).  Every call goes through the C ABI of libsmcgpu.so with host
buffers; the checker is the CPU oracle (oracle/liboracle.so) and the committed golden vectors that came from the
unmodified reference (tests/golden/).  Tolerances: integer/index work bit-exact; likelihoods bit-exact (1e-12 is the
stated bar); floating-point reductions whose order differs (ln normaliser, tallies) 1e-10 relative."""
import ctypes
import os

import numpy as np
import pytest

import _libs
import _cases
from _libs import P

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def S():
    import smcsmc_b200
    return smcsmc_b200


@pytest.mark.parametrize("n", [4, 8, 16])
def test_dephase_vs_reference_golden(S, n):
    """-dephase: the driver recodes genotypes whose haplotypes differ as unphased pairs; the device likelihood of the
    recoded data equals the reference's likelihood with dephase = true on the phased data (tests/golden/dephase.npz)"""
    from test_oracle_golden import _dephased
    g, d = np.load(os.path.join(G, "prune.npz")), np.load(os.path.join(G, "dephase.npz"))
    mu = float(g[f"n{n}_mu"])
    data = np.array([_dephased(row) for row in g[f"n{n}_phased"]])
    assert (data == 2).any()
    for aa, key in ((0, "site_dephase"), (1, "site_dephase_aa")):
        got = S.site_likelihood(g[f"n{n}_parent"], g[f"n{n}_height"], data, mu, ancestral_aware=aa)
        np.testing.assert_allclose(got, d[f"n{n}_{key}"], rtol=1e-12, atol=0)


def test_fastexp_bit_exact(S):
    L = _libs.oracle()
    x = -np.concatenate([np.logspace(-12, 1.5, 4001), [0.0, 0.7184, 0.71845, 0.7185]])
    got = S.fastexp(x)
    want = np.array([L.orc_fastexp(float(v)) for v in x])
    rational = x * x < 0.516167859
    assert np.array_equal(got[rational], want[rational])          # same rational expression, no contraction
    np.testing.assert_allclose(got[~rational], want[~rational], rtol=2.3e-16)   # libm exp vs CUDA exp: <= 1 ulp


@pytest.mark.parametrize("n", [2, 3, 4, 8, 16])
def test_site_likelihood_and_branch_lengths_vs_reference_golden(S, n):
    g = np.load(os.path.join(G, "prune.npz"))
    mu = float(g[f"n{n}_mu"])
    parent, height = g[f"n{n}_parent"], g[f"n{n}_height"]
    for aa, key_ph, key_site in ((0, "lik", "site"), (1, "lik_aa", "site_aa")):
        got = S.site_likelihood(parent, height, g[f"n{n}_phased"], mu, ancestral_aware=aa)
        want = g[f"n{n}_{key_ph}"]
        # the trees stretched beyond the fastexp switch-over use exp(): CUDA's exp is within 1 ulp of libm's
        np.testing.assert_allclose(got, want, rtol=1e-12, atol=0)
        assert (got == want).mean() > 0.45
        got = S.site_likelihood(parent, height, g[f"n{n}_data"], mu, ancestral_aware=aa)
        np.testing.assert_allclose(got, g[f"n{n}_{key_site}"], rtol=1e-12, atol=0)
    bl, tl = S.branch_lengths(parent, height, g[f"n{n}_phased"])
    np.testing.assert_allclose(bl, g[f"n{n}_bl"], rtol=1e-12, atol=0)
    np.testing.assert_allclose(tl, g[f"n{n}_tl"], rtol=1e-12, atol=0)
    # and bit-exact against the oracle's summation order
    L = _libs.oracle()
    for t in range(parent.shape[0]):
        par, hei = np.ascontiguousarray(parent[t]), np.ascontiguousarray(height[t])
        ph = np.ascontiguousarray(g[f"n{n}_phased"][t])
        assert bl[t] == L.orc_branch_length_with_data(n, P(par), P(hei), P(ph))
        assert tl[t] == L.orc_tree_length(n, P(par), P(hei))


def test_reference_example_tree(S):
    g = np.load(os.path.join(G, "prune.npz"))
    pats = g["ex_pats"].astype(np.int8)
    k = len(pats)
    par = np.tile(g["ex_parent"], (k, 1))
    hei = np.tile(g["ex_height"], (k, 1))
    bl, tl = S.branch_lengths(par, hei, pats)
    assert list(tl) == [24.0] * k
    assert np.array_equal(bl, g["ex_bl"])
    lik = S.site_likelihood(par, hei * 1e4, pats, float(g["ex_mu"]))
    assert np.array_equal(lik, g["ex_lik_scaled1e4"])


def test_systematic_offspring_bit_exact_vs_reference_golden(S):
    g = np.load(os.path.join(G, "resampling.npz"))
    for k in range(int(g["ncases"])):
        got = S.systematic_offspring(g[f"c{k}_ps"], float(g[f"c{k}_u"]))
        assert np.array_equal(got, g[f"c{k}_counts"]), k


def test_resample_counts_from_weights(S):
    """bit-exact offspring given the same WEIGHTS and uniform: order-faithful partial sums on the device"""
    g = np.load(os.path.join(G, "resampling.npz"))
    L = _libs.oracle()
    for j in range(int(g["npaths"])):
        w = g[f"p{j}_w"].copy()
        N = len(w)
        L.orc_normalize(P(w), N)
        u = float(g[f"p{j}_u"])
        ess, counts = S.resample_counts(w, u, parity_scan=True)
        ps = np.zeros(N + 1)
        ess_o = L.orc_ess_partial_sums(P(w), N, P(ps))
        assert ess == ess_o
        if int(g[f"p{j}_resampled"]):
            assert np.array_equal(counts, g[f"p{j}_mult"])
        # the blocked scan may move an offspring only when a boundary lies within rounding of some u_j
        ess2, counts2 = S.resample_counts(w, u, parity_scan=False)
        assert counts2.sum() == N and np.abs(counts2 - counts).sum() <= 2
        assert ess2 == pytest.approx(ess_o, rel=1e-12)


@pytest.mark.parametrize("N", [1, 2, 3, 1000, 65537, 1 << 20])
def test_offspring_chain_sizes(S, N):
    """the closed-form u_j chain against the explicit chain, edge sizes included"""
    L = _libs.oracle()
    rng = np.random.default_rng(N)
    w = np.exp(2 * rng.standard_normal(N))
    w /= w.sum()
    ps = np.concatenate([[0.0], np.cumsum(w)])
    ps[-1] = max(ps[-1], ps[-2])
    for u in (1e-300, 0.37, 0.99999999):   # u within rounding of 1 runs the reference loop off its array
        want = np.zeros(N, dtype=np.int32)
        L.orc_systematic_resampling(P(ps), N, N, u, P(want))
        got = S.systematic_offspring(ps, u)
        assert np.array_equal(got, want)


def test_normalize_log(S):
    L = _libs.oracle()
    rng = np.random.default_rng(3)
    for N in (1, 5, 1000, 100003):
        lw = rng.standard_normal(N) * 30 - 500
        got, lnZ, ess = S.normalize_log(lw)
        ref = lw.copy()
        lz = L.orc_normalize_log(P(ref), N)
        assert lnZ == pytest.approx(lz, rel=1e-13)
        np.testing.assert_allclose(got, ref, rtol=0, atol=1e-9)
        w = np.exp(ref)
        assert ess == pytest.approx(1.0 / np.sum((w / w.sum()) ** 2), rel=1e-10)
    with pytest.raises(S.SmcError):
        S.normalize_log(np.full(4, -np.inf))


def _rows(n, R, seqlen, seed, p_missing=0.0, p_unphased=0.0, p_var=0.3, partial=0.0, dup=False):
    rng = np.random.default_rng(seed)
    x_end = np.sort(rng.choice(np.arange(50, int(seqlen)), size=R - 1, replace=False)).astype(np.float64)
    x_end = np.append(x_end, float(seqlen))
    if dup:      # two typed sites at one position: zero-length rows are legal .seg input
        x_end[5] = x_end[4]
        x_end[R // 2] = x_end[R // 2 - 1]
    al = (rng.random((R, n)) < p_var).astype(np.int8)
    for r in range(R):
        for i in range(0, n - 1, 2):
            if rng.random() < p_unphased:
                al[r, i] = al[r, i + 1] = 2
        for i in range(0, n - 1, 2):
            if rng.random() < p_missing:
                al[r, i] = al[r, i + 1] = -1
    al[R - 1] = -1
    is_site = (rng.random(R) >= partial).astype(np.uint8)
    return x_end, al, is_site


CASES = {
    # name: (n, N, R, seqlen, t_start, inv2N, lags, row kwargs)
    "const4": (4, 96, 60, 60000.0, [0.0], [1 / 2e4], [5000.0], dict(dup=True)),
    "epochs3_n4_missing": (4, 130, 80, 80000.0, [0.0, 4000.0, 20000.0], [1 / 2e4, 1 / 8e3, 1 / 3e4],
                           [9000.0, 2500.0, 800.0], dict(p_missing=0.2, partial=0.2)),
    "zigzag11_n8": (8, 200, 50, 40000.0, [4e4 * t for t in (0, .0025, .005, .01, .02, .04, .08, .16, .32, .64, 1.28)],
                    [1 / (2e4 * s) for s in (1, .2, 1, .2, 1, .2, 1, .2, 1, .2, 1)],
                    [20000.0, 15000.0, 9000.0, 7000.0, 5000.0, 3000.0, 2000.0, 1500.0, 1000.0, 600.0, 300.0],
                    dict(p_unphased=0.15)),
    "n16_small": (16, 40, 25, 20000.0, [0.0, 8000.0], [1 / 2e4, 1 / 4e4], [4000.0, 1000.0], dict(p_missing=0.1)),
    "n2": (2, 64, 40, 50000.0, [0.0, 10000.0], [1 / 2e4, 1 / 2e4], [6000.0, 6000.0], {}),
    # the limits of the ABI (SMC_MAX_HAPLOTYPES = 32, SMC_MAX_EPOCHS = 64) and degenerate shapes
    "n32_max": (32, 24, 12, 6000.0, [0.0, 6000.0], [1 / 2e4, 1 / 3e4], [1500.0, 700.0], dict(p_missing=0.05, p_unphased=0.05)),
    "E64_max": (4, 48, 30, 30000.0, [250.0 * k * (1 + 0.05 * k) for k in range(64)], [1 / (2e4 * (1 + (k % 3))) for k in range(64)],
                [4000.0 - 50.0 * k for k in range(64)], {}),
    "N1_single_particle": (4, 1, 25, 25000.0, [0.0], [1 / 2e4], [3000.0], {}),
    "R1_single_row": (4, 64, 1, 5000.0, [0.0, 9000.0], [1 / 2e4, 1 / 2e4], [1000.0, 1000.0], {}),
    "all_missing": (4, 64, 30, 30000.0, [0.0, 9000.0], [1 / 2e4, 1 / 2e4], [4000.0, 1000.0], dict(p_missing=1.0)),
}
NO_RESAMPLING = {"N1_single_particle", "R1_single_row", "all_missing"}   # uniform or single weights: ESS never drops


@pytest.mark.parametrize("name", sorted(CASES))
def test_full_filter_vs_oracle(S, name):
    """The whole per-chunk sweep (prior trees, SMC' extension, pruning, normalisation, ESS, systematic resampling +
    gather + re-draw, lagged tallies) against the CPU restatement on the same Philox streams."""
    n, N, R, seqlen, t_start, inv2N, lags, kw = CASES[name]
    mu, rho = 2.5e-8, 1e-8
    if name == "zigzag11_n8":
        mu, rho = 1e-7, 4e-8    # more events per row
    x_end, al, is_site = _rows(n, R, seqlen, seed=len(name), **kw)
    rng = np.random.default_rng(1)
    max_rec = None
    if name == "epochs3_n4_missing":
        max_rec = rng.integers(-1, 3, size=R).astype(np.int32)
    pf = S.ParticleFilter(n, N, t_start, inv2N, mu, rho, seqlen, lags, seed=2024, chunk=3, parity_scan=True,
                          ancestral_aware=(name == "n2"))
    pf.upload_rows(x_end, al, is_site, max_rec)
    pf.init_particles()
    half = R // 2
    pf.sweep(0, half)
    pf.sweep(half, R)
    pf.finish()
    parent, height, logw, nb = pf.particles()
    ref = _libs.oracle_filter(_libs.make_model(n, t_start, inv2N, mu, rho, seqlen), N, 2024, 3, 0.0, 0.5, lags, x_end,
                              al.astype(np.int32), is_site.astype(np.int32), max_rec,
                              ancestral_aware=int(name == "n2"))
    c = pf.counters()
    assert c["err"] == 0 and c["rows"] == R
    assert c["resamples"] == ref["n_resample"] and (ref["n_resample"] > 0 or name in NO_RESAMPLING)
    assert c["recombinations"] == ref["n_recomb"] or max_rec is not None   # the oracle counts RECORDED events
    assert np.array_equal(parent, ref["parent"])
    np.testing.assert_allclose(height, ref["height"], rtol=1e-11, atol=0)
    np.testing.assert_allclose(nb, ref["next_base"], rtol=1e-11, atol=0)
    np.testing.assert_allclose(logw, ref["logw"], rtol=0, atol=1e-9)
    assert pf.ln_normalization_factor() == pytest.approx(ref["lnZ"], rel=1e-11)
    pos, ess = pf.resample_log()
    assert np.array_equal(pos, ref["rs_pos"])
    np.testing.assert_allclose(ess, ref["rs_ess"], rtol=1e-9)
    t = pf.tallies()
    for q in (0, 1, 3, 4):       # counts and opportunities (linear in the weights)
        np.testing.assert_allclose(t[q], ref["tallies"][q], rtol=1e-10, atol=1e-13)
    assert t[4].sum() > 0 and (name == "R1_single_row" or (t[1].sum() > 0 and t[3].sum() > 0))
    if max_rec is None:          # per-row opportunity of the local recombination map (filtering weights)
        ro = pf.row_opportunity()
        np.testing.assert_allclose(ro, ref["row_opp"], rtol=1e-10, atol=0)
        assert np.all(ro[np.diff(x_end, prepend=0.0) > 0] > 0) and np.all(ro[np.diff(x_end, prepend=0.0) == 0] == 0)
    # weight^2 tallies: copies do not share records, so sum(w^2 opp) is a lower bound of the reference's value
    assert np.all(t[2] <= ref["tallies"][2] * (1 + 1e-9) + 1e-300)
    assert np.all(t[5] <= ref["tallies"][5] * (1 + 1e-9) + 1e-300)
    # ... and close to it since the weights are pooled per ancestor slot of every closed bucket (what is missed: sharing
    # among siblings born while a bucket was open, and the bucket open at the end -- large on these short sequences)
    for q in (2, 5):
        if c["resamples"] > 0 and ref["tallies"][q].sum() > 0:
            assert t[q].sum() >= 0.3 * ref["tallies"][q].sum(), (q, t[q].sum(), ref["tallies"][q].sum())
    # exactly what the engine defines (oracle/smc_tally_oracle.c, pooled at the row in which the record's update window
    # was passed): the pooling kernels, the two-launch hand-over and the end-of-data passes add up to the same numbers
    if max_rec is None:
        np.testing.assert_allclose(t[2], ref["w2_pooled_at_close"][0], rtol=1e-9, atol=1e-300)
        np.testing.assert_allclose(t[5], ref["w2_pooled_at_close"][1], rtol=1e-9, atol=1e-300)
    pf.close()


def test_mid_sweep_state_and_ancestry(S):
    """particles() between sweeps applies the pending resampling; the ancestry it reports is the oracle's"""
    n, N, R, seqlen, t_start, inv2N, lags, kw = CASES["const4"]
    x_end, al, is_site = _rows(n, R, seqlen, seed=77)
    mu, rho = 2.5e-8, 1e-8
    m = _libs.make_model(n, t_start, inv2N, mu, rho, seqlen)
    for stop in (7, 23):
        pf = S.ParticleFilter(n, N, t_start, inv2N, mu, rho, seqlen, lags, seed=5, parity_scan=True)
        pf.upload_rows(x_end, al, is_site)
        pf.init_particles()
        pf.sweep(0, stop)
        parent, height, logw, nb = pf.particles()
        # the oracle stops BEFORE applying a resampling decided at its last row; compare when none was pending
        ref2 = _libs.oracle_filter(m, N, 5, 0, 0.0, 0.5, lags, x_end[:stop], al[:stop].astype(np.int32),
                                   is_site[:stop].astype(np.int32))
        if ref2["n_resample"] == 0 or not np.isclose(ref2["rs_pos"][-1], x_end[stop - 1]):
            assert np.array_equal(parent, ref2["parent"])
            np.testing.assert_allclose(logw, ref2["logw"], atol=1e-9)
        assert np.exp(logw).sum() == pytest.approx(1.0, rel=1e-9)
        # continuing after the read-out gives the same end state as an uninterrupted sweep
        pf.sweep(stop, R)
        pf.finish()
        full = _libs.oracle_filter(m, N, 5, 0, 0.0, 0.5, lags, x_end, al.astype(np.int32), is_site.astype(np.int32))
        p2, h2, lw2, _ = pf.particles()
        assert np.array_equal(p2, full["parent"])
        np.testing.assert_allclose(pf.tallies()[[0, 1, 3, 4]], full["tallies"][[0, 1, 3, 4]], rtol=1e-10, atol=1e-13)
        pf.close()


@pytest.mark.parametrize("name", ["const4", "epochs3_n4_missing", "zigzag11_n8"])
def test_getter_between_last_row_and_finish(S, name):
    """sweep(0, R); particles(); finish() -- the header does not forbid it.  The read-out after the last row must not
    consume the buckets smc_finish's flush is about to consume (they were counted twice), and when it applies a
    resampling decided at the last row the ancestor slots of those buckets must travel with the offspring."""
    n, N, R, seqlen, t_start, inv2N, lags, kw = CASES[name]
    mu, rho = (1e-7, 4e-8) if name == "zigzag11_n8" else (2.5e-8, 1e-8)
    x_end, al, is_site = _rows(n, R, seqlen, seed=len(name), **kw)
    out = {}
    for getter in (False, True):
        pf = S.ParticleFilter(n, N, t_start, inv2N, mu, rho, seqlen, lags, seed=2024, chunk=3, parity_scan=True)
        pf.upload_rows(x_end, al, is_site)
        pf.init_particles()
        pf.sweep(0, R)
        if getter:
            parent, height, logw, nb = pf.particles()
            assert np.exp(logw).sum() == pytest.approx(1.0, rel=1e-9)
            pf.last_ancestry()
        pf.finish()
        pos, _ = pf.resample_log()
        out[getter] = (pf.tallies().copy(), pf.ln_normalization_factor(), len(pos) > 0 and pos[-1] == x_end[-1])
        pf.close()
    plain, read = out[False], out[True]
    assert read[1] == plain[1]
    if not plain[2]:
        # no resampling pending at the last row: the read-out only normalised the stored weights
        for q in (0, 1, 3, 4):
            np.testing.assert_allclose(read[0][q], plain[0][q], rtol=1e-10, atol=1e-13)
    else:
        # the read-out applied the pending resampling, so the flush weighted the remaining buckets with the offspring's
        # equal weights (as the reference does, smcsmc.cpp:346-373) instead of the parents': equal in expectation
        for q in (1, 4):
            assert read[0][q].sum() == pytest.approx(plain[0][q].sum(), rel=0.05)


def test_set_particles_fixed_tree_set(S):
    """pruning log-likelihoods and tallies over a FIXED tree set: rho so small that no recombination happens, so the
    weights are pure pruning/no-mutation factors and the tallies pure recombination opportunity of the given trees"""
    n, N = 8, 257
    parent, height = _cases.random_trees(n, N, seed=9)
    t_start, inv2N, lags = [0.0, 5000.0, 30000.0], [1 / 2e4] * 3, [3000.0, 2000.0, 1000.0]
    mu, rho, seqlen = 2.5e-8, 1e-30, 20000.0
    x_end, al, is_site = _rows(n, 30, seqlen, seed=4, p_missing=0.15, p_unphased=0.1)
    pf = S.ParticleFilter(n, N, t_start, inv2N, mu, rho, seqlen, lags, seed=1, ess_fraction=0.0)
    pf.upload_rows(x_end, al, is_site)
    logw0 = np.full(N, -np.log(N))
    pf.set_particles(parent, height, logw0, np.full(N, seqlen))
    pf.sweep()
    pf.finish()
    L = _libs.oracle()
    m = _libs.make_model(n, t_start, inv2N, mu, rho, seqlen)
    lw = logw0.copy()
    x0 = 0.0
    for r in range(len(x_end)):
        d = np.ascontiguousarray(al[r].astype(np.int32))
        for i in range(N):
            par, hei = np.ascontiguousarray(parent[i]), np.ascontiguousarray(height[i])
            nmiss = int((d == -1).sum())
            lobs = 0.0 if nmiss == n else (L.orc_tree_length(n, P(par), P(hei)) if nmiss == 0 else
                                           L.orc_branch_length_with_data(n, P(par), P(hei), P(d)))
            lw[i] += np.log(L.orc_fastexp(-mu * lobs * (x_end[r] - x0)))
            if is_site[r]:
                lw[i] += np.log(L.orc_site_likelihood(n, P(par), P(hei), P(d), mu, 0, 0))
        x0 = x_end[r]
    lnZ = np.log(np.exp(lw - lw.max()).sum()) + lw.max()
    _, _, logw, _ = pf.particles()
    np.testing.assert_allclose(logw, lw - lnZ, rtol=0, atol=1e-10)
    assert pf.ln_normalization_factor() == pytest.approx(lnZ, rel=1e-12)
    assert pf.counters()["resamples"] == 0 and pf.counters()["recombinations"] == 0
    pf.close()


def test_blocked_scan_mode_matches_parity_mode(S):
    """production mode (blocked parallel partial sums) gives the same filter as the order-faithful mode except where a
    resampling boundary falls within rounding of some u_j -- never, at these sizes and seeds"""
    n, N, R, seqlen, t_start, inv2N, lags, kw = CASES["epochs3_n4_missing"]
    x_end, al, is_site = _rows(n, R, seqlen, seed=12, **kw)
    out = []
    for parity in (True, False):
        pf = S.ParticleFilter(n, N, t_start, inv2N, 2.5e-8, 1e-8, seqlen, lags, seed=8, parity_scan=parity)
        pf.upload_rows(x_end, al, is_site)
        pf.init_particles()
        pf.sweep()
        pf.finish()
        out.append((pf.particles(), pf.ln_normalization_factor(), pf.tallies(), pf.counters()))
        pf.close()
    assert np.array_equal(out[0][0][0], out[1][0][0])
    np.testing.assert_allclose(out[0][0][1], out[1][0][1], rtol=1e-12)
    assert out[0][1] == pytest.approx(out[1][1], rel=1e-12)
    np.testing.assert_allclose(out[0][2][[0, 1, 3, 4]], out[1][2][[0, 1, 3, 4]], rtol=1e-10, atol=1e-13)
    assert out[0][3] == out[1][3]


def test_two_level_parent_search_equals_flat_search(S, monkeypatch):
    """production mode finds an offspring's parent by a search over block offsets followed by a search inside one
    block; SMC_FLAT_SEARCH=1 selects the plain binary search over all particles it must reproduce: identical
    ancestry, trees and weights (20 000 particles = 157 blocks, many resampling events)"""
    n, N = 4, 20000
    t_start, inv2N, lags = [0.0, 4000.0, 20000.0], [1 / 2e4, 1 / 8e3, 1 / 3e4], [3000.0, 1500.0, 800.0]
    R, seqlen = 60, 30000.0
    x_end, al, is_site = _rows(n, R, seqlen, seed=21)
    out = []
    for flat in (False, True):
        if flat:
            monkeypatch.setenv("SMC_FLAT_SEARCH", "1")
        pf = S.ParticleFilter(n, N, t_start, inv2N, 2.5e-8, 1e-8, seqlen, lags, seed=5, ess_fraction=0.9)
        pf.upload_rows(x_end, al, is_site)
        pf.init_particles()
        pf.sweep()
        out.append((pf.last_ancestry(), pf.particles(), pf.counters()))
        pf.close()
    assert out[0][2]["resamples"] > 10 and out[0][2] == out[1][2]
    assert np.array_equal(out[0][0], out[1][0])
    for a, b in zip(out[0][1], out[1][1]):
        assert np.array_equal(a, b)


def test_large_population_properties(S):
    """size-independent properties at a BASELINE-scale particle count (n=8, N=100k): weights normalised, trees valid
    (height-sorted binary trees), resampling ancestry monotone, opportunity tallies positive and consistent"""
    n, N = 8, 100000
    t_start = [4e4 * t for t in (0, .0025, .005, .01, .02, .04, .08, .16, .32, .64, 1.28)]
    inv2N = [1 / 2e4] * 11
    lags = [4000.0] * 11
    R, seqlen = 40, 16000.0
    x_end, al, is_site = _rows(n, R, seqlen, seed=2)
    pf = S.ParticleFilter(n, N, t_start, inv2N, 2.5e-8, 1e-8, seqlen, lags, seed=3)
    pf.upload_rows(x_end, al, is_site)
    pf.init_particles()
    pf.sweep(0, R - 1)
    parent, height, logw, nb = pf.particles()
    assert np.exp(logw).sum() == pytest.approx(1.0, rel=1e-9)
    nn = 2 * n - 1
    assert np.all(np.diff(height[:, n:], axis=1) >= 0) and np.all(height[:, :n] == 0)
    assert np.all(parent[:, nn - 1] == -1) and np.all(parent[:, :nn - 1] >= n)
    kids = np.zeros((N, nn), dtype=np.int32)
    for v in range(nn - 1):
        np.add.at(kids, (np.arange(N), parent[:, v]), 1)
    assert np.all(kids[:, n:] == 2)
    assert np.all(height[np.arange(N)[:, None], parent[:, :nn - 1]] >= height[:, :nn - 1])
    assert np.all(nb > x_end[R - 2]) and np.all(nb <= seqlen)
    c = pf.counters()
    assert c["resamples"] > 0
    anc = pf.last_ancestry()
    assert np.all(np.diff(anc) >= 0) and anc.min() >= 0 and anc.max() < N
    pf.sweep(R - 1, R)
    pf.finish()
    t = pf.tallies()
    # total recombination opportunity = E[tree length] * bp swept: sum over epochs of sum_i w_i len_e * dx
    assert t[4].sum() > 0 and t[3].sum() > 0 and t[1].sum() > 0
    mean_len = t[4].sum() / seqlen
    assert 2e4 * 2 < mean_len < 2e4 * 12          # E[tree length] = 4N sum_{k<n} 1/k = 2e4 * 2 * 2.59 for n=8 (prior)
    pf.close()


def test_sweep_many_concurrent_chunks(S):
    """several chunks swept concurrently on their own streams give exactly what separate sweeps give"""
    n, N, R, seqlen, t_start, inv2N, lags, kw = CASES["epochs3_n4_missing"]
    sets = [_rows(n, R - 7 * c, seqlen, seed=40 + c, **kw) for c in range(3)]

    def make(c):
        pf = S.ParticleFilter(n, N, t_start, inv2N, 2.5e-8, 1e-8, seqlen, lags, seed=6, chunk=c)
        pf.upload_rows(*sets[c])
        pf.init_particles()
        return pf
    a = [make(c) for c in range(3)]
    S.sweep_many(a, rows_per_turn=3)
    b = [make(c) for c in range(3)]
    for pf in b:
        pf.sweep()
    for x, y in zip(a, b):
        x.finish(); y.finish()
        assert x.counters() == y.counters()
        px, py = x.particles(), y.particles()
        assert np.array_equal(px[0], py[0]) and np.array_equal(px[1], py[1]) and np.array_equal(px[2], py[2])
        assert x.ln_normalization_factor() == y.ln_normalization_factor()
        np.testing.assert_allclose(x.tallies(), y.tallies(), rtol=1e-12, atol=1e-300)
        x.close(); y.close()


def test_errors(S):
    with pytest.raises(S.SmcError):
        S.ParticleFilter(1, 10, [0.0], [1.0], 1e-8, 1e-8, 1e5, [1e3])
    with pytest.raises(S.SmcError):
        S.ParticleFilter(4, 10, [0.0, 0.0], [1.0, 1.0], 1e-8, 1e-8, 1e5, [1e3, 1e3])
    pf = S.ParticleFilter(4, 10, [0.0], [1 / 2e4], 2.5e-8, 1e-8, 1e4, [1e3])
    with pytest.raises(S.SmcError):
        pf.init_particles()                                     # rows first
    with pytest.raises(S.SmcError):
        pf.upload_rows([10.0, 5.0], np.zeros((2, 4)), [1, 1])   # not increasing
    with pytest.raises(S.SmcError):
        pf.upload_rows([10.0], np.array([[2, 0, 0, 0]]), [1])   # unpaired unphased het
    pf.upload_rows([10.0, 1e4], np.zeros((2, 4)), [1, 1])
    with pytest.raises(S.SmcError):
        pf.sweep()                                              # particles not initialised
    pf.init_particles()
    with pytest.raises(S.SmcError):
        pf.sweep(1, 2)                                          # out of order
    with pytest.raises(S.SmcError):
        pf.finish()                                             # before the last row
    pf.sweep()
    pf.finish()
    with pytest.raises(S.SmcError):
        S.systematic_offspring(np.array([0.0, 1.0]), 1.5)
    pf.close()


def test_terminal_branch_quantiles_vs_reference_prior_trees(S):
    """calculate_terminal_branch_length_quantiles (smcsmc.cpp:128-167) on the device against the same statistic of the
    reference's own prior trees (tests/golden/tbl_quantiles.json, 2e5 trees through libsmcref): zigzag model, n = 8.
    In law: the quantile of an order statistic has relative error ~ 1 / sqrt(q * count)."""
    import json
    g = json.load(open(os.path.join(G, "tbl_quantiles.json")))
    t = [4e4 * x for x in (0, .0025, .005, .01, .02, .04, .08, .16, .32, .64, 1.28)]
    inv = [1 / (2e4 * s) for s in (1, .2, 1, .2, 1, .2, 1, .2, 1, .2, 1)]
    q = g["quantiles"]
    lengths, mean_tl = S.terminal_branch_quantiles(8, t, inv, quantiles=q, n_trees=1000000, seed=3)
    ref = np.array(g["lengths"])
    assert lengths.shape == ref.shape == (8, len(q))
    assert mean_tl == pytest.approx(g["mean_total_length"], rel=0.01)
    pooled, pooled_ref = lengths.mean(axis=0), ref.mean(axis=0)           # samples are exchangeable
    for k, x in enumerate(q):
        tol = 4.0 / np.sqrt(min(x, 1 - x) * g["count"]) + 0.01
        assert pooled[k] == pytest.approx(pooled_ref[k], rel=tol), (x, pooled[k], pooled_ref[k])
        np.testing.assert_allclose(lengths[:, k], ref[:, k], rtol=3 * tol)
    assert np.all(np.diff(lengths, axis=1) > 0)
    with pytest.raises(S.SmcError):
        S.terminal_branch_quantiles(8, t, inv, quantiles=[1.5], n_trees=1000)



def test_fp64_fma_probe(S):
    """smc_measure_fp64_fma (bench.py's fp64 denominator): a B200 sustains 30-40 TFLOP/s of DFMA; a value far outside
    means the probe kernel was optimised away or the clocks are capped"""
    v = S.measure_fp64_fma()
    assert 5.0 < v < 100.0, v
