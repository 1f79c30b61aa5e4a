#!/usr/bin/env python
"""Regenerate tests/golden/*.npz from the UNMODIFIED reference compiled into oracle/_ref/libsmcref.so
(oracle/build_ref.sh, needs /root/reference).  Test tooling only.  Run:  python tests/golden/make_golden.py

Every fixture stores the seeded inputs next to the reference's outputs, so the tests need neither the reference nor
this script at run time.  Reference entry points: see oracle/ref_harness.cpp.
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _libs  # noqa: E402
import _cases  # noqa: E402
from _libs import P  # noqa: E402

CMD = "smcsmc -nsam {n} -Np 10 -N0 10000 -t {t} -r {r} {L} -seed 1 {extra}"


def open_ref(n, extra="", L=1000000, mu=2.5e-8, rho=1e-8):
    return _libs.ref_open(CMD.format(n=n, t=4e4 * mu * L, r=4e4 * rho * L, L=L, extra=extra))


def golden_prune():
    R = _libs.ref()
    out = {}
    for n in (2, 3, 4, 8, 16):
        h = open_ref(n)
        mu = R.ref_mutation_rate(h)
        count = 64
        parent, height = _cases.random_trees(n, count, seed=100 + n)
        # stretch some trees so that branch*mu crosses the fastexp switch-over (x^2 >= 0.516)
        height[count // 2:] *= np.logspace(0, 4.5, count - count // 2)[:, None]
        data = _cases.random_alleles(n, count, seed=200 + n, p_unphased=0.3 if n >= 4 else 0.0)
        phased = _cases.random_alleles(n, count, seed=300 + n)
        nn = 2 * n - 1
        lik = np.zeros(count); lik_aa = np.zeros(count); site = np.zeros(count); site_aa = np.zeros(count)
        bl = np.zeros(count); tl = np.zeros(count); nomut = np.zeros(count)
        for t in range(count):
            v = ctypes.c_double()
            par, hei = np.ascontiguousarray(parent[t]), np.ascontiguousarray(height[t])
            assert R.ref_prune(h, n, nn, P(par), P(hei), P(phased[t]), 0, ctypes.byref(v)) == 0
            lik[t] = v.value
            assert R.ref_prune(h, n, nn, P(par), P(hei), P(phased[t]), 1, ctypes.byref(v)) == 0
            lik_aa[t] = v.value
            assert R.ref_site_likelihood(h, n, nn, P(par), P(hei), P(data[t]), 0, 0, ctypes.byref(v)) == 0
            site[t] = v.value
            assert R.ref_site_likelihood(h, n, nn, P(par), P(hei), P(data[t]), 1, 0, ctypes.byref(v)) == 0
            site_aa[t] = v.value
            a, b = ctypes.c_double(), ctypes.c_double()
            assert R.ref_branch_lengths(h, n, nn, P(par), P(hei), P(phased[t]), ctypes.byref(a), ctypes.byref(b)) == 0
            bl[t], tl[t] = a.value, b.value
            assert R.ref_extend_nomut(h, n, nn, P(par), P(hei), P(phased[t]), 10.0, 10.0 + 37.0 * (t + 1),
                                      ctypes.byref(v)) == 0
            nomut[t] = v.value
        out.update({f"n{n}_parent": parent, f"n{n}_height": height, f"n{n}_data": data, f"n{n}_phased": phased,
                    f"n{n}_lik": lik, f"n{n}_lik_aa": lik_aa, f"n{n}_site": site, f"n{n}_site_aa": site_aa,
                    f"n{n}_bl": bl, f"n{n}_tl": tl, f"n{n}_nomut": nomut, f"n{n}_mu": np.array(mu)})
        R.ref_close(h)
    # the reference's own example tree (scrm/src/forest-debug.cc:57-121) and the allele patterns of
    # test/old/unittest/test_FroestState.cpp:135-199
    h = open_ref(4)
    par = np.array([4, 4, 5, 5, 6, 6, -1], dtype=np.int32)
    hei = np.array([0, 0, 0, 0, 1, 3, 10], dtype=np.float64)
    pats = np.array([[1, 1, 1, 1], [0, 1, 1, 1], [-1, 1, 1, 1], [0, 1, 0, 1], [0, 1, -1, 1], [0, 1, -1, -1],
                     [-1, -1, 0, 1], [-1, -1, 0, -1], [-1, 1, 0, -1], [-1, -1, -1, -1], [1, -1, -1, 0]], dtype=np.int32)
    ex_bl = np.zeros(len(pats)); ex_tl = np.zeros(len(pats)); ex_lik = np.zeros(len(pats))
    for i, d in enumerate(pats):
        a, b, v = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
        d = np.ascontiguousarray(d)
        assert R.ref_branch_lengths(h, 4, 7, P(par), P(hei), P(d), ctypes.byref(a), ctypes.byref(b)) == 0
        assert R.ref_prune(h, 4, 7, P(par), P(hei * 1e4), P(d), 0, ctypes.byref(v)) == 0
        ex_bl[i], ex_tl[i], ex_lik[i] = a.value, b.value, v.value
    out.update({"ex_parent": par, "ex_height": hei, "ex_pats": pats, "ex_bl": ex_bl, "ex_tl": ex_tl,
                "ex_lik_scaled1e4": ex_lik, "ex_mu": np.array(R.ref_mutation_rate(h))})
    R.ref_close(h)
    np.savez_compressed(os.path.join(HERE, "prune.npz"), **out)


def golden_resampling():
    R = _libs.ref()
    h = open_ref(4, "-ESS 0.5")
    out = {}
    k = 0
    for N in (1, 2, 7, 100, 1000, 4096):
        for kind in ("uniform", "mild", "lognormal", "degenerate", "zeros"):
            w = _cases.skewed_weights(N, seed=1000 + k, kind=kind)
            ps = _cases.sequential_partial_sums(w)
            for u in (1e-300, 0.25, 0.5, 0.9999999999, float(np.random.default_rng(k).random())):
                counts = np.zeros(N, dtype=np.int32)
                assert R.ref_systematic(h, P(ps), N, N, ctypes.c_double(u), P(counts)) == 0
                assert counts.sum() == N
                out[f"c{k}_ps"] = ps; out[f"c{k}_u"] = np.array(u); out[f"c{k}_counts"] = counts
                k += 1
    out["ncases"] = np.array(k)
    # full weight path: normalise -> ESS test -> systematic resampling -> multiplicities and child weights
    j = 0
    for N in (10, 333, 2000):
        for kind in ("uniform", "mild", "lognormal"):
            rng = np.random.default_rng(5000 + j)
            w = _cases.skewed_weights(N, seed=2000 + j, kind=kind) * 0.37     # un-normalised on purpose
            u = float(rng.random())
            lnZ, res = ctypes.c_double(), ctypes.c_int()
            mult = np.zeros(N, dtype=np.int32); wout = np.zeros(N)
            assert R.ref_weights_path(h, N, P(w), P(w), ctypes.c_double(u), ctypes.byref(lnZ), ctypes.byref(res),
                                      P(mult), P(wout)) == 0
            out[f"p{j}_w"] = w; out[f"p{j}_u"] = np.array(u); out[f"p{j}_lnZ"] = np.array(lnZ.value)
            out[f"p{j}_resampled"] = np.array(res.value); out[f"p{j}_mult"] = mult; out[f"p{j}_wout"] = wout
            j += 1
    out["npaths"] = np.array(j)
    R.ref_close(h)
    np.savez_compressed(os.path.join(HERE, "resampling.npz"), **out)


def golden_tallies():
    R = _libs.ref()
    out = {}
    k = 0
    for (N, E, Rows, seed) in ((4, 1, 30, 1), (6, 3, 60, 2), (16, 5, 80, 3)):
        times = " ".join("-eN {} 1".format(0.1 * e) for e in range(1, E))
        h = open_ref(4, times)
        assert R.ref_num_epochs(h) == E
        seqlen = 1e6
        lags = np.array([2e5 / (e + 1) for e in range(E)])
        ts, inv = np.zeros(E), np.zeros(E)
        R.ref_epoch_table(h, P(ts), P(inv))
        script, row_end, weights, anc, w_final, evs = _cases.tally_script(N, E, Rows, seed, lags, seqlen, ts)
        res = np.zeros(6 * E)
        assert R.ref_tally_script(h, N, P(lags), P(script), len(script), P(res)) == 0, R.ref_last_error()
        out[f"t{k}_N"] = np.array(N); out[f"t{k}_E"] = np.array(E); out[f"t{k}_lags"] = lags
        out[f"t{k}_row_end"] = row_end; out[f"t{k}_weights"] = weights; out[f"t{k}_anc"] = anc
        out[f"t{k}_w_final"] = w_final; out[f"t{k}_seqlen"] = np.array(seqlen)
        out[f"t{k}_evs"] = np.array(evs, dtype=np.float64)
        out[f"t{k}_tallies"] = res.reshape(6, E)
        R.ref_close(h)
        k += 1
    out["ncases"] = np.array(k)
    np.savez_compressed(os.path.join(HERE, "tallies.npz"), **out)


def golden_dephase():
    """-dephase (particleContainer.cpp:138-181): site likelihoods of the reference with dephase = true on PHASED data, on
    the trees of prune.npz; the driver implements the flag by recoding differing pairs as unphased hets"""
    R = _libs.ref()
    g = np.load(os.path.join(HERE, "prune.npz"))
    out = {}
    for n in (4, 8, 16):
        h = open_ref(n)
        parent, height, phased = g[f"n{n}_parent"], g[f"n{n}_height"], g[f"n{n}_phased"]
        nn = 2 * n - 1
        site = np.zeros(len(parent)); site_aa = np.zeros(len(parent))
        for t in range(len(parent)):
            v = ctypes.c_double()
            par, hei, d = (np.ascontiguousarray(x) for x in (parent[t], height[t], phased[t]))
            assert R.ref_site_likelihood(h, n, nn, P(par), P(hei), P(d), 0, 1, ctypes.byref(v)) == 0
            site[t] = v.value
            assert R.ref_site_likelihood(h, n, nn, P(par), P(hei), P(d), 1, 1, ctypes.byref(v)) == 0
            site_aa[t] = v.value
        out[f"n{n}_site_dephase"] = site; out[f"n{n}_site_dephase_aa"] = site_aa
        R.ref_close(h)
    np.savez_compressed(os.path.join(HERE, "dephase.npz"), **out)


def golden_smc2_cmdlines():
    """Command lines exactly as the reference's Python frontend emits them (smcsmc/populationmodels.py:
    Population.core_command_line, with and without -vb, for its own model classes) and the tables of the reference's
    parsed Model for each (scrm Param::parse + Model::finalize through libsmcref.so).  The CPU test feeds the same
    strings to the drop-in driver's -dump_model.  populationmodels.py is loaded on its own (the package __init__ pulls
    matplotlib, which this image lacks)."""
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location("populationmodels", "/root/reference/smcsmc/populationmodels.py")
    pm = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(pm)
    R = _libs.ref()
    cases = []
    two = dict(sample_populations=[1, 1, 1, 1, 2, 2, 2, 2])
    for cls, kw in ((pm.PopSingleConst, {}), (pm.PopSingleBottleneck, {}), (pm.PopSingleExpand, {}), (pm.Pop4, {}),
                    (pm.TwoPopUniDirMigr, two), (pm.TwoPopBiDirMigr, two), (pm.TwoPopSplitNoMigr, two)):
        for vb in (False, True):
            pop = cls(**kw)
            cmd = pop.core_command_line(vb=vb)
            n = pop.num_samples
            h = _libs.ref_open("smcsmc -nsam %d -Np 10 -tmax 4 %s -seed 1" % (n, cmd))
            E, Pn = R.ref_num_epochs(h), R.ref_num_pops(h)
            ts, inv = np.zeros(E), np.zeros((E, Pn))
            R.ref_epoch_table(h, P(ts), P(inv))
            mig, single = np.zeros((E, Pn, Pn)), np.zeros((E, Pn, Pn))
            R.ref_mig_table(h, P(mig), P(single))
            pops = np.zeros(n, dtype=np.int32)
            R.ref_sample_pops(h, P(pops))
            cases.append({"model": cls.__name__, "vb": vb, "nsam": n, "cmd": cmd, "t_start": ts.tolist(),
                          "inv2N": inv.tolist(), "mig": mig.tolist(), "single_mig": single.tolist(),
                          "leaf_pop": pops.tolist(), "mu": R.ref_mutation_rate(h), "rho": R.ref_recombination_rate(h)})
            R.ref_close(h)
    # further command lines a user (or smc2 -P / -p) may write: PSMC-style patterns, options out of time order,
    # per-population sizes, symmetric and single migration rates, partial joins, admixture
    extra = [
        (4, "-N0 10000 -t 1000 -r 400 1000000 -p 1*3+15*4+1 -tmax 4"),
        (4, "-N0 10000 -t 1000 -r 400 1000000 -p 4+25*2+4+6 -tmax 2"),
        (4, "-N0 14312 -t 1000 -r 400 1000000 -eN 0.5 2 -eN 0.1 0.5 -eN 0 3"),
        (6, "-N0 10000 -t 1000 -r 400 1000000 -I 2 3 3 2.0 -en 0.05 2 0.25 -eM 0.1 0.5 -en 0.2 1 4 -ej 0.7 2 1"),
        (6, "-N0 10000 -t 1000 -r 400 1000000 -I 3 2 2 2 -ema 0 x 1 2 3 x 4 5 6 x -eps 0.2 3 1 0.4 -ej 0.5 3 2 -ej 0.9 2 1"),
        (6, "-N0 10000 -t 1000 -r 400 1000000 -I 2 3 3 -em 0 1 2 0.7 -es 0.05 2 0.6 -en 0.05 3 0.3 -ej 0.3 3 1 -ej 0.6 2 1"),
        (8, "-N0 10000 -t 1000 -r 400 1000000 -I 2 4 4 -ej 0.5 2 1 -eN 0.5 2 -eN 1.0 1"),
        # -N0 is read as an integer (Model::set_default_pop_size(int))
        (4, "-N0 14312.7 -t 1000 -r 400 1000000 -eN 0.1 2 -eN 0.3 0.5"),
        # -r before -t: the mutation rate is divided by the final locus length directly
        (4, "-N0 5000 -r 400 1000000 -t 1000 -eN 0.2 3"),
    ]
    for n, cmd in extra:
        h = _libs.ref_open("smcsmc -nsam %d -Np 10 %s -seed 1" % (n, cmd))
        E, Pn = R.ref_num_epochs(h), R.ref_num_pops(h)
        ts, inv = np.zeros(E), np.zeros((E, Pn))
        R.ref_epoch_table(h, P(ts), P(inv))
        mig, single = np.zeros((E, Pn, Pn)), np.zeros((E, Pn, Pn))
        R.ref_mig_table(h, P(mig), P(single))
        pops = np.zeros(n, dtype=np.int32)
        R.ref_sample_pops(h, P(pops))
        cases.append({"model": "extra", "vb": False, "nsam": n, "cmd": cmd, "tmax_in_cmd": True, "t_start": ts.tolist(),
                      "inv2N": inv.tolist(), "mig": mig.tolist(), "single_mig": single.tolist(),
                      "leaf_pop": pops.tolist(), "mu": R.ref_mutation_rate(h), "rho": R.ref_recombination_rate(h)})
        R.ref_close(h)
    json.dump(cases, open(os.path.join(HERE, "smc2_cmdlines.json"), "w"), indent=1)


MOMENT_MODELS = {
    # name: (n, scrm demography flags)
    "const4": (4, ""),
    "bottleneck8": (8, "-eN 0.05 0.2 -eN 0.2 2"),
    "zigzag8": (8, "-eN 0.0025 0.2 -eN 0.005 1 -eN 0.01 0.2 -eN 0.02 1 -eN 0.04 0.2 -eN 0.08 1 -eN 0.16 0.2 "
                   "-eN 0.32 1 -eN 0.64 0.2 -eN 1.28 1"),
}


STRUCT_MODELS = {
    # name: (n, scrm flags) -- structured models (SURVEY.md 8d config 3 and simpler ones around it)
    "island2": (4, "-I 2 2 2 1.0"),
    "split2": (4, "-I 2 2 2 -ej 0.5 2 1"),
    "config3": (8, "-I 2 4 4 -em 0 2 1 1 -ej 0.5 2 1"),
    "island2_epochs": (6, "-I 2 3 3 0.5 -en 0.05 1 0.3 -en 0.05 2 2 -eM 0.2 3.0 -eN 0.6 1.5"),
    "three": (6, "-I 3 2 2 2 -em 0.1 2 1 2.0 -ej 0.3 3 2 -ej 0.8 2 1"),
    # admixture: at 0.1 a lineage of population 1 moves to a new population 3 with probability 0.7 (-es t i p keeps p)
    "admix3": (6, "-I 2 3 3 0.5 -es 0.1 1 0.3 -ej 0.4 3 2 -ej 0.9 2 1"),
}


def golden_struct_moments(count=40000, walk=20000):
    """the reference's prior trees and ARG walks under structured models: TMRCA, local tree length, recombination
    heights, gaps -- mean / standard error (batch means for the autocorrelated walk) for tests/test_struct_oracle.py;
    the model tables the oracle needs are stored alongside"""
    R = _libs.ref()
    out = {}
    for name, (n, dem) in STRUCT_MODELS.items():
        h = open_ref(n, dem, L=100000000)
        E, Pn = R.ref_num_epochs(h), R.ref_num_pops(h)
        ts, inv = np.zeros(E), np.zeros((E, Pn))
        R.ref_epoch_table(h, P(ts), P(inv))
        mig, single = np.zeros((E, Pn, Pn)), np.zeros((E, Pn, Pn))
        R.ref_mig_table(h, P(mig), P(single))
        pops = np.zeros(n, dtype=np.int32)
        R.ref_sample_pops(h, P(pops))
        tm, tl = np.zeros(count), np.zeros(count)
        assert R.ref_prior_stats(h, count, P(tm), P(tl)) == 0, R.ref_last_error()
        out[f"{name}_n"] = np.array(n); out[f"{name}_t_start"] = ts; out[f"{name}_inv2N"] = inv
        out[f"{name}_mig"] = mig; out[f"{name}_single"] = single; out[f"{name}_leaf_pop"] = pops
        out[f"{name}_prior"] = np.array([[x.mean(), x.std(ddof=1) / np.sqrt(count)] for x in (tm, tl, tm ** 2, tl ** 2)])
        pos = np.zeros(walk); rech = np.zeros(walk); tmr = np.zeros(walk); tln = np.zeros(walk)
        got = R.ref_arg_walk(h, walk, P(pos), P(rech), P(tmr), P(tln))
        assert got > 1000, got
        pos, rech, tmr, tln = pos[:got], rech[:got], tmr[:got], tln[:got]

        def bm(x):
            b = np.array_split(x, 50)
            mm = np.array([c.mean() for c in b])
            return [x.mean(), mm.std(ddof=1) / np.sqrt(len(mm))]
        out[f"{name}_walk"] = np.array([bm(np.diff(pos)), bm(rech), bm(tmr[1:]), bm(tln[1:])])
        out[f"{name}_rho"] = np.array(R.ref_recombination_rate(h)); out[f"{name}_mu"] = np.array(R.ref_mutation_rate(h))
        R.ref_close(h)
    np.savez_compressed(os.path.join(HERE, "struct_moments.npz"), **out)


def golden_moments(count=40000, reps=40000):
    """Sample moments of the reference's stochastic pieces (its own mt19937_64 stream): prior trees
    (Forest::buildInitialTree), one SMC' move from a fixed start tree (Forest::sampleNextGenealogy) and the prior ARG
    walk (sampleNextGenealogy + sampleRecSeqPosition).  Stored as mean / standard error for z-tests."""
    R = _libs.ref()
    out = {}
    for name, (n, dem) in MOMENT_MODELS.items():
        h = open_ref(n, dem, L=100000000)
        E = R.ref_num_epochs(h)
        ts, inv = np.zeros(E), np.zeros(E)
        R.ref_epoch_table(h, P(ts), P(inv))
        nn = 2 * n - 1
        parent = np.zeros((count, nn), dtype=np.int32); height = np.zeros((count, nn)); tl = np.zeros(count)
        assert R.ref_prior_trees(h, count, P(parent), P(height), P(tl)) == 0, R.ref_last_error()
        tm = height.max(axis=1)
        first = height[:, n]
        out[f"{name}_n"] = np.array(n); out[f"{name}_t_start"] = ts; out[f"{name}_inv2N"] = inv
        out[f"{name}_prior"] = np.array([[x.mean(), x.std(ddof=1) / np.sqrt(count)] for x in (tm, tl, first, tm ** 2)])
        # one SMC' move from a fixed start tree (the first prior tree)
        p0, h0 = np.ascontiguousarray(parent[0]), np.ascontiguousarray(height[0])
        mp = np.zeros((reps, nn), dtype=np.int32); mh = np.zeros((reps, nn)); rh = np.zeros(reps)
        assert R.ref_smc_moves(h, n, P(p0), P(h0), reps, P(mp), P(mh), P(rh)) == 0, R.ref_last_error()
        mtl = np.array([(mh[i][mp[i][mp[i] >= 0]] - mh[i][mp[i] >= 0]).sum() for i in range(reps)])
        mtm = mh.max(axis=1)
        unchanged = (np.abs(mh - h0[None, :]).max(axis=1) < 1e-9 * h0.max()).astype(float)
        newroot = (mtm > h0.max() * (1 + 1e-12)).astype(float)
        out[f"{name}_move_start_parent"] = p0; out[f"{name}_move_start_height"] = h0
        out[f"{name}_move"] = np.array([[x.mean(), x.std(ddof=1) / np.sqrt(reps)] for x in (rh, mtl, mtm, unchanged, newroot)])
        # ARG walk along the sequence
        cnt = 20000
        pos = np.zeros(cnt); rech = np.zeros(cnt); tmr = np.zeros(cnt); tln = np.zeros(cnt)
        got = R.ref_arg_walk(h, cnt, P(pos), P(rech), P(tmr), P(tln))
        assert got > 1000, got
        pos, rech, tmr, tln = pos[:got], rech[:got], tmr[:got], tln[:got]
        gaps = np.diff(pos)
        # autocorrelated along the sequence: use batch means (50 batches) for the standard errors
        def bm(x):
            b = np.array_split(x, 50)
            m = np.array([c.mean() for c in b])
            return [x.mean(), m.std(ddof=1) / np.sqrt(len(m))]
        out[f"{name}_walk"] = np.array([bm(gaps), bm(rech), bm(tmr[1:]), bm(tln[1:])])
        out[f"{name}_rho"] = np.array(R.ref_recombination_rate(h)); out[f"{name}_mu"] = np.array(R.ref_mutation_rate(h))
        R.ref_close(h)
    np.savez_compressed(os.path.join(HERE, "moments.npz"), **out)


if __name__ == "__main__":
    if _libs.ref() is None:
        sys.exit("oracle/_ref/libsmcref.so not available (needs /root/reference)")
    golden_prune()
    golden_resampling()
    golden_tallies()
    golden_moments()
    print("wrote", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


EM_BASE = ("-nsam 4 -Np 500 -N0 10000 -t 2000 -r 800 2000000 -eN 0 1 -eN 0.05 1 -eN 0.2 1 -eN 0.6 1 {extra} -seed {seed} "
           "-seg {seg} -o {out}")
EM_VARIANTS = {
    # one sweep, everything recorded (the smc2 mode: -EM 0 per chunk and iteration)
    "sweep": "-EM 0",
    # three in-binary EM iterations on the population sizes; recombination excluded from the EM (-xr) so that the
    # recombination rate stays at its input value (see DESIGN.md 7a on the reference's recombination opportunity)
    "em_ne": "-xr 0-3 -EM 2",
    # three in-binary EM iterations on the population sizes AND the recombination rate (CountModel::reset_recomb_rate,
    # count.cpp:296-325) -- comparable since the corrected reference tallies the full recombination opportunity
    "em_rho": "-EM 2",
}


def ref_binary(fixed=True):
    """oracle/_ref/smcsmc_fixed: the reference with ForestState::max_epoch_to_record_ initialised (two added lines,
    tools/patch_reference_uninit.py, built by oracle/build_fixed_ref.sh).  The unmodified binary reads that field
    uninitialised when a particle with multiplicity > 1 recombines and loses 15-20 % of the recombination opportunity,
    by an amount that changes with the process environment -- it cannot serve as a fixture source for
    anything that depends on the recombination opportunity."""
    d = os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle", "_ref")
    return os.path.join(d, "smcsmc_fixed" if fixed else "smcsmc")


def parse_out(path):
    rows = []
    with open(path) as f:
        hdr = f.readline().split()
        for line in f:
            rows.append(dict(zip(hdr, line.split())))
    return rows


# BASELINE.json configs[0] at full size: 4 haplotypes x 10 Mb simulated under a constant size, inferred with 8 log-spaced
# epochs from all-ones sizes, 1000 particles, 3 EM iterations (4 sweeps), seeds 1..10; EM on sizes and recombination rate
C0_BASE = ("-nsam 4 -Np 1000 -N0 10000 -t 10000 -r 4000 10000000 -eN 0 1 -eN 0.01 1 -eN 0.02 1 -eN 0.04 1 -eN 0.08 1 "
           "-eN 0.16 1 -eN 0.32 1 -eN 0.64 1 -EM 3 -seed {seed} -seg {seg} -o {out}")


def golden_configs0(seeds=range(1, 11)):
    """reference ensemble for tests/test_em_parity.py::test_configs0_full_size (about 5 core-minutes per seed)"""
    import gzip
    import json
    import shutil
    import subprocess
    import tempfile
    from concurrent.futures import ThreadPoolExecutor
    segz = os.path.join(os.path.dirname(HERE), "data", "const_n4_10Mb_seed11.seg.gz")
    binp = ref_binary()
    out = {"base": C0_BASE, "seeds": list(seeds), "binary": "smcsmc_fixed"}
    with tempfile.TemporaryDirectory() as tmp:
        seg = os.path.join(tmp, "c0.seg")
        with gzip.open(segz, "rb") as f, open(seg, "wb") as g:
            shutil.copyfileobj(f, g)

        def run(seed):
            o = os.path.join(tmp, "s%d" % seed)
            subprocess.run([binp] + C0_BASE.format(seed=seed, seg=seg, out=o).split(), check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            return parse_out(o + ".out")
        with ThreadPoolExecutor(8) as ex:
            out["runs"] = list(ex.map(run, seeds))
    json.dump(out, open(os.path.join(HERE, "configs0_reference.json"), "w"))


EM_STRUCT_BASE = ("-nsam 8 -Np 500 -N0 10000 -t 2000 -r 800 2000000 -I 2 4 4 -em 0 2 1 1 -eN 0.1 1 -ej 0.5 2 1 {extra} "
                  "-seed {seed} -seg {seg} -o {out}")
EM_STRUCT_VARIANTS = {
    # BASELINE.json configs[2] in small: two populations of 4 haplotypes, migration 2 -> 1 backwards in time, joined at
    # 0.5 * 4N0; inference epochs {0, 0.1, 0.5}; one sweep with everything recorded
    "sweep": "-EM 0",
    # three in-binary EM iterations on population sizes AND migration rates (recombination excluded)
    "em": "-xr 0-2 -EM 2",
    # the same with the recombination rate in the EM as well
    "em_all": "-EM 2",
}


def golden_em_struct(seeds=range(1, 13)):
    """The same for a structured model (tests/data/split_mig_n8_2Mb_seed13.seg, simulated by the reference's scrm under
    `-I 2 4 4 -em 0 2 1 1 -ej 0.5 2 1`): per-population Coal rows and the Migr rows of the unmodified binary."""
    import json
    import subprocess
    import tempfile
    from concurrent.futures import ThreadPoolExecutor
    seg = os.path.join(os.path.dirname(HERE), "data", "split_mig_n8_2Mb_seed13.seg")
    binp = ref_binary()
    out = {"base": EM_STRUCT_BASE, "variants": EM_STRUCT_VARIANTS, "seeds": list(seeds), "runs": {}, "binary": "smcsmc_fixed"}
    with tempfile.TemporaryDirectory() as tmp:
        for name, extra in EM_STRUCT_VARIANTS.items():
            def run(seed):
                o = os.path.join(tmp, "%s_s%d" % (name, seed))
                subprocess.run([binp] + EM_STRUCT_BASE.format(extra=extra, seed=seed, seg=seg, out=o).split(), check=True,
                               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                return parse_out(o + ".out")
            with ThreadPoolExecutor(8) as ex:
                out["runs"][name] = list(ex.map(run, seeds))
    json.dump(out, open(os.path.join(HERE, "em_struct_reference.json"), "w"))


def golden_em(seeds=range(1, 25)):
    """End-to-end reference runs (the unmodified smcsmc binary, oracle/_ref/smcsmc) on tests/data/const_n4_2Mb_seed7.seg:
    12 seeds per command variant; stores every .out row per seed (tests/test_em_parity.py compares the B200 driver's
    seed ensemble with this one)."""
    import json
    import subprocess
    import tempfile
    from concurrent.futures import ThreadPoolExecutor
    seg = os.path.join(os.path.dirname(HERE), "data", "const_n4_2Mb_seed7.seg")
    binp = ref_binary()
    out = {"base": EM_BASE, "variants": EM_VARIANTS, "seeds": list(seeds), "runs": {}, "binary": "smcsmc_fixed"}
    with tempfile.TemporaryDirectory() as tmp:
        for name, extra in EM_VARIANTS.items():
            def run(seed):
                o = os.path.join(tmp, "%s_s%d" % (name, seed))
                subprocess.run([binp] + EM_BASE.format(extra=extra, seed=seed, seg=seg, out=o).split(), check=True,
                               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                return parse_out(o + ".out")
            with ThreadPoolExecutor(8) as ex:
                out["runs"][name] = list(ex.map(run, seeds))

        # the UNMODIFIED binary on the sweep, twice per seed under environments of different size: its recombination
        # opportunity is low and not even a function of its inputs (the Coal rows and the recombination count are)
        def run_unmodified(arg):
            seed, pad = arg
            o = os.path.join(tmp, "u_s%d_%d" % (seed, pad))
            env = dict(os.environ, SMC_ENV_PAD="x" * pad)
            subprocess.run([ref_binary(False)] + EM_BASE.format(extra="-EM 0", seed=seed, seg=seg, out=o).split(),
                           check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, env=env)
            return parse_out(o + ".out")
        with ThreadPoolExecutor(8) as ex:
            out["unmodified_sweep"] = {str(pad): list(ex.map(run_unmodified, [(sd, pad) for sd in seeds]))
                                       for pad in (0, 64)}
    json.dump(out, open(os.path.join(HERE, "em_reference.json"), "w"))


def golden_spawn_probe():
    """oracle/ref_harness.cpp::ref_spawn_probe on the unmodified reference: what a spawned particle records for its
    first stretch as a function of the bytes its heap block held (tests/test_oracle_golden.py)."""
    import json
    cmd = "-nsam 4 -Np 10 -N0 10000 -t 2000 -r 800 2000000 -eN 0 1 -eN 0.05 1 -eN 0.2 1 -eN 0.6 1 -seed 3"
    h = _libs.ref_open("smcsmc " + cmd)
    out = {"cmd": cmd, "probes": []}
    for fill in (0x00, 0xFF, 0x7F, 0x01):
        for _ in range(4):
            r = _libs.ref_spawn_probe(h, fill)
            r["fill"] = fill
            out["probes"].append(r)
    json.dump(out, open(os.path.join(HERE, "spawn_probe.json"), "w"), indent=1)


# BASELINE.json configs[1] in medium: 8 haplotypes, the first 10 Mb of the zigzag bench data, the 11 zigzag epochs from
# all-ones sizes, 1000 particles, 5 EM iterations (6 sweeps) on sizes and recombination rate, seeds 1..10
C1_TIMES = [0, .0025, .005, .01, .02, .04, .08, .16, .32, .64, 1.28]
C1_BASE = ("-nsam 8 -Np 1000 -N0 10000 -t 10000 -r 4000 10000000 " + " ".join("-eN %r 1" % t for t in C1_TIMES) +
           " -EM 5 -seed {seed} -seg {seg} -o {out}")
# BASELINE.json configs[2] in medium: 4 + 4 haplotypes, the first 5 Mb of bench_data/split_mig_n8_10Mb_seed3.seg.gz,
# 2000 particles, 2 EM iterations (3 sweeps) on sizes, migration and recombination rates, seeds 1..10
C2_BASE = ("-nsam 8 -Np 2000 -N0 10000 -t 5000 -r 2000 5000000 -I 2 4 4 -em 0 2 1 1 -eN 0.1 1 -ej 0.5 2 1 -EM 2 "
           "-seed {seed} -seg {seg} -o {out}")


def cut_seg(gz_path, out_path, length):
    """the rows of a .seg.gz that start before `length` (the driver clips the last one to the locus length)"""
    import gzip
    with gzip.open(gz_path, "rt") as f, open(out_path, "w") as g:
        for line in f:
            if not line.startswith("#") and int(line.split("\t", 1)[0]) > length:
                break
            g.write(line)


def _golden_medium(name, base, gz, length, seeds):
    import json
    import subprocess
    import tempfile
    from concurrent.futures import ThreadPoolExecutor
    root = os.path.dirname(os.path.dirname(HERE))
    out = {"base": base, "seeds": list(seeds), "binary": "smcsmc_fixed", "data": gz, "length": length}
    with tempfile.TemporaryDirectory() as tmp:
        seg = os.path.join(tmp, "d.seg")
        cut_seg(os.path.join(root, gz), seg, length)

        def run(seed):
            o = os.path.join(tmp, "s%d" % seed)
            subprocess.run([ref_binary()] + base.format(seed=seed, seg=seg, out=o).split(), check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            return parse_out(o + ".out")
        with ThreadPoolExecutor(8) as ex:
            out["runs"] = list(ex.map(run, seeds))
        # the lags the reference calibrates for the first sweep (its calibration has its own fixed seed, so they depend
        # on the model only): a 2-particle run, "Lags set to:" on stderr
        o = os.path.join(tmp, "lagprobe")
        cmd = base.format(seed=1, seg=seg, out=o).split()
        cmd[cmd.index("-Np") + 1] = "2"
        cmd[cmd.index("-EM") + 1] = "0"
        r = subprocess.run([ref_binary()] + cmd, check=True, capture_output=True, text=True)
        line = [ln for ln in r.stderr.splitlines() if "Lags set to:" in ln][0]
        out["lags_iter0"] = [float(x) for x in line.split(":")[1].split()]
    json.dump(out, open(os.path.join(HERE, name), "w"))


def golden_configs1_medium(seeds=range(1, 11)):
    _golden_medium("configs1_medium_reference.json", C1_BASE, "bench_data/zigzag_n8_100Mb_seed1.seg.gz", 10000000, seeds)


def golden_configs2_medium(seeds=range(1, 11)):
    _golden_medium("configs2_medium_reference.json", C2_BASE, "bench_data/split_mig_n8_10Mb_seed3.seg.gz", 5000000, seeds)


def golden_smc2_frontend():
    """result.out of the reference's own frontend driving the reference binary (tests/smc2_helper.py; the same arguments
    the GPU test passes with -smcsmcpath smcsmc_b200/bin/smcsmc)"""
    import shutil
    import subprocess
    import tempfile
    root = os.path.dirname(os.path.dirname(HERE))
    seg = os.path.join(os.path.dirname(HERE), "data", "const_n4_2Mb_seed7.seg")
    with tempfile.TemporaryDirectory() as tmp:
        out = os.path.join(tmp, "run")
        subprocess.run([sys.executable, os.path.join(root, "tests", "smc2_helper.py"), ref_binary(), out, seg], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        shutil.copy(os.path.join(out, "result.out"), os.path.join(HERE, "smc2_frontend_reference.out"))


def golden_tbl_quantiles(count=200000):
    """Terminal branch lengths of the reference's own prior trees (Forest::buildInitialTree through libsmcref), sorted and
    read at the reference's quantile indices (calculate_terminal_branch_length_quantiles, smcsmc.cpp:128-167), for the
    zigzag model with 8 samples; tests/test_gpu_parity.py compares smc_terminal_branch_quantiles with it in law."""
    import json
    q = [0.001, 0.003, 0.01, 0.03, 0.1, 0.5, 0.95]
    cmd = ("-nsam 8 -Np 10 -N0 10000 -t 1000 -r 400 1000000 " +
           " ".join("-eN %r %r" % (t, s) for t, s in zip(C1_TIMES, [1, .2, 1, .2, 1, .2, 1, .2, 1, .2, 1])) + " -seed 5")
    h = _libs.ref_open("smcsmc " + cmd)
    L = _libs.ref()
    n, nn = 8, 15
    parent = np.zeros((count, nn), dtype=np.int32)
    height = np.zeros((count, nn))
    tl = np.zeros(count)
    assert L.ref_prior_trees(h, count, _libs.P(parent), _libs.P(height), _libs.P(tl)) == 0
    tbl = np.take_along_axis(height, parent[:, :n], axis=1)          # parent height of every leaf
    lengths = [[float(np.sort(tbl[:, v])[int(x * count)]) for x in q] for v in range(n)]
    json.dump({"cmd": cmd, "quantiles": q, "count": count, "lengths": lengths, "mean_total_length": float(tl.mean())},
              open(os.path.join(HERE, "tbl_quantiles.json"), "w"), indent=1)


def read_recomb_map(path, iteration=0):
    """<prefix>.recomb.gz -> (locus, opp_per_nt, rest) of one iteration (CountModel::dump_local_recomb_logs)"""
    import gzip
    loc, opp, rest = [], [], []
    with gzip.open(path, "rt") as f:
        header = f.readline().rstrip("\n").split("\t")
        for line in f:
            c = line.rstrip("\n").split("\t")
            if int(c[0]) != iteration:
                continue
            loc.append(float(c[1])); opp.append(float(c[3])); rest.append([float(v) for v in c[4:]])
    return header, np.array(loc), np.array(opp), np.array(rest)


def golden_recomb_map(seeds=range(1, 13), window=50000):
    """The opportunity column of the reference's local recombination map (<prefix>.recomb.gz) on
    tests/data/const_n4_2Mb_seed7.seg, one sweep per seed with the corrected binary, averaged over 50-kb windows;
    tests/test_gpu_parity.py compares the B200 driver's table with it window by window."""
    import json
    import subprocess
    import tempfile
    from concurrent.futures import ThreadPoolExecutor
    seg = os.path.join(os.path.dirname(HERE), "data", "const_n4_2Mb_seed7.seg")
    out = {"base": EM_BASE, "extra": "-EM 0", "seeds": list(seeds), "window": window, "binary": "smcsmc_fixed"}
    with tempfile.TemporaryDirectory() as tmp:
        def run(seed):
            o = os.path.join(tmp, "m_s%d" % seed)
            subprocess.run([ref_binary()] + EM_BASE.format(extra="-EM 0", seed=seed, seg=seg, out=o).split(), check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            header, loc, opp, rest = read_recomb_map(o + ".recomb.gz")
            rows = parse_out(o + ".out")
            total = [r for r in rows if r["Type"] == "Recomb"]
            k = int(window // 100)
            nw = len(opp) // k
            return {"header": header, "bins": int(len(opp)), "first_locus": float(loc[0]), "last_locus": float(loc[-1]),
                    "windows": opp[:nw * k].reshape(nw, k).mean(axis=1).tolist(), "sum_opp": float(opp.sum() * 100.0),
                    "out_recomb_opp": float(sum(float(r["Opp"]) for r in total))}
        with ThreadPoolExecutor(8) as ex:
            out["runs"] = list(ex.map(run, seeds))
    json.dump(out, open(os.path.join(HERE, "recomb_map_reference.json"), "w"))
