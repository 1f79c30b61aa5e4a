"""Seeded input generators shared by the golden-fixture script and the parity tests (checker side only)."""
import ctypes

import numpy as np

import _libs
from _libs import P, OrcParticle, OrcRng


def random_trees(n, count, seed, t_start=(0.0,), inv2N=(1.0 / 20000,), jitter=False):
    """`count` prior coalescent trees of n leaves from the oracle's Philox-driven sampler -> parent[count,2n-1] (int32),
    height[count,2n-1] (float64)."""
    L = _libs.oracle()
    m = _libs.make_model(n, list(t_start), list(inv2N), 2.5e-8, 1e-8, 1e6)
    nn = 2 * n - 1
    parent = np.zeros((count, nn), dtype=np.int32)
    height = np.zeros((count, nn), dtype=np.float64)
    p, r = OrcParticle(), OrcRng()
    for t in range(count):
        L.orc_rng_init(ctypes.byref(r), seed, t, 0, 0, 7)
        L.orc_init_particle(ctypes.byref(m), ctypes.byref(p), ctypes.byref(r), 1.0)
        parent[t] = np.ctypeslib.as_array(p.parent)[:nn]
        height[t] = np.ctypeslib.as_array(p.height)[:nn]
    return parent, height


def random_alleles(n, count, seed, p_missing=0.15, p_unphased=0.0):
    """allele vectors in {0,1,-1(missing),2(unphased het, in pairs)}; missing in an odd column implies missing in the
    preceding one is NOT enforced here (the pruning functions do not care)."""
    rng = np.random.default_rng(seed)
    d = rng.integers(0, 2, size=(count, n)).astype(np.int32)
    miss = rng.random((count, n)) < p_missing
    d[miss] = -1
    if p_unphased > 0:
        for c in range(count):
            for i in range(0, n - 1, 2):
                if rng.random() < p_unphased:
                    d[c, i] = d[c, i + 1] = 2
    return d


def skewed_weights(N, seed, kind):
    rng = np.random.default_rng(seed)
    if kind == "uniform":
        w = np.full(N, 1.0 / N)
    elif kind == "mild":
        w = rng.random(N) + 0.5
    elif kind == "lognormal":
        w = np.exp(3.0 * rng.standard_normal(N))
    elif kind == "degenerate":
        w = np.full(N, 1e-300)
        w[rng.integers(0, N)] = 1.0
    elif kind == "zeros":
        w = rng.random(N)
        w[rng.random(N) < 0.5] = 0.0
        w[0] = 0.3
    else:
        raise ValueError(kind)
    return w / w.sum()


def sequential_partial_sums(w):
    """left-to-right fp64 partial sums, as ParticleContainer::resample builds them (particleContainer.cpp:256-262)"""
    ps = np.zeros(len(w) + 1)
    s = 0.0
    for i, x in enumerate(w):
        ps[i] = s
        s += float(x)
    ps[len(w)] = s
    return ps


def tally_script(N, E, R, seed, lags, seqlen, t_start):
    """A random scripted particle history: per row some coal/recomb records per slot, an update with random weights,
    and (sometimes) a resampling with a random ancestry.  Returns the script for ref_tally_script (8 doubles per op)
    and the equivalent inputs for orc_lagged_tallies."""
    rng = np.random.default_rng(seed)
    ops = []
    evs = []   # (row, slot, type, epoch, cnt, x0, x1, opp)
    row_end = np.sort(rng.random(R)) * seqlen * 0.98 + 0.01 * seqlen
    weights = np.zeros((R, N))
    anc = np.full((R, N), -1, dtype=np.int32)
    t_bounds = [(float(t_start[e]) + 1.0, float(t_start[e]) + 11.0) for e in range(E)]   # time slice inside epoch e
    x_prev = 0.0
    for r in range(R):
        x1 = row_end[r]
        for s in range(N):
            e = int(rng.integers(0, E))
            t0, t1 = t_bounds[e]
            # recombination opportunity over the whole row, k lineages, slice [t0,t1)
            k = int(rng.integers(2, 6))
            code = 2 if rng.random() < 0.3 else 1
            ops.append([code, s, e, t0, t1, x_prev, x1, k])
            evs.append((r, s, 1, e, 0, x_prev, x1, k * (t1 - t0)))
            if code == 2:
                evs.append((r, s, 2, e, 1, x_prev, x_prev, 0.0))
            if rng.random() < 0.5:
                e2 = int(rng.integers(0, E))
                a, b = t_bounds[e2]
                kk = int(rng.integers(1, 5))
                has = int(rng.random() < 0.6)
                x = x_prev + (x1 - x_prev) * rng.random()
                ops.append([0, s, e2, a, b, x, kk, has])
                evs.append((r, s, 0, e2, has, x, x, kk * (b - a)))
        w = rng.random(N) + 0.05
        w /= w.sum()
        weights[r] = w
        ops.append([4, x1, 0, 0, 0, 0, 0, 0])
        ops.extend(_pack(w))
        if rng.random() < 0.5 and r + 1 < R:
            a = np.sort(rng.integers(0, N, size=N)).astype(np.int32)
            anc[r] = a
            ops.append([3, 0, 0, 0, 0, 0, 0, 0])
            ops.extend(_pack(a.astype(np.float64)))
        x_prev = x1
    w_final = rng.random(N) + 0.05
    w_final /= w_final.sum()
    ops.append([4, seqlen, 1, 0, 0, 0, 0, 0])
    ops.extend(_pack(w_final))
    return (np.array(ops, dtype=np.float64), row_end, weights, anc, w_final, evs)


def _pack(vals):
    out = []
    vals = list(vals)
    for i in range(0, len(vals), 7):
        chunk = vals[i:i + 7]
        out.append([0.0] + [float(v) for v in chunk] + [0.0] * (7 - len(chunk)))
    return out


def rec_events_array(evs):
    arr = (_libs.OrcRecEvent * len(evs))()
    for i, (r, s, ty, e, cnt, x0, x1, opp) in enumerate(evs):
        arr[i].row, arr[i].slot = r, s
        arr[i].ev.type, arr[i].ev.epoch, arr[i].ev.cnt = ty, e, cnt
        arr[i].ev.x0, arr[i].ev.x1, arr[i].ev.opp = x0, x1, opp
    return arr
