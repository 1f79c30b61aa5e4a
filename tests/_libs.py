"""ctypes loaders used by the tests (checker side only):

  oracle()   -> oracle/liboracle.so   the plain-C restatement (built on demand with gcc)
  ref()      -> oracle/_ref/libsmcref.so  the UNMODIFIED reference compiled from /root/reference (None if not built
                and /root/reference is absent, e.g. never on the GPU box unless it travelled with the snapshot)
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
MAXN, MAXE = 64, 64
MAXNODES = 2 * MAXN - 1

c_dp = ctypes.POINTER(ctypes.c_double)
c_ip = ctypes.POINTER(ctypes.c_int)


def P(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class OrcModel(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int), ("E", ctypes.c_int), ("mu", ctypes.c_double), ("rho", ctypes.c_double),
                ("seqlen", ctypes.c_double), ("t_start", ctypes.c_double * MAXE), ("inv2N", ctypes.c_double * MAXE),
                ("vb_logfac", ctypes.c_double * MAXE), ("rec_mask", ctypes.c_uint8 * MAXE)]


class OrcParticle(ctypes.Structure):
    _fields_ = [("parent", ctypes.c_int * MAXNODES), ("height", ctypes.c_double * MAXNODES),
                ("logw", ctypes.c_double), ("updated_to", ctypes.c_double), ("cur_base", ctypes.c_double),
                ("next_base", ctypes.c_double)]


class OrcRng(ctypes.Structure):
    _fields_ = [("key", ctypes.c_uint32 * 2), ("ctr", ctypes.c_uint32 * 4), ("buf", ctypes.c_double * 2),
                ("have", ctypes.c_int)]


class OrcEvent(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int), ("epoch", ctypes.c_int), ("cnt", ctypes.c_int),
                ("x0", ctypes.c_double), ("x1", ctypes.c_double), ("opp", ctypes.c_double)]


class OrcEvents(ctypes.Structure):
    _fields_ = [("ev", ctypes.POINTER(OrcEvent)), ("count", ctypes.c_int), ("cap", ctypes.c_int),
                ("overflow", ctypes.c_int)]


class OrcPfConfig(ctypes.Structure):
    _fields_ = [("N", ctypes.c_int), ("seed", ctypes.c_uint64), ("chunk", ctypes.c_uint32),
                ("slot_offset", ctypes.c_uint32), ("start_pos", ctypes.c_double), ("ess_fraction", ctypes.c_double),
                ("ancestral_aware", ctypes.c_int), ("lags", ctypes.c_void_p)]


class OrcRecEvent(ctypes.Structure):
    _fields_ = [("row", ctypes.c_int), ("slot", ctypes.c_int), ("ev", OrcEvent)]


_oracle = None
_ref = None


def build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("smc_oracle.c", "smc_tally_oracle.c", "smc_filter_oracle.c",
                                                   "smc_oracle.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def oracle():
    global _oracle
    if _oracle is None:
        L = ctypes.CDLL(build_oracle())
        for name in ("orc_fastexp", "orc_exp_digamma", "orc_prune_likelihood", "orc_site_likelihood",
                     "orc_tree_length", "orc_branch_length_with_data", "orc_normalize", "orc_ess_partial_sums",
                     "orc_normalize_log", "orc_rng_uniform"):
            getattr(L, name).restype = ctypes.c_double
        L.orc_fastexp.argtypes = [ctypes.c_double]
        L.orc_exp_digamma.argtypes = [ctypes.c_double]
        L.orc_prune_likelihood.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_double, ctypes.c_int]
        L.orc_site_likelihood.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_double, ctypes.c_int, ctypes.c_int]
        L.orc_tree_length.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        L.orc_branch_length_with_data.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.orc_systematic_resampling.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                                ctypes.c_void_p]
        L.orc_normalize.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_normalize_log.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_ess_partial_sums.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.orc_rng_init.argtypes = [ctypes.POINTER(OrcRng), ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32,
                                   ctypes.c_uint32, ctypes.c_uint32]
        L.orc_rng_uniform.argtypes = [ctypes.POINTER(OrcRng)]
        L.orc_init_particle.argtypes = [ctypes.POINTER(OrcModel), ctypes.POINTER(OrcParticle), ctypes.POINTER(OrcRng),
                                        ctypes.c_double]
        L.orc_smc_move.argtypes = [ctypes.POINTER(OrcModel), ctypes.POINTER(OrcParticle), ctypes.POINTER(OrcRng),
                                   ctypes.c_int, ctypes.POINTER(OrcEvents), c_dp]
        L.orc_extend_row.argtypes = [ctypes.POINTER(OrcModel), ctypes.POINTER(OrcParticle), ctypes.POINTER(OrcRng),
                                     ctypes.c_double, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                     ctypes.POINTER(OrcEvents)]
        L.orc_redraw_next_base.argtypes = [ctypes.POINTER(OrcModel), ctypes.POINTER(OrcParticle),
                                           ctypes.POINTER(OrcRng)]
        L.orc_update_schedule.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double,
                                          ctypes.c_int, ctypes.c_void_p]
        L.orc_max_epoch_to_record.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_double]
        L.orc_lagged_tallies.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double,
                                         ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p]
        L.orc_particle_filter.argtypes = [ctypes.POINTER(OrcModel), ctypes.POINTER(OrcPfConfig), ctypes.c_int,
                                          ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_void_p, c_dp, ctypes.c_void_p, c_ip, ctypes.c_void_p,
                                          ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_long)]
        L.orc_set_row_opportunity_buffer.argtypes = [ctypes.c_void_p]
        L.orc_set_row_opportunity_buffer.restype = None
        L.orc_set_pooled_at_close_buffer.argtypes = [ctypes.c_void_p]
        L.orc_set_pooled_at_close_buffer.restype = None
        _oracle = L
    return _oracle


def oracle_filter(model, N, seed, chunk, start_pos, ess_fraction, lags, row_end, data, is_site, max_rec=None,
                  ancestral_aware=0, slot_offset=0):
    """Run the CPU restatement of the whole per-chunk filter.  Returns a dict with parent[N,2n-1], height, logw,
    next_base, lnZ, tallies[6,E], n_resample, rs_pos, rs_ess, last_anc, n_recomb."""
    L = oracle()
    n, E = model.n, model.E
    nn = 2 * n - 1
    R = len(row_end)
    lags = np.ascontiguousarray(lags, dtype=np.float64)
    cfg = OrcPfConfig(N, seed, chunk, slot_offset, start_pos, ess_fraction, ancestral_aware, lags.ctypes.data)
    parts = (OrcParticle * N)()
    row_end = np.ascontiguousarray(row_end, dtype=np.float64)
    data = np.ascontiguousarray(data, dtype=np.int32)
    is_site = np.ascontiguousarray(is_site, dtype=np.int32)
    mr = None if max_rec is None else np.ascontiguousarray(max_rec, dtype=np.int32)
    lnZ = ctypes.c_double()
    tallies = np.zeros(6 * E)
    nres = ctypes.c_int()
    nrec = ctypes.c_long()
    rs_pos, rs_ess = np.zeros(R), np.zeros(R)
    last_anc = np.zeros(N, dtype=np.int32)
    row_opp = np.zeros(R)
    L.orc_set_row_opportunity_buffer(P(row_opp))
    pooled = np.zeros(2 * E)
    L.orc_set_pooled_at_close_buffer(P(pooled))
    rc = L.orc_particle_filter(ctypes.byref(model), ctypes.byref(cfg), R, P(row_end), P(data), P(is_site),
                               None if mr is None else P(mr), parts, ctypes.byref(lnZ), P(tallies),
                               ctypes.byref(nres), P(rs_pos), P(rs_ess), P(last_anc), ctypes.byref(nrec))
    L.orc_set_row_opportunity_buffer(None)
    L.orc_set_pooled_at_close_buffer(None)
    if rc:
        raise RuntimeError("oracle filter failed (event buffer overflow or zero weight)")
    parent = np.zeros((N, nn), dtype=np.int32)
    height = np.zeros((N, nn))
    logw, nb = np.zeros(N), np.zeros(N)
    for i in range(N):
        parent[i] = np.ctypeslib.as_array(parts[i].parent)[:nn]
        height[i] = np.ctypeslib.as_array(parts[i].height)[:nn]
        logw[i], nb[i] = parts[i].logw, parts[i].next_base
    return dict(parent=parent, height=height, logw=logw, next_base=nb, lnZ=lnZ.value, tallies=tallies.reshape(6, E),
                n_resample=nres.value, rs_pos=rs_pos[:nres.value], rs_ess=rs_ess[:nres.value], last_anc=last_anc,
                n_recomb=nrec.value, row_opp=row_opp, w2_pooled_at_close=pooled.reshape(2, E))


def make_model(n, t_start, inv2N, mu, rho, seqlen, vb_logfac=None, rec_mask=None):
    m = OrcModel()
    m.n, m.E, m.mu, m.rho, m.seqlen = n, len(t_start), mu, rho, seqlen
    for e, (t, v) in enumerate(zip(t_start, inv2N)):
        m.t_start[e] = t
        m.inv2N[e] = v
        m.vb_logfac[e] = 0.0 if vb_logfac is None else vb_logfac[e]
        m.rec_mask[e] = 3 if rec_mask is None else rec_mask[e]
    return m


def ref():
    """The compiled reference (or None).  Built from /root/reference when that is present."""
    global _ref
    if _ref is None:
        so = os.path.join(ORACLE_DIR, "_ref", "libsmcref.so")
        if not os.path.exists(so) and os.path.isdir("/root/reference/src"):
            subprocess.check_call([os.path.join(ORACLE_DIR, "build_ref.sh")], stdout=subprocess.DEVNULL,
                                  stderr=subprocess.DEVNULL)
        if not os.path.exists(so):
            return None
        L = ctypes.CDLL(so)
        L.ref_open.restype = ctypes.c_void_p
        L.ref_last_error.restype = ctypes.c_char_p
        L.ref_mutation_rate.restype = ctypes.c_double
        L.ref_recombination_rate.restype = ctypes.c_double
        L.ref_max_seg_len.restype = ctypes.c_double
        for name in ("ref_close", "ref_num_epochs", "ref_mutation_rate", "ref_recombination_rate", "ref_max_seg_len"):
            getattr(L, name).argtypes = [ctypes.c_void_p]
        L.ref_epoch_table.argtypes = [ctypes.c_void_p] * 3
        L.ref_prune.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                ctypes.c_void_p, ctypes.c_int, c_dp]
        L.ref_site_likelihood.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                          ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, c_dp]
        L.ref_branch_lengths.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p, c_dp, c_dp]
        L.ref_extend_nomut.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_void_p, ctypes.c_double, ctypes.c_double, c_dp]
        L.ref_systematic.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                     ctypes.c_void_p]
        L.ref_weights_path.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_double, c_dp, c_ip, ctypes.c_void_p, ctypes.c_void_p]
        L.ref_prior_trees.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                      ctypes.c_void_p]
        L.ref_smc_moves.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                    ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.ref_arg_walk.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 4
        L.ref_tally_script.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_long, ctypes.c_void_p]
        if hasattr(L, "ref_spawn_probe"):
            L.ref_spawn_probe.argtypes = [ctypes.c_void_p, ctypes.c_int, c_ip, c_ip, c_ip, c_dp, c_dp]
        _ref = L
    return _ref


def ref_open(cmdline):
    """Open a reference handle from an smcsmc command line string (quiet)."""
    L = ref()
    args = cmdline.split()
    av = (ctypes.c_char_p * len(args))(*[a.encode() for a in args])
    # the reference prints its command line on stdout while parsing; silence fd 1 around the call
    import sys
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    try:
        h = L.ref_open(len(args), av)
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    if not h:
        raise RuntimeError("ref_open failed: %s" % L.ref_last_error())
    return ctypes.c_void_p(h)


def ref_spawn_probe(h, fill_byte):
    """oracle/ref_harness.cpp::ref_spawn_probe -> dict(field, epochs_recorded, epochs_extended, recorded, expected)"""
    L = ref()
    f, a, b = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    r, e = ctypes.c_double(), ctypes.c_double()
    if L.ref_spawn_probe(h, fill_byte, ctypes.byref(f), ctypes.byref(a), ctypes.byref(b), ctypes.byref(r), ctypes.byref(e)):
        raise RuntimeError("ref_spawn_probe failed: %s" % L.ref_last_error())
    return {"field": f.value, "epochs_recorded": a.value, "epochs_extended": b.value, "recorded": r.value,
            "expected": e.value}


def ref_model(h, n, seqlen):
    """orc model taken from the reference's own parsed Model (single population)."""
    L = ref()
    E = L.ref_num_epochs(h)
    t = np.zeros(E)
    inv = np.zeros(E)
    L.ref_epoch_table(h, P(t), P(inv))
    return make_model(n, list(t), list(inv), L.ref_mutation_rate(h), L.ref_recombination_rate(h), seqlen)


# ---- structured-population oracle (oracle/smc_struct_oracle.h) -----------------------------------------------------
ORC_MAXP, ORC_MAXMARKS = 4, 256


class OrcSModel(ctypes.Structure):
    _fields_ = [("base", OrcModel), ("P", ctypes.c_int),
                ("inv2N", (ctypes.c_double * ORC_MAXP) * 64),
                ("mig", ((ctypes.c_double * ORC_MAXP) * ORC_MAXP) * 64),
                ("single_mig", ((ctypes.c_double * ORC_MAXP) * ORC_MAXP) * 64),
                ("leaf_pop", ctypes.c_int * 64),
                ("vb_coal", (ctypes.c_double * ORC_MAXP) * 64),
                ("vb_mig", ((ctypes.c_double * ORC_MAXP) * ORC_MAXP) * 64)]


class OrcSParticle(ctypes.Structure):
    _fields_ = [("tree", OrcParticle), ("pop", ctypes.c_int * 127), ("nmarks", ctypes.c_int),
                ("mark_branch", ctypes.c_int * ORC_MAXMARKS), ("mark_time", ctypes.c_double * ORC_MAXMARKS),
                ("mark_pop", ctypes.c_int * ORC_MAXMARKS), ("overflow", ctypes.c_int)]


def make_smodel(n, t_start, inv2N, mig, single, leaf_pop, mu, rho, seqlen, vb_coal=None, vb_mig=None):
    """inv2N [E][P], mig / single [E][P][P] (backward rates per generation / fixed-time move probabilities);
    vb_coal [E][P] / vb_mig [E][P][P]: -vb log factors (default none)"""
    import numpy as np
    inv2N, mig, single = np.asarray(inv2N), np.asarray(mig), np.asarray(single)
    E, Pn = inv2N.shape
    m = OrcSModel()
    m.base.n, m.base.E, m.base.mu, m.base.rho, m.base.seqlen = n, E, mu, rho, seqlen
    m.P = Pn
    for e in range(E):
        m.base.t_start[e] = t_start[e]
        m.base.inv2N[e] = inv2N[e][0]
        m.base.rec_mask[e] = 3
        for i in range(Pn):
            m.inv2N[e][i] = inv2N[e][i]
            for j in range(Pn):
                m.mig[e][i][j] = mig[e][i][j]
                m.single_mig[e][i][j] = single[e][i][j]
                if vb_mig is not None:
                    m.vb_mig[e][i][j] = vb_mig[e][i][j]
            if vb_coal is not None:
                m.vb_coal[e][i] = vb_coal[e][i]
    for i in range(n):
        m.leaf_pop[i] = int(leaf_pop[i])
    return m


def oracle_sfilter(smodel, N, seed, chunk, start_pos, ess_fraction, lags, row_end, data, is_site, max_rec=None,
                   ancestral_aware=0, slot_offset=0):
    """The CPU restatement of the whole filter with populations (oracle/smc_struct_oracle.c: orc_sparticle_filter).
    Returns parent, height, pop [N,2n-1], marks (per particle: sorted list of (branch, time, pop)), logw, next_base,
    lnZ, tallies [2,Q,E], n_resample, last_anc, n_recomb."""
    L = oracle()
    L.orc_schannels.restype = ctypes.c_int
    n, E = smodel.base.n, smodel.base.E
    nn = 2 * n - 1
    Q = L.orc_schannels(smodel.P)
    R = len(row_end)
    lags = np.ascontiguousarray(lags, dtype=np.float64)
    parts = (OrcSParticle * N)()
    row_end = np.ascontiguousarray(row_end, dtype=np.float64)
    data = np.ascontiguousarray(data, dtype=np.int32)
    is_site = np.ascontiguousarray(is_site, dtype=np.int32)
    mr = None if max_rec is None else np.ascontiguousarray(max_rec, dtype=np.int32)
    lnZ = ctypes.c_double()
    tallies = np.zeros(2 * Q * E)
    nres = ctypes.c_int()
    nrec = ctypes.c_long()
    last_anc = np.zeros(N, dtype=np.int32)
    pooled = np.zeros(Q * E)
    L.orc_set_pooled_at_close_buffer(P(pooled))
    rc = L.orc_sparticle_filter(ctypes.byref(smodel), N, ctypes.c_uint64(seed), ctypes.c_uint32(chunk),
                                ctypes.c_uint32(slot_offset), ctypes.c_double(start_pos), ctypes.c_double(ess_fraction),
                                ancestral_aware, P(lags), R, P(row_end), P(data), P(is_site),
                                None if mr is None else P(mr), parts, ctypes.byref(lnZ), P(tallies), ctypes.byref(nres),
                                P(last_anc), ctypes.byref(nrec))
    L.orc_set_pooled_at_close_buffer(None)
    if rc:
        raise RuntimeError("structured oracle filter failed: %d" % rc)
    parent = np.zeros((N, nn), dtype=np.int32)
    height = np.zeros((N, nn))
    pop = np.zeros((N, nn), dtype=np.int32)
    logw, nb = np.zeros(N), np.zeros(N)
    marks = []
    for i in range(N):
        parent[i] = np.ctypeslib.as_array(parts[i].tree.parent)[:nn]
        height[i] = np.ctypeslib.as_array(parts[i].tree.height)[:nn]
        pop[i] = np.ctypeslib.as_array(parts[i].pop)[:nn]
        logw[i], nb[i] = parts[i].tree.logw, parts[i].tree.next_base
        k = parts[i].nmarks
        marks.append(sorted((int(parts[i].mark_branch[q]), float(parts[i].mark_time[q]), int(parts[i].mark_pop[q]))
                            for q in range(k)))
    return dict(parent=parent, height=height, pop=pop, marks=marks, logw=logw, next_base=nb, lnZ=lnZ.value,
                tallies=tallies.reshape(2, Q, E), n_resample=nres.value, last_anc=last_anc, n_recomb=nrec.value,
                w2_pooled_at_close=pooled.reshape(Q, E))
