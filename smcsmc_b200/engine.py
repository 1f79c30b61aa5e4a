"""Python mirror of the reference's ParticleContainer surface for the hot path, over the C ABI (host buffers in/out).

ParticleFilter plays the role of ParticleContainer + CountModel inside pfARG_core (reference src/smcsmc.cpp:278-400):
construct (prior trees), update over rows, read ln_normalization_factor and the lagged tallies.
"""
import ctypes

import numpy as np

from ._abi import lib, check, SmcConfig

_P = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def shard_unique_id():
    """128-byte NCCL id for one sharded population (call on rank 0, pass to every rank's ParticleFilter)"""
    buf = np.zeros(128, dtype=np.uint8)
    check(lib().smc_shard_unique_id(_P(buf)))
    return buf.tobytes()


class ParticleFilter:
    def __init__(self, n, N, t_start, inv2N, mu, rho, seqlen, lags, seed=1, chunk=0, start_pos=0.0, ess_fraction=0.5,
                 ancestral_aware=False, parity_scan=False, vb_logfac=None, rec_mask=None, device=0, slot_offset=0,
                 shard=None, populations=None):
        """N is the number of particles THIS engine holds.  populations = dict(inv2N[E][P], mig[E][P][P] or None,
        single_mig[E][P][P] or None, leaf_pop[n], max_marks=0) switches the structured transition on
        (smc_set_populations; scrm -I / -ej / -em); the argument inv2N is then ignored.  shard = (rank, world, unique_id) splits a population of
        world * N particles over `world` engines, one per process and GPU (smc_shard_attach); unique_id comes from
        shard_unique_id() on rank 0 and is handed to the other ranks by the caller (smcsmc_b200.sharding)."""
        self.n, self.N, self.E = int(n), int(N), len(t_start)
        self._keep = [_f64(t_start), _f64(inv2N), _f64(lags),
                      None if vb_logfac is None else _f64(vb_logfac),
                      None if rec_mask is None else np.ascontiguousarray(rec_mask, dtype=np.uint8)]
        c = SmcConfig()
        c.n, c.N, c.E, c.device = self.n, self.N, self.E, device
        c.seed, c.chunk, c.slot_offset = seed, chunk, slot_offset
        c.mu, c.rho, c.seqlen, c.start_pos, c.ess_fraction = mu, rho, seqlen, start_pos, ess_fraction
        c.ancestral_aware, c.parity_scan = int(ancestral_aware), int(parity_scan)
        c.t_start, c.inv2N, c.lags = (self._keep[0].ctypes.data, self._keep[1].ctypes.data, self._keep[2].ctypes.data)
        c.vb_logfac = None if self._keep[3] is None else self._keep[3].ctypes.data
        c.rec_mask = None if self._keep[4] is None else self._keep[4].ctypes.data
        self._h = ctypes.c_void_p()
        check(lib().smc_create(ctypes.byref(c), ctypes.byref(self._h)))
        self.R = 0
        self.shard = None
        self.P, self.max_marks = 1, 0
        if populations is not None:
            self.set_populations(**populations)
        if shard is not None:
            rank, world, uid = shard
            uid = np.ascontiguousarray(np.frombuffer(bytes(uid), dtype=np.uint8))
            assert uid.size == 128
            check(lib().smc_shard_attach(self._h, int(rank), int(world), _P(uid)))
            self.shard = (int(rank), int(world))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().smc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_populations(self, inv2N, leaf_pop, mig=None, single_mig=None, max_marks=0):
        """Model with several populations; call again after reset() to replace the rates (M-step)"""
        inv = _f64(inv2N)
        E, P = inv.shape
        assert E == self.E
        m = None if mig is None else _f64(mig).reshape(E, P, P)
        sm = None if single_mig is None else _f64(single_mig).reshape(E, P, P)
        lp = np.ascontiguousarray(leaf_pop, dtype=np.int32)
        assert lp.shape == (self.n,)
        check(lib().smc_set_populations(self._h, P, _P(inv), None if m is None else _P(m),
                                        None if sm is None else _P(sm), _P(lp), int(max_marks)))
        self.P = P
        self.max_marks = int(max_marks) if max_marks > 0 else 2 * self.n + 8

    def set_population_vb(self, coal_logfac=None, mig_logfac=None):
        """-vb with populations: log(exp_digamma(c)/c) per coalescence [E][P] and per migration [E][P][P]"""
        c = None if coal_logfac is None else _f64(coal_logfac).reshape(self.E, self.P)
        m = None if mig_logfac is None else _f64(mig_logfac).reshape(self.E, self.P, self.P)
        check(lib().smc_set_population_vb(self._h, None if c is None else _P(c), None if m is None else _P(m)))

    def population_state(self):
        """(pop[N][2n-1], marks) with marks[i] = list of (branch, time, population above) of particle i"""
        nn, K = 2 * self.n - 1, self.max_marks
        pop = np.zeros((self.N, nn), dtype=np.int32)
        nm = np.zeros(self.N, dtype=np.int32)
        mb, mp = np.zeros((self.N, K), dtype=np.int32), np.zeros((self.N, K), dtype=np.int32)
        mt = np.zeros((self.N, K))
        check(lib().smc_get_population_state(self._h, _P(pop), _P(nm), _P(mb), _P(mt), _P(mp)))
        marks = [[(int(mb[i, k]), float(mt[i, k]), int(mp[i, k])) for k in range(nm[i])] for i in range(self.N)]
        return pop, marks

    # --- Segment rows ---
    def upload_rows(self, x_end, alleles, is_site, max_rec_epoch=None):
        x_end = _f64(x_end)
        alleles = np.ascontiguousarray(alleles, dtype=np.int8)
        is_site = np.ascontiguousarray(is_site, dtype=np.uint8)
        mr = None if max_rec_epoch is None else np.ascontiguousarray(max_rec_epoch, dtype=np.int32)
        assert alleles.shape == (len(x_end), self.n)
        check(lib().smc_upload_rows(self._h, _P(x_end), _P(alleles), _P(is_site), None if mr is None else _P(mr),
                                    len(x_end)))
        self.R = len(x_end)

    def init_particles(self):
        check(lib().smc_init_particles(self._h))

    def set_particles(self, parent, height, logw, next_base):
        parent = np.ascontiguousarray(parent, dtype=np.int32)
        check(lib().smc_set_particles(self._h, _P(parent), _P(_f64(height)), _P(_f64(logw)), _P(_f64(next_base))))

    # --- the hot loop ---
    def sweep(self, row_begin=0, row_end=None):
        check(lib().smc_sweep(self._h, row_begin, self.R if row_end is None else row_end))

    def finish(self):
        check(lib().smc_finish(self._h))

    def reset(self, inv2N=None, rho=0.0, vb_logfac=None, seed=1):
        """next EM iteration on the same rows, optionally with the rates the M-step produced"""
        a = None if inv2N is None else _f64(inv2N)
        b = None if vb_logfac is None else _f64(vb_logfac)
        check(lib().smc_reset(self._h, None if a is None else _P(a), rho, None if b is None else _P(b), seed))

    # --- results ---
    def particles(self):
        nn = 2 * self.n - 1
        parent = np.zeros((self.N, nn), dtype=np.int32)
        height = np.zeros((self.N, nn))
        logw, nb = np.zeros(self.N), np.zeros(self.N)
        check(lib().smc_get_particles(self._h, _P(parent), _P(height), _P(logw), _P(nb)))
        return parent, height, logw, nb

    def ln_normalization_factor(self):
        v = ctypes.c_double()
        check(lib().smc_get_lnZ(self._h, ctypes.byref(v)))
        return v.value

    def tallies(self):
        """[6][E] (coal count, opp, w^2-opp, rec count, opp, w^2-opp); with populations [2][Q][E] (see the header)"""
        rows = lib().smc_tally_rows(self._h)
        out = np.zeros((rows, self.E))
        check(lib().smc_get_tallies(self._h, _P(out)))
        return out if self.P == 1 else out.reshape(2, rows // 2, self.E)

    def counters(self):
        c = np.zeros(4, dtype=np.int64)
        check(lib().smc_get_counters(self._h, _P(c)))
        return dict(rows=int(c[0]), resamples=int(c[1]), recombinations=int(c[2]), err=int(c[3]))

    def row_opportunity(self):
        """per row: sum_i w_i * (local tree length x bp over the row), the local recombination map's opportunity"""
        out = np.zeros(self.R)
        check(lib().smc_get_row_opportunity(self._h, _P(out)))
        return out

    def resample_log(self):
        pos, ess = np.zeros(max(self.R, 1)), np.zeros(max(self.R, 1))
        k = ctypes.c_int64()
        check(lib().smc_get_resample_log(self._h, _P(pos), _P(ess), len(pos), ctypes.byref(k)))
        return pos[:k.value], ess[:k.value]

    def last_ancestry(self):
        a = np.zeros(self.N, dtype=np.int32)
        check(lib().smc_get_last_ancestry(self._h, _P(a)))
        return a

    def timing(self):
        ms, k = ctypes.c_double(), ctypes.c_int64()
        check(lib().smc_get_timing(self._h, ctypes.byref(ms), ctypes.byref(k)))
        return ms.value, k.value


def sweep_many(filters, rows_per_turn=8):
    """sweep all remaining rows of several ParticleFilters (chunks) concurrently on the device"""
    arr = (ctypes.c_void_p * len(filters))(*[f._h for f in filters])
    check(lib().smc_sweep_many(arr, len(filters), rows_per_turn))


# ---- stand-alone functions (each mirrors one reference member function) ----
def site_likelihood(parent, height, alleles, mu, ancestral_aware=False, device=0):
    """ForestState::calculate_likelihood averaged over phasings (particle.cpp:667, particleContainer.cpp:204-221)."""
    parent = np.ascontiguousarray(parent, dtype=np.int32)
    height = _f64(height)
    alleles = np.ascontiguousarray(alleles, dtype=np.int8)
    count, n = alleles.shape
    out = np.zeros(count)
    check(lib().smc_site_likelihood(device, n, count, _P(parent), _P(height), _P(alleles), mu, int(ancestral_aware),
                                    _P(out)))
    return out


def branch_lengths(parent, height, alleles, device=0):
    """(trackLocalTreeBranchLength, getLocalTreeLength)  (particle.cpp:699, forest.h:188)."""
    parent = np.ascontiguousarray(parent, dtype=np.int32)
    height = _f64(height)
    alleles = np.ascontiguousarray(alleles, dtype=np.int8)
    count, n = alleles.shape
    a, b = np.zeros(count), np.zeros(count)
    check(lib().smc_branch_lengths(device, n, count, _P(parent), _P(height), _P(alleles), _P(a), _P(b)))
    return a, b


def systematic_offspring(partial_sum, u, device=0):
    """ParticleContainer::systematic_resampling (particleContainer.cpp:474)."""
    ps = _f64(partial_sum)
    N = len(ps) - 1
    counts = np.zeros(N, dtype=np.int32)
    check(lib().smc_systematic_offspring(device, _P(ps), N, u, _P(counts)))
    return counts


def resample_counts(w, u, parity_scan=True, device=0):
    w = _f64(w)
    counts = np.zeros(len(w), dtype=np.int32)
    ess = ctypes.c_double()
    check(lib().smc_resample_counts(device, _P(w), len(w), u, int(parity_scan), ctypes.byref(ess), _P(counts)))
    return ess.value, counts


def normalize_log(logw, device=0):
    lw = _f64(logw).copy()
    lnZ, ess = ctypes.c_double(), ctypes.c_double()
    check(lib().smc_normalize_log(device, _P(lw), len(lw), ctypes.byref(lnZ), ctypes.byref(ess)))
    return lw, lnZ.value, ess.value


def fastexp(x, device=0):
    x = _f64(x)
    out = np.zeros_like(x)
    check(lib().smc_fastexp(device, _P(x), x.size, _P(out)))
    return out


def calibrate_survival(n, t_start, inv2N, rho, seqlen, seed=1, n_args=8192, device=0):
    """calculate_median_survival_distances (smcsmc.cpp:169): per-epoch median node survival distance and sample counts"""
    ts, iv = _f64(t_start), _f64(inv2N)
    E = len(ts)
    med = np.zeros(E)
    cnt = np.zeros(E, dtype=np.int64)
    check(lib().smc_calibrate_survival(device, n, E, _P(ts), _P(iv), rho, seqlen, seed, n_args, _P(med), _P(cnt)))
    return med, cnt


def calibrate_survival_struct(n, t_start, inv2N, mig, single_mig, leaf_pop, rho, seqlen, seed=1, n_args=8192, device=0):
    """the same under a structured model: inv2N[E][P], mig / single_mig [E][P][P] (or None), leaf_pop[n]"""
    ts, iv = _f64(t_start), _f64(inv2N)
    E, npop = iv.shape
    mg = None if mig is None else _f64(mig)
    sg = None if single_mig is None else _f64(single_mig)
    lp = np.ascontiguousarray(leaf_pop, dtype=np.int32)
    med = np.zeros(E)
    cnt = np.zeros(E, dtype=np.int64)
    check(lib().smc_calibrate_survival_struct(device, n, E, _P(ts), npop, _P(iv), None if mg is None else _P(mg),
                                              None if sg is None else _P(sg), _P(lp), rho, seqlen, seed, n_args,
                                              _P(med), _P(cnt)))
    return med, cnt


def measure_fp64_fma(device=0):
    """sustained fp64 FMA rate of the device in TFLOP/s (register-only probe kernel)"""
    v = ctypes.c_double()
    check(lib().smc_measure_fp64_fma(device, ctypes.byref(v)))
    return v.value


def terminal_branch_quantiles(n, t_start, inv2N, quantiles=(0.001, 0.003, 0.01, 0.03, 0.1, 0.5, 0.95), n_trees=1000000, seed=1,
                              mig=None, single_mig=None, leaf_pop=None, device=0):
    """calculate_terminal_branch_length_quantiles (smcsmc.cpp:128): (lengths[n][len(quantiles)], mean tree length) from
    n_trees prior trees drawn on the device; inv2N[E] for one population, [E][P] with leaf_pop for several"""
    ts, iv, q = _f64(t_start), _f64(inv2N), _f64(quantiles)
    E = len(ts)
    npop = 1 if iv.ndim == 1 else iv.shape[1]
    out = np.zeros((n, len(q)))
    mean = ctypes.c_double()
    mg = None if mig is None else _f64(mig)
    sg = None if single_mig is None else _f64(single_mig)
    lp = None if leaf_pop is None else np.ascontiguousarray(leaf_pop, dtype=np.int32)
    check(lib().smc_terminal_branch_quantiles(device, n, E, _P(ts), npop, _P(iv), None if mg is None else _P(mg),
                                              None if sg is None else _P(sg), None if lp is None else _P(lp), seed,
                                              n_trees, len(q), _P(q), _P(out), ctypes.byref(mean)))
    return out, mean.value

