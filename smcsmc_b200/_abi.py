"""ctypes declarations for include/smcsmc_b200.h (libsmcgpu.so)."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_dp = ctypes.POINTER(ctypes.c_double)


class SmcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libsmcgpu error %d: %s" % (code, msg))
        self.code = code


class SmcConfig(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int32), ("N", ctypes.c_int32), ("E", ctypes.c_int32), ("device", ctypes.c_int32),
                ("seed", ctypes.c_uint64), ("chunk", ctypes.c_uint32), ("slot_offset", ctypes.c_uint32),
                ("mu", ctypes.c_double), ("rho", ctypes.c_double), ("seqlen", ctypes.c_double),
                ("start_pos", ctypes.c_double), ("ess_fraction", ctypes.c_double),
                ("ancestral_aware", ctypes.c_int32), ("parity_scan", ctypes.c_int32),
                ("t_start", ctypes.c_void_p), ("inv2N", ctypes.c_void_p), ("vb_logfac", ctypes.c_void_p),
                ("rec_mask", ctypes.c_void_p), ("lags", ctypes.c_void_p)]


# every symbol include/smcsmc_b200.h declares: name -> (restype, argtypes)
_V, _I, _I64, _D = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double
SYMBOLS = {
    "smc_create": (_I, [ctypes.POINTER(SmcConfig), ctypes.POINTER(_V)]),
    "smc_destroy": (None, [_V]),
    "smc_last_error": (ctypes.c_char_p, []),
    "smc_version": (ctypes.c_char_p, []),
    "smc_upload_rows": (_I, [_V, _V, _V, _V, _V, _I64]),
    "smc_init_particles": (_I, [_V]),
    "smc_sweep": (_I, [_V, _I64, _I64]),
    "smc_finish": (_I, [_V]),
    "smc_sweep_many": (_I, [_V, ctypes.c_int32, _I64]),
    "smc_shard_unique_id": (_I, [_V]),
    "smc_shard_attach": (_I, [_V, ctypes.c_int32, ctypes.c_int32, _V]),
    "smc_set_populations": (_I, [_V, ctypes.c_int32, _V, _V, _V, _V, ctypes.c_int32]),
    "smc_set_population_vb": (_I, [_V, _V, _V]),
    "smc_tally_rows": (_I, [_V]),
    "smc_get_population_state": (_I, [_V, _V, _V, _V, _V, _V]),
    "smc_reset": (_I, [_V, _V, _D, _V, ctypes.c_uint64]),
    "smc_get_tallies": (_I, [_V, _V]),
    "smc_get_lnZ": (_I, [_V, c_dp]),
    "smc_get_counters": (_I, [_V, _V]),
    "smc_get_resample_log": (_I, [_V, _V, _V, _I64, ctypes.POINTER(_I64)]),
    "smc_get_particles": (_I, [_V, _V, _V, _V, _V]),
    "smc_set_particles": (_I, [_V, _V, _V, _V, _V]),
    "smc_get_last_ancestry": (_I, [_V, _V]),
    "smc_get_timing": (_I, [_V, c_dp, ctypes.POINTER(_I64)]),
    "smc_site_likelihood": (_I, [_I, ctypes.c_int32, _I64, _V, _V, _V, _D, ctypes.c_int32, _V]),
    "smc_branch_lengths": (_I, [_I, ctypes.c_int32, _I64, _V, _V, _V, _V, _V]),
    "smc_systematic_offspring": (_I, [_I, _V, _I64, _D, _V]),
    "smc_resample_counts": (_I, [_I, _V, _I64, _D, ctypes.c_int32, c_dp, _V]),
    "smc_get_row_opportunity": (_I, [_V, _V]),
    "smc_measure_fp64_fma": (_I, [_I, _V]),
    "smc_terminal_branch_quantiles": (_I, [_I, ctypes.c_int32, ctypes.c_int32, _V, ctypes.c_int32, _V, _V, _V, _V,
                                            ctypes.c_uint64, _I64, ctypes.c_int32, _V, _V, c_dp]),
    "smc_resample_counts_noalloc": (_I, [_I, _V, _I64, _D, ctypes.c_int32, _V, _V, _V]),
    "smc_normalize_log": (_I, [_I, _V, _I64, c_dp, c_dp]),
    "smc_fastexp": (_I, [_I, _V, _I64, _V]),
    "smc_calibrate_survival": (_I, [_I, ctypes.c_int32, ctypes.c_int32, _V, _V, _D, _D, ctypes.c_uint64, _I64, _V, _V]),
    "smc_calibrate_survival_struct": (_I, [_I, ctypes.c_int32, ctypes.c_int32, _V, ctypes.c_int32, _V, _V, _V, _V, _D, _D,
                                           ctypes.c_uint64, _I64, _V, _V]),
}


def lib_path():
    # SMCSMC_B200_LIB: another build of the same library (kernel experiments); never a fallback
    return os.environ.get("SMCSMC_B200_LIB") or os.path.join(HERE, "lib", "libsmcgpu.so")


def lib():
    """Load libsmcgpu.so.  No fallback: a missing library is an error (build it with __graft_entry__.build())."""
    global _LIB
    if _LIB is None:
        path = lib_path()
        if not os.path.exists(path):
            raise SmcError(-1, "%s not built -- run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)" % path)
        L = ctypes.CDLL(path)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != 0:
        raise SmcError(rc, lib().smc_last_error().decode())
