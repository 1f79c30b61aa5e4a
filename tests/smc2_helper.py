"""Run the reference's UNMODIFIED Python frontend (smc2 / smcsmc.run_smcsmc, reference smcsmc/utils.py:184-217) against a
given `smcsmc` executable.  Test infrastructure: the package is the pip install of /root/reference into baseline/_ref
(__graft_entry__.build(); git-ignored, travels with gpurun).  Only plotting / tree-sequence libraries the run does not
touch are stubbed (matplotlib, tskit, msprime, stdpopsim): nothing of the frontend's own code is replaced.

usage: python smc2_helper.py <smcsmc executable> <output dir> <seg file> [Np]"""
import importlib
import os
import sys
from unittest import mock

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FRONTEND = os.path.join(ROOT, "baseline", "_ref")


def available():
    return os.path.isdir(os.path.join(FRONTEND, "smcsmc"))


def main(binp, out, seg, Np="500"):
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.collections", "matplotlib.colors", "matplotlib.cm", "tskit",
                 "msprime", "stdpopsim", "seaborn"):
        try:
            importlib.import_module(name)
        except Exception:
            sys.modules[name] = mock.MagicMock(name=name)
    sys.path.insert(0, FRONTEND)
    import smcsmc
    # two chunks, two EM iterations, nine log-spaced epochs, population sizes only (the frontend's -no_infer_recomb
    # passes -xr), chunks run one after the other in this process (-nothreads)
    args = {"seg": seg, "nsam": "4", "Np": str(Np), "EM": "1", "chunks": "2", "o": out, "smcsmcpath": binp,
            "mu": "2.5e-8", "rho": "1e-8", "N0": "10000", "tmax": "3.5", "P": "133 133000 7*1", "no_infer_recomb": "",
            "seed": "3", "nothreads": ""}
    smcsmc.run_smcsmc(args)


if __name__ == "__main__":
    main(*sys.argv[1:])
