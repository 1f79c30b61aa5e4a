"""Worker of tests/test_sharded.py -- run under `python -m torch.distributed.run --nproc-per-node G` on a box with
G GPUs.  Every rank holds N/G particles of ONE population (smc_shard_attach); rank 0 also sweeps the same population
unsharded on its own GPU.  Philox streams are keyed by the global particle slot, so the two runs must agree: trees,
ancestry and resampling log exactly, weights / ln normaliser / tallies to the order of the floating-point sums.
Prints one JSON line from rank 0 and exits non-zero on any mismatch."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import smcsmc_b200 as S                                    # noqa: E402
from smcsmc_b200.sharding import broadcast_unique_id       # noqa: E402


def rows(n, R, seqlen, seed):
    rng = np.random.default_rng(seed)
    x_end = np.sort(rng.choice(np.arange(50, int(seqlen)), size=R - 1, replace=False)).astype(np.float64)
    x_end = np.append(x_end, float(seqlen))
    al = (rng.random((R, n)) < 0.3).astype(np.int8)
    al[R - 1] = -1
    return x_end, al, np.ones(R, dtype=np.uint8)


CASES = {
    # name: n, N_total, R, seqlen, t_start, inv2N, lags, mu, rho
    "n4_small": (4, 256, 120, 120000.0, [0.0, 4000.0, 20000.0], [1 / 2e4, 1 / 8e3, 1 / 3e4], [9000.0, 2500.0, 800.0],
                 2.5e-8, 1e-8),
    "zigzag11_n8": (8, 4096, 150, 150000.0, [4e4 * t for t in (0, .0025, .005, .01, .02, .04, .08, .16, .32, .64, 1.28)],
                    [1 / (2e4 * s) for s in (1, .2, 1, .2, 1, .2, 1, .2, 1, .2, 1)],
                    [20000.0, 15000.0, 9000.0, 7000.0, 5000.0, 3000.0, 2000.0, 1500.0, 1000.0, 600.0, 300.0], 1e-7, 4e-8),
}
# the first case once more: the new engine finds its slabs in the block pool and the peers' mappings in the cache, the
# communicator is the one of the first engine (same id), and it sweeps twice (smc_reset in between)
CASES["n4_small_again"] = CASES["n4_small"]
# two populations of 4 haplotypes, migration 2 -> 1 backwards in time, joined in the last epoch (BASELINE configs[2] in
# small): the populations and migration marks of a parent on another rank travel over NVLink as well
CASES["struct_2pop"] = (8, 2048, 100, 100000.0, [0.0, 4000.0, 20000.0], [1 / 2e4] * 3, [9000.0, 3000.0, 1000.0], 5e-8, 2e-8)


def struct_model(E=3, P=2):
    inv = np.full((E, P), 1 / 2e4)
    mig = np.zeros((E, P, P))
    mig[:2, 1, 0] = 2.5e-5
    single = np.zeros((E, P, P))
    single[2, 1, 0] = 1.0
    return dict(inv2N=inv, mig=mig, single_mig=single, leaf_pop=[0, 0, 0, 0, 1, 1, 1, 1], max_marks=24)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    report = {}
    ok = True
    for name, (n, Ntot, R, seqlen, t_start, inv2N, lags, mu, rho) in CASES.items():
        x_end, al, is_site = rows(n, R, seqlen, seed=len(name.replace("_again", "")))
        uid = broadcast_unique_id(device="cuda")
        N = Ntot // world
        pops = struct_model() if name.startswith("struct") else None
        pf = S.ParticleFilter(n, N, t_start, inv2N, mu, rho, seqlen, lags, seed=77, chunk=5, device=local,
                              shard=(rank, world, uid), populations=pops)
        pf.upload_rows(x_end, al, is_site)
        pf.init_particles()
        pf.sweep(0, R // 2)
        pf.sweep(R // 2, R)
        pf.finish()
        if name.endswith("_again"):
            # a second sweep on the same engine (the next EM iteration): smc_reset starts a new mailbox generation
            pf.reset(seed=78)
            pf.init_particles()
            pf.sweep()
            pf.finish()
        parent, height, logw, nb = pf.particles()
        mine = dict(parent=parent, height=height, logw=logw, nb=nb, lnZ=pf.ln_normalization_factor(),
                    tallies=pf.tallies(), counters=pf.counters(), rs=pf.resample_log(), anc=pf.last_ancestry(),
                    row_opp=pf.row_opportunity())
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        pf.close()
        if rank != 0:
            continue
        # a sharded population keeps per-particle squares in the *_weight rows (include/smcsmc_b200.h, smc_get_tallies):
        # the single-GPU engine it is compared with is created with the same rule (the knob is read by smc_create)
        os.environ["SMC_NO_POOL"] = "1"
        one = S.ParticleFilter(n, Ntot, t_start, inv2N, mu, rho, seqlen, lags, seed=77, chunk=5, device=local,
                               populations=pops)
        del os.environ["SMC_NO_POOL"]
        one.upload_rows(x_end, al, is_site)
        one.init_particles()
        one.sweep()
        one.finish()
        if name.endswith("_again"):
            one.reset(seed=78)
            one.init_particles()
            one.sweep()
            one.finish()
        p1, h1, w1, b1 = one.particles()
        res = {}
        try:
            for g in gathered:     # every rank ends with the global scalars
                assert g["counters"]["resamples"] == one.counters()["resamples"] > 0, (g["counters"], one.counters())
                assert g["counters"]["recombinations"] == one.counters()["recombinations"]
                assert g["counters"]["err"] == 0
                assert abs(g["lnZ"] - one.ln_normalization_factor()) <= 1e-9 * abs(one.ln_normalization_factor())
                np.testing.assert_allclose(g["tallies"], one.tallies(), rtol=1e-8, atol=1e-12)
                np.testing.assert_allclose(g["row_opp"], one.row_opportunity(), rtol=1e-9, atol=0)
                assert np.array_equal(g["rs"][0], one.resample_log()[0])
                np.testing.assert_allclose(g["rs"][1], one.resample_log()[1], rtol=1e-9)
            assert np.array_equal(np.concatenate([g["parent"] for g in gathered]), p1)
            assert np.array_equal(np.concatenate([g["anc"] for g in gathered]), one.last_ancestry())
            np.testing.assert_allclose(np.concatenate([g["height"] for g in gathered]), h1, rtol=1e-12, atol=0)
            np.testing.assert_allclose(np.concatenate([g["nb"] for g in gathered]), b1, rtol=1e-12, atol=0)
            np.testing.assert_allclose(np.concatenate([g["logw"] for g in gathered]), w1, rtol=0, atol=1e-9)
            anc = one.last_ancestry()
            res = {"ok": True, "resamples": one.counters()["resamples"],
                   "cross_rank_parents": int(np.sum(anc // N != np.arange(Ntot) // N))}
        except AssertionError as ex:
            ok = False
            res = {"ok": False, "error": str(ex)[:2000]}
        one.close()
        report[name] = res
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, src=0)
    if rank == 0:
        print(json.dumps({"world": world, "cases": report}), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
