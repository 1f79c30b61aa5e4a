#!/usr/bin/env python
"""bench.py -- particle*sites/s of the SMCSMC particle-filter sweep on B200 (BASELINE.json metric), one JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our engine (libsmcgpu.so through the C ABI)
  python bench.py --impl reference [...]                          the reference's own CPU implementation (oracle/_ref)

Workload (BASELINE.json configs[1]): 8 haplotypes x 100 Mb scrm-simulated under the zigzag
piecewise-constant demography (bench_data/zigzag_n8_100Mb_seed1.seg.gz, made by tools/make_seg.py with the reference's
scrm), inferred under 11 epochs at the zigzag change times starting from all-ones sizes, 10 000 particles.  A "step" is
one EM iteration's sweep over the whole chunk: prior trees, every .seg row (extend + prune + normalise + tally +
resample), end-of-data flush.  With --gpus N every rank sweeps its own 100 Mb chunk (the reference's own decomposition:
independent chunks, tallies summed; weak scaling), the tallies are all-reduced once per step over NCCL.

  value  particle*sites/s with the rows already resident in HBM (engine re-used, smc_reset between steps)
  e2e    the same through the public C-ABI calls with HOST buffers, everything the drop-in driver does per EM iteration:
         lag calibration on the device (smc_calibrate_survival -- the reference's step pays its own calibration too,
         smcsmc.cpp:287-288), max_epoch_to_update per row from those lags, create + upload rows (H2D from pinned
         memory) + sweep + finish + read tallies/lnZ (D2H), engines destroyed afterwards

Extra records on the same line (not the headline): `target_shape` (ONE population of 10^5 particles x 16 Mb, the
north-star shape per GPU), `struct` (configs[2] at full size through the drop-in driver binary) at --gpus 1, and `shard`
(ONE population of 10^6 particles x 4 Mb split over the ranks, strong scaling) at --gpus N > 1.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N0, MU, RHO = 1e4, 2.5e-8, 1e-8
ZIGZAG_TIMES = [0, .0025, .005, .01, .02, .04, .08, .16, .32, .64, 1.28]       # units of 4*N0 generations
# CountModel lags for this model as calibrated by the reference itself (calculate_median_survival_distances,
# smcsmc.cpp:169-263, lag_fraction 2; run of oracle/_ref/smcsmc recorded in DESIGN.md).  Device calibration is row f1.
ZIGZAG_LAGS = [1.48082e+06, 489354, 240986, 120483, 67150.3, 38647.3, 21660.1, 14104.1, 8775.52, 5010.28, 2619.84]
SEG = os.path.join(ROOT, "bench_data", "zigzag_n8_100Mb_seed1.seg.gz")
NSAM = 8


def workload(len_mb, chunks):
    """rows of the `chunks` equal windows of the first len_mb Mb -- the -startpos / -chunks decomposition of the
    reference frontend (smcsmc/model.py:563-662): each chunk is an independent sweep that restarts from the prior"""
    from smcsmc_b200.segdata import parse_seg, window_rows
    seqlen = len_mb * 1e6 / chunks
    max_seg_len = int(2.0 / (RHO * 4 * N0))                    # pfparam.cpp:364
    raw = parse_seg(SEG, NSAM)
    rows = [window_rows(raw, seqlen, data_start=1 + int(c * seqlen), max_segment_length=max_seg_len)
            for c in range(chunks)]
    t_start = [4 * N0 * t for t in ZIGZAG_TIMES]
    inv2N = [1.0 / (2 * N0)] * len(t_start)
    return rows, t_start, inv2N, seqlen


class Clocks(threading.Thread):
    """nvidia-smi clock / throttle-reason samples during the timed region (B200_PROFILING.md recipe)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.lines, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.lines.append(line.strip())
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, smax, reasons = [], 0.0, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax = max(smax, float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes_per_launch(n, N, E, resample_fraction):
    """DESIGN.md 'roofline': per particle and row the state is read and written once, S = 8(n-1) + (2n-1) + 16 bytes,
    plus one read-modify-write of the open recombination-opportunity cell per epoch (16 E) and the 8-byte weight prefix;
    a resampling row additionally gathers the parent's S bytes.  What the gather moves of the tallies (32 B of open
    cells per epoch and a 4-byte ancestor slot per pending lag bucket) is NOT counted: it is overhead of the design,
    visible as traffic > algorithmic."""
    S = 8 * (n - 1) + (2 * n - 1) + 16
    return N * (2 * S + 16 * E + 8 + resample_fraction * S)


def run_ours(a):
    import torch
    import torch.distributed as dist
    import smcsmc_b200 as S
    from smcsmc_b200.sharding import allreduce_tallies
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    shard = a.shard and world > 1
    C = 1 if shard else a.chunks
    rows, t_start, inv2N, seqlen = workload(a.len_mb, C)
    R, E, N = sum(len(r.x_end) for r in rows), len(t_start), a.particles
    if shard:
        # ONE population of a.particles split over the ranks (smc_shard_attach): total work fixed => strong scaling
        from smcsmc_b200.sharding import broadcast_unique_id
        assert a.particles % world == 0
        N = a.particles // world
    # pinned host staging of the step's inputs (the e2e leg copies from here)
    pin = []
    for r in rows:
        d = dict(x_end=r.x_end, alleles=r.alleles, is_site=r.is_site, max_rec=r.max_epoch_to_record(ZIGZAG_LAGS))
        pin.append({k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in d.items()})
    h2d = sum(t.numel() * t.element_size() for d in pin for t in d.values())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2

    def make_engines(seed, lags=ZIGZAG_LAGS, max_rec=None):
        pfs = []
        for c in range(C):
            if shard:
                pf = S.ParticleFilter(NSAM, N, t_start, inv2N, MU, RHO, seqlen, lags, seed=seed, chunk=c,
                                      device=local, ess_fraction=a.ess,
                                      shard=(rank, world, broadcast_unique_id(device="cuda")))
            else:
                pf = S.ParticleFilter(NSAM, N, t_start, inv2N, MU, RHO, seqlen, lags, seed=seed,
                                      chunk=rank * C + c, device=local, ess_fraction=a.ess)
            d = pin[c]
            pf.upload_rows(d["x_end"].numpy(), d["alleles"].numpy(), d["is_site"].numpy(),
                           d["max_rec"].numpy() if max_rec is None else max_rec[c])
            pfs.append(pf)
        return pfs

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def collect(pfs):
        t = np.sum([pf.tallies() for pf in pfs], axis=0)          # chunk tallies are summed (model.py:1176-1184)
        lnz = float(np.sum([pf.ln_normalization_factor() for pf in pfs]))
        if shard:
            return t, lnz                                           # smc_finish already made them global
        return allreduce_tallies(t, lnz, device="cuda")           # the one exchange per EM iteration (NCCL when N > 1)

    def timing(pfs):
        ms, k = 0.0, 0
        for pf in pfs:
            m_, k_ = pf.timing()
            ms += m_; k += k_
        return ms, k

    def sweep_all(pfs):
        for pf in pfs:
            pf.init_particles()
        if shard:
            for pf in pfs:
                pf.sweep()
        else:
            S.sweep_many(pfs, a.rows_per_turn)
        for pf in pfs:
            pf.finish()

    # ---- resident leg: rows in HBM, engines re-used ----
    pfs = make_engines(1)
    def step_resident(seed):
        flush.fill_(seed & 0xff)                     # flush L2 between iterations
        torch.cuda.synchronize()
        for pf in pfs:
            pf.reset(seed=seed)
        sweep_all(pfs)
        return collect(pfs)
    for w in range(a.warmup):
        step_resident(100 + w)
    clk = Clocks(local)
    clk.start()
    time.sleep(0.3)
    ms0, l0 = timing(pfs)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    barrier()
    t0 = time.perf_counter()
    ev[0].record()
    for k in range(a.steps):
        tallies, lnZ = step_resident(200 + k)
    ev[1].record()
    barrier()
    wall_host = time.perf_counter() - t0
    wall = ev[0].elapsed_time(ev[1]) * 1e-3       # device clock; every step ends with a device-wide synchronize
    ms1, l1 = timing(pfs)
    cs = [pf.counters() for pf in pfs]
    counters = {k: sum(c[k] for c in cs) for k in cs[0]}
    for pf in pfs:
        pf.close()
    dev_ms = ms1 - ms0
    launches = l1 - l0
    # ---- e2e leg: host buffers through the public calls, H2D + D2H inside the timed region ----
    cal_events = []

    def step_e2e(seed):
        tr = [time.perf_counter()]
        # lag calibration as the drop-in driver does it (smcsmc_main.cpp: 8192 prior ARGs on the device, medians per
        # epoch, lag = 2 x median) and the rows' max_epoch_to_update under those lags -- inside the timed region
        med, cnt = S.calibrate_survival(NSAM, t_start, inv2N, RHO, seqlen, seed=1, n_args=8192, device=local)
        lags = [2.0 * x for x in med]
        cal_events.append(int(min(cnt)))
        max_rec = [r.max_epoch_to_record(lags) for r in rows]
        p = make_engines(seed, lags, max_rec)
        tr.append(time.perf_counter())
        sweep_all(p)
        tr.append(time.perf_counter())
        out = collect(p)
        tr.append(time.perf_counter())
        for pf in p:
            pf.close()
        tr.append(time.perf_counter())
        if os.environ.get("BENCH_TRACE") and rank == 0:
            print("[bench] e2e step: calibrate+create+upload %.1f ms, sweep %.1f ms, read-back %.1f ms, destroy %.1f ms"
                  % tuple((tr[k + 1] - tr[k]) * 1e3 for k in range(4)), file=sys.stderr, flush=True)
        return out
    step_e2e(300)
    barrier()
    ev[2].record()
    for k in range(a.steps):
        step_e2e(400 + k)
    ev[3].record()
    barrier()
    wall_e2e = ev[2].elapsed_time(ev[3]) * 1e-3
    clocks = clk.stop()
    # max over ranks
    if world > 1:
        tt = torch.tensor([wall, wall_e2e, dev_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        wall, wall_e2e, dev_ms = [float(x) for x in tt.cpu()]
    units = float(N) * seqlen * C * a.steps * world        # sharded: N is the per-rank share, so this is the whole population
    rs_frac = counters["resamples"] / max(counters["rows"], 1)
    avg_launch_s = dev_ms * 1e-3 / max(launches, 1)
    abytes = algorithmic_bytes_per_launch(NSAM, N, E, rs_frac)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = abytes / avg_launch_s / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum per launch, mean of the six launches in
    # profiles/r1_row_kernel_final_ncu_full.csv (n = 8, E = 11, N = 10 000; ncu runs every launch cold-cache, so the
    # ancestor slots and open cells a resampling row gathers all come from DRAM: 4.8 - 8.1 MB per launch)
    traffic = None
    if (NSAM, E, N) == (8, 11, 10000):
        traffic = 6.78e6
    line = {
        "metric": "particle*sites/sec (s per EM iteration in ms_per_step)", "value": units / wall,
        "unit": "particle*sites/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": wall / a.steps * 1e3, "higher_is_better": True, "scaling": "strong" if shard else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic (scrm-simulated zigzag, bench_data/, seed 1)",
        "config": {"workload": ("8 haplotypes x %g Mb zigzag, ONE population of %d particles sharded over %d GPUs, 11 epochs"
                                % (a.len_mb, a.particles, world)) if shard else
                               "configs[1]: 8 haplotypes x %g Mb zigzag, %d particles, 11 epochs, 1 EM iteration per step, "
                               "swept as %d concurrent chunks (smc2 -chunks)" % (a.len_mb, N, C), "chunks": C, "rows": R, "particles": N, "haplotypes": NSAM, "epochs": E,
                   "per_gpu": "the whole %g Mb data set per GPU" % a.len_mb, "l2": "flushed between steps (256 MiB write)",
                   "lags": "value: reference-calibrated (DESIGN.md); e2e: calibrated on the device every step"},
        "e2e": {"value": units / wall_e2e, "unit": "particle*sites/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": C * (6 * E * 8 + 8) + 2 * E * 8, "ms_per_step": wall_e2e / a.steps * 1e3,
                "includes": "device lag calibration (8192 ARGs, fewest events in an epoch: %d), max_epoch_to_update, "
                            "engine create/upload/sweep/finish/read-back/destroy" % (min(cal_events) if cal_events else -1)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "kernel": "smc_row_kernel", "avg_launch_us": avg_launch_s * 1e6,
                     "algorithmic_bytes_per_launch": abytes, "peak_source": "measured" if peaks else "fallback",
                     "note": "one launch per .seg row and chunk (%d rows/step), %d chunks concurrent; avg_launch_us = "
                             "device time / launches" % (R, C)},
        "roofline_issue": issue_roofline(S, local, clocks, float(N) * R * a.steps * world / wall / world, rs_frac,
                                         (NSAM, E) == (8, 11)),
        "clocks": clocks,
        "device_ms_per_step": dev_ms / a.steps, "host_wall_ms_per_step": wall_host / a.steps * 1e3,
        "particle_rows_per_s": float(N) * R * a.steps * world / wall,
        "resample_fraction": rs_frac, "lnZ": lnZ,
    }
    if not a.no_extras and not shard:
        if world == 1:
            line["target_shape"] = extra_target_shape(S, torch, local, a)
            line["struct"] = extra_struct_driver(local)
        else:
            line["shard"] = extra_shard(S, torch, dist, rank, world, local, a)
    if rank == 0:
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_reference_sample(a)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _timed(torch, fn):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record()
    out = fn()
    ev1.record()
    torch.cuda.synchronize()
    return ev0.elapsed_time(ev1) * 1e-3, out


def extra_target_shape(S, torch, local, a, particles=100000, len_mb=16.0):
    """the north-star shape per GPU: ONE population of 10^5 particles (n = 8, 11 epochs) over 16 Mb in one chunk"""
    rows, t_start, inv2N, seqlen = workload(len_mb, 1)
    r = rows[0]
    pf = S.ParticleFilter(NSAM, particles, t_start, inv2N, MU, RHO, seqlen, ZIGZAG_LAGS, seed=7, device=local,
                          ess_fraction=a.ess)
    pf.upload_rows(r.x_end, r.alleles, r.is_site, r.max_epoch_to_record(ZIGZAG_LAGS))

    def step(seed):
        pf.reset(seed=seed)
        pf.init_particles()
        pf.sweep()
        pf.finish()
        return pf.ln_normalization_factor()
    step(1)                                                    # warm-up
    secs, lnz = _timed(torch, lambda: step(2))
    c = pf.counters()
    pf.close()
    return {"workload": "8 haplotypes x %g Mb zigzag, ONE population of %d particles, 11 epochs, one chunk" % (len_mb, particles),
            "value": particles * seqlen / secs, "unit": "particle*sites/s", "ms_per_step": secs * 1e3,
            "rows": len(r.x_end), "us_per_row": secs / len(r.x_end) * 1e6, "resamples": c["resamples"], "lnZ": lnz}


STRUCT_SEG = os.path.join(ROOT, "bench_data", "split_mig_n8_10Mb_seed3.seg.gz")
DRIVER_BIN = os.path.join(ROOT, "smcsmc_b200", "bin", "smcsmc")


def extra_struct_driver(local, particles=50000, length=10000000):
    """BASELINE.json configs[2] at full size through the drop-in driver itself (reference command line, .seg parsed from
    disk, device lag calibration, .out written): 4 + 4 haplotypes with split and directional migration, 50 000 particles
    x 10 Mb, one sweep.  Wall time of the process; the device time of the sweep is what the driver logs."""
    import gzip
    import shutil
    if not os.path.exists(DRIVER_BIN):
        return {"unavailable": "smcsmc_b200/bin/smcsmc not built"}
    with tempfile.TemporaryDirectory() as tmp:
        seg = os.path.join(tmp, "c2.seg")
        with gzip.open(STRUCT_SEG, "rb") as f, open(seg, "wb") as g:
            shutil.copyfileobj(f, g)
        theta, R = 4 * N0 * MU * length, 4 * N0 * RHO * length
        cmd = [DRIVER_BIN, "-nsam", "8", "-Np", str(particles), "-N0", repr(N0), "-t", repr(theta), "-r", repr(R),
               str(length), "-I", "2", "4", "4", "-em", "0", "2", "1", "1", "-eN", "0.1", "1", "-ej", "0.5", "2", "1",
               "-EM", "0", "-seed", "1", "-seg", seg, "-o", os.path.join(tmp, "c2")]
        env = dict(os.environ, SMCSMC_DEVICE=str(local))
        t0 = time.perf_counter()
        r = subprocess.run(cmd, capture_output=True, text=True, env=env)
        wall = time.perf_counter() - t0
        if r.returncode != 0:
            return {"unavailable": "driver exited with %d: %s" % (r.returncode, r.stderr[-300:])}
        dev_ms = None
        for ln in r.stderr.splitlines():
            if "device time" in ln:
                try:
                    dev_ms = float(ln.split("device time")[1].split("ms")[0].strip(" :="))
                except Exception:
                    pass
    out = {"workload": "configs[2]: 4+4 haplotypes, split at 0.5 and migration 2->1, %d particles x %g Mb, 3 epochs, one sweep "
                       "through smcsmc_b200/bin/smcsmc" % (particles, length / 1e6),
           "wall_s": wall, "value": particles * length / wall, "unit": "particle*sites/s (process wall: parse + calibrate + sweep + write)"}
    if dev_ms:
        out["sweep_device_ms"] = dev_ms
        out["value_sweep"] = particles * length / (dev_ms * 1e-3)
    return out


def extra_shard(S, torch, dist, rank, world, local, a, particles=1000000, len_mb=4.0):
    """ONE population of 10^6 particles (n = 8, 11 epochs) x 4 Mb split over the ranks (smc_shard_attach): strong scaling.
    value = resident sweep (engine re-used), e2e = attach + upload + sweep + finish + read-back + destroy."""
    from smcsmc_b200.sharding import broadcast_unique_id
    rows, t_start, inv2N, seqlen = workload(len_mb, 1)
    r = rows[0]
    n_loc = particles // world
    max_rec = r.max_epoch_to_record(ZIGZAG_LAGS)

    def make(seed):
        pf = S.ParticleFilter(NSAM, n_loc, t_start, inv2N, MU, RHO, seqlen, ZIGZAG_LAGS, seed=seed, device=local,
                              ess_fraction=a.ess, shard=(rank, world, broadcast_unique_id(device="cuda")))
        pf.upload_rows(r.x_end, r.alleles, r.is_site, max_rec)
        return pf

    def sweep(pf, seed=None):
        if seed is not None:
            pf.reset(seed=seed)
        pf.init_particles()
        pf.sweep()
        pf.finish()
        return pf.ln_normalization_factor()

    def tmax(x):
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.cpu())
    pf = make(11)
    sweep(pf)                                                  # warm-up
    dist.barrier()
    secs, lnz = _timed(torch, lambda: sweep(pf, 12))
    secs = tmax(secs)
    c = pf.counters()
    pf.close()
    dist.barrier()

    def e2e():
        p = make(13)
        z = sweep(p)
        p.tallies()
        p.close()
        return z
    secs_e2e, _ = _timed(torch, e2e)
    secs_e2e = tmax(secs_e2e)
    units = float(n_loc) * world * seqlen
    return {"workload": "8 haplotypes x %g Mb zigzag, ONE population of %d particles sharded over %d GPUs, 11 epochs"
                        % (len_mb, n_loc * world, world), "scaling": "strong", "value": units / secs,
            "unit": "particle*sites/s", "ms_per_step": secs * 1e3, "rows": len(r.x_end),
            "us_per_row": secs / len(r.x_end) * 1e6, "resamples": c["resamples"], "lnZ": lnz,
            "e2e": {"value": units / secs_e2e, "ms_per_step": secs_e2e * 1e3}}


# ---------------------------------------------------------------------------------------------------------------
# the reference's own CPU implementation of the path (oracle/_ref/smcsmc, built unmodified from the reference)
# ---------------------------------------------------------------------------------------------------------------
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "smcsmc")


def _host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _ref_chunk_cmd(seg_path, out_prefix, start, length, particles, seed):
    theta, R = 4 * N0 * MU * length, 4 * N0 * RHO * length
    cmd = [REF_BIN, "-nsam", str(NSAM), "-Np", str(particles), "-N0", repr(N0), "-t", repr(theta), "-r", repr(R),
           str(int(length))]
    for t in ZIGZAG_TIMES:
        cmd += ["-eN", repr(t), "1"]
    cmd += ["-EM", "0", "-seed", str(seed), "-seg", seg_path, "-startpos", str(int(start)), "-o", out_prefix]
    return cmd


def reference_sample(particles, chunk_mb, cores, seed=1):
    """Time the reference binary the way smc2 parallelises it (model.py:563-662,1094-1100): `cores` independent chunk
    processes of chunk_mb Mb each, one per host core, wall time of the slowest.  Returns (particle*sites/s, seconds)."""
    import gzip
    if not os.path.exists(REF_BIN):
        raise RuntimeError("oracle/_ref/smcsmc missing (built by __graft_entry__.build() where /root/reference exists)")
    length = chunk_mb * 1e6
    with tempfile.TemporaryDirectory() as tmp:
        seg = os.path.join(tmp, "sample.seg")
        need = cores * length + 1
        with gzip.open(SEG, "rt") as f, open(seg, "w") as g:
            for line in f:
                g.write(line)
                if int(line.split("\t", 1)[0]) > need:
                    break
        t0 = time.perf_counter()
        procs = [subprocess.Popen(_ref_chunk_cmd(seg, os.path.join(tmp, "c%d" % c), 1 + c * length, length, particles,
                                                 seed + c), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                 for c in range(cores)]
        rcs = [p.wait() for p in procs]
        wall = time.perf_counter() - t0
        if any(rcs):
            raise RuntimeError("reference smcsmc exited with %s" % rcs)
        ok = all("LogL" in open(os.path.join(tmp, "c%d.out" % c)).read() for c in range(cores))
        if not ok:
            raise RuntimeError("reference run wrote no LogL row")
    return cores * particles * length / wall, wall


def cpu_reference_sample(a):
    cores = _host_cores()
    particles, chunk_mb = a.cpu_particles, a.cpu_chunk_mb
    v, wall = reference_sample(particles, chunk_mb, cores)
    # the reference's fixed per-process lag calibration (smcsmc.cpp:287-288), timed with a single particle, so that the
    # sweep alone can be read off as well (SURVEY.md 8d reports it separately)
    _, cal = reference_sample(1, chunk_mb, cores, seed=7)
    sweep = max(wall - cal, 1e-9)
    return {"value": v, "unit": "particle*sites/s", "cores": cores, "kind": "reference",
            "calibration_s": cal, "value_excluding_calibration": cores * particles * chunk_mb * 1e6 / sweep,
            "sample": "%d chunk processes x %g Mb of the same .seg x %d particles, same 11-epoch model, -EM 0, wall of the "
                      "slowest (%.1f s, of which %.1f s is the reference's per-process lag calibration)"
                      % (cores, chunk_mb, particles, wall, cal)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = _host_cores()
    for w in range(min(a.warmup, 1)):                       # one warm-up pass is enough for a CPU binary (page cache)
        reference_sample(max(a.cpu_particles // 10, 10), a.cpu_chunk_mb, cores, seed=50)
    t0 = time.perf_counter()
    vals = []
    for k in range(a.steps):
        v, wall = reference_sample(a.cpu_particles, a.cpu_chunk_mb, cores, seed=60 + k)
        vals.append(v)
    total = time.perf_counter() - t0
    value = cores * a.cpu_particles * a.cpu_chunk_mb * 1e6 * a.steps / total
    sample = ("%d chunk processes x %g Mb x %d particles per step (bounded sample of configs[1]; the reference is linear "
              "in particles x length)" % (cores, a.cpu_chunk_mb, a.cpu_particles))
    line = {"impl": "reference", "metric": "particle*sites/sec (s per EM iteration in ms_per_step)", "value": value,
            "unit": "particle*sites/s", "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": total / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic (scrm-simulated zigzag, bench_data/, seed 1)",
            "config": {"workload": "configs[1]: 8 haplotypes x 100 Mb zigzag, 11 epochs -- bounded sample: " + sample},
            "cpu_baseline": {"value": value, "unit": "particle*sites/s", "cores": cores, "kind": "reference",
                             "sample": sample},
            "e2e": {"value": value, "unit": "particle*sites/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def issue_roofline(S, device, clocks, particle_rows_per_s_per_gpu, rs_frac, profiled_shape):
    """The row kernel against the resource that binds it.  It moves 3 % of what HBM could (roofline above) and keeps the
    fp64 pipe 5 % busy; what it uses up is instruction issue: ~10 k SASS instructions of dependent, divergent fp64 /
    shared-memory code per particle and row.  achieved = warp instructions per second, from the per-launch instruction
    count of the ncu capture of this kernel (profiles/r2_final_row_kernel_N160k_ncu_full.csv, smsp__inst_executed.sum
    / particles: 273 per particle-row on a plain row, 369 on a resampling row; n = 8, E = 11) x the particle-rows per
    second measured here; peak = SMs x 4 schedulers x the SM clock sampled during the run (one warp instruction per
    scheduler and cycle).  The fp64 FMA peak is measured live by a register-only probe kernel (smc_measure_fp64_fma)."""
    import torch
    sms = torch.cuda.get_device_properties(device).multi_processor_count
    mhz = float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0)
    peak = sms * 4 * mhz * 1e6
    out = {"bound": "issue", "unit": "warp-inst/s", "peak": peak, "sms": sms, "sm_mhz": mhz,
           "dfma_peak_tflops": S.measure_fp64_fma(device), "fp64_pipe_active_frac": 0.056,
           "source": "instruction counts and fp64 pipe utilisation: ncu capture under profiles/ (static); rate: this run"}
    if profiled_shape:
        per = 273.0 * (1.0 - rs_frac) + 369.0 * rs_frac
        out.update({"warp_inst_per_particle_row": per, "achieved": per * particle_rows_per_s_per_gpu,
                    "frac": per * particle_rows_per_s_per_gpu / peak})
    else:
        out.update({"achieved": None, "frac": None})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=10000)
    ap.add_argument("--len-mb", type=float, default=100.0)
    ap.add_argument("--chunks", type=int, default=16, help="independent chunks swept concurrently (smc2 -chunks)")
    ap.add_argument("--rows-per-turn", type=int, default=1)
    ap.add_argument("--ess", type=float, default=0.5, help="-ESS resampling threshold fraction (reference default 0.5)")
    ap.add_argument("--cpu-particles", type=int, default=1000)
    ap.add_argument("--cpu-chunk-mb", type=float, default=6.25,
                    help="chunk length of the bounded CPU sample (6.25 Mb = configs[1]'s 100 Mb in 16 chunks)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the target_shape / struct / shard records")
    ap.add_argument("--shard", action="store_true",
                    help="with N > 1 ranks: split ONE population of --particles over the GPUs (smc_shard_attach) instead "
                         "of giving every GPU its own chunks; a single chunk of --len-mb is swept")
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
