"""Multi-GPU plumbing for the chunk decomposition: chunks are independent units -- the reference runs
them as separate processes and sums the Opp/Count columns (smcsmc/model.py:563-662,1176-1184) -- so ranks own disjoint
chunk sets, there is no data-path collective, and the only exchange is one all-reduce of the tallies and of the
log-likelihood per EM iteration.  torch.distributed is used for the plumbing only (NCCL on GPUs, gloo in the CPU tests).
"""
import numpy as np


def chunks_for_rank(num_chunks, rank, world):
    """contiguous, balanced assignment of chunk ids to ranks (sizes differ by at most one)"""
    if not (0 <= rank < world) or num_chunks < 0:
        raise ValueError("bad rank/world/num_chunks")
    base, extra = divmod(num_chunks, world)
    lo = rank * base + min(rank, extra)
    return list(range(lo, lo + base + (1 if rank < extra else 0)))


def chunk_windows(total_len, num_chunks, data_start=1):
    """(data_start, seqlen) of each chunk: equal windows, the -startpos / locus-length pairs smc2 passes per chunk"""
    seqlen = float(total_len) / num_chunks
    return [(data_start + int(c * seqlen), seqlen) for c in range(num_chunks)]


def allreduce_tallies(tallies, lnZ, device=None):
    """sum the [6, E] tallies and the log-likelihood over ranks (identity when torch.distributed is not initialised)"""
    import torch
    import torch.distributed as dist
    buf = np.concatenate([np.asarray(tallies, dtype=np.float64).ravel(), [float(lnZ)]])
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return buf[:-1].reshape(np.shape(tallies)), float(buf[-1])
    t = torch.from_numpy(buf)
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t)
    out = t.cpu().numpy()
    return out[:-1].reshape(np.shape(tallies)), float(out[-1])


_ID_CACHE = {}


def broadcast_unique_id(device=None, fresh=False):
    """rank 0 draws the id of a sharded population (smc_shard_unique_id) and every rank of the default process group
    receives it; returns the 128 bytes to pass as ParticleFilter(shard=(rank, world, id)).  The id is drawn once per
    process group and handed out again on later calls (every rank must call in step): engines attached with the same id
    share one NCCL communicator and the peers' memory mappings inside libsmcgpu, so creating the next engine costs an
    allocation instead of a communicator.  fresh=True draws a new one."""
    import torch
    import torch.distributed as dist
    from .engine import shard_unique_id
    key = (dist.get_world_size(), dist.get_rank())
    if not fresh and key in _ID_CACHE:
        return _ID_CACHE[key]
    t = torch.zeros(128, dtype=torch.uint8)
    if dist.get_rank() == 0:
        t = torch.frombuffer(bytearray(shard_unique_id()), dtype=torch.uint8).clone()
    if device is not None:
        t = t.to(device)
    dist.broadcast(t, src=0)
    _ID_CACHE[key] = bytes(t.cpu().numpy().tobytes())
    return _ID_CACHE[key]
