"""Particle sharding over GPUs (smc_shard_attach): needs a box with >= 2 GPUs (`gpurun --gpus 2`); skipped elsewhere.
The ranks are separate processes (torch.distributed.run), as in bench.py."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_population_equals_single_gpu(world):
    if _gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29620 + world), os.path.join(ROOT, "tests", "sharded_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and line, (r.stdout[-3000:], r.stderr[-3000:])
    rep = json.loads(line[-1])
    assert rep["world"] == world
    for name, res in rep["cases"].items():
        assert res["ok"], (name, res)
        assert res["cross_rank_parents"] > 0, "no offspring ever crossed a rank boundary: the peer path went untested"


def test_attach_argument_checks():
    import numpy as np
    import smcsmc_b200 as S
    from smcsmc_b200._abi import check, lib
    pf = S.ParticleFilter(4, 8, [0.0], [1 / 2e4], 2.5e-8, 1e-8, 1e4, [1000.0])
    zero = np.zeros(128, dtype=np.uint8)
    with pytest.raises(S.SmcError, match="rank/world"):
        check(lib().smc_shard_attach(pf._h, 0, 1, zero.ctypes.data))
    with pytest.raises(S.SmcError, match="rank/world"):
        check(lib().smc_shard_attach(pf._h, 3, 2, zero.ctypes.data))
    pf.close()
