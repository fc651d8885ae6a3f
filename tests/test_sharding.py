"""N > 1 path on CPU: world_size-2 gloo process group. Chains shard with no data-path collective; the
only exchange is the all-reduce of observable moments. The GPU engine is not involved
here (no CUDA in this container): the per-chain observables are synthetic functions of the GLOBAL
chain id, so the sharded reduction must reproduce the single-process result exactly."""
import os
import socket

import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _fake_observables(offset, count):
    ids = np.arange(offset, offset + count, dtype=np.float64)
    return dict(occupation=1.0 + 0.01 * ids, G=np.sin(ids)[:, None, None] * np.ones((count, 3, 3)),
                cdc=np.stack([ids, ids * ids]).T if count else np.zeros((0, 2)))


def _worker(rank, world, port, total, q):
    import torch.distributed as dist
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r, lr, w = sharding.dist_env()
        offset, count = sharding.shard(total, r, w)
        res, n = sharding.reduce_observables(_fake_observables(offset, count))
        # HS fields are keyed by the global chain id: shard-local generation == global generation
        model = mcb.HubbardModelAttractive(L=2, U=1.0)
        confs = [model.rand_conf(10, 123, offset + c) for c in range(count)]
        q.put((rank, offset, count, n, {k: (v[0].copy(), v[1].copy()) for k, v in res.items()},
               [c.tobytes() for c in confs]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 16])
def test_world2_gloo_reduction_matches_single_process(total):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=120) for _ in procs], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # partition covers every chain exactly once
    assert out[0][1] == 0 and out[0][1] + out[0][2] == out[1][1] and out[1][1] + out[1][2] == total
    want, n_want = sharding.reduce_observables(_fake_observables(0, total))  # no process group here
    assert n_want == total
    for rank, _, _, n, res, _ in out:
        assert n == total
        for k in want:
            assert np.allclose(res[k][0], want[k][0], rtol=1e-13, atol=1e-13)
            assert np.allclose(res[k][1], want[k][1], rtol=1e-10, atol=1e-13)
    model = mcb.HubbardModelAttractive(L=2, U=1.0)
    glob = [model.rand_conf(10, 123, c).tobytes() for c in range(total)]
    assert out[0][5] + out[1][5] == glob


def test_shard_edge_cases():
    assert sharding.shard(0, 0, 2) == (0, 0)                       # empty input
    assert [sharding.shard(3, r, 8)[1] for r in range(8)] == [1, 1, 1, 0, 0, 0, 0, 0]  # ragged: empty shards
    assert [sharding.shard(1024, r, 8) for r in range(8)] == [(128 * r, 128) for r in range(8)]
    assert [sharding.shard(256, r, 8)[1] for r in range(8)] == [32] * 8
    with pytest.raises(ValueError):
        sharding.shard(4, 2, 2)
    res, n = sharding.reduce_observables({})                       # a rank with no chains
    assert n == 0 and res == {}


def test_error_bars_are_standard_errors_over_chains():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((64, 5))
    res, n = sharding.reduce_observables(dict(x=x))
    assert n == 64
    assert np.allclose(res["x"][0], x.mean(0))
    assert np.allclose(res["x"][1], x.std(0, ddof=1) / 8.0)


class _FakeDQMC:
    """Stands in for the CUDA-backed DQMC inside the CPU workers: observables are functions of the global chain id,
    in the real layout (DQMC.observable_shapes)."""
    observable_shapes = staticmethod(mcb.DQMC.observable_shapes)

    def __init__(self, model, n_chains=1, seed=123, chain_offset=0, device=0, **kw):
        self.model, self.n, self.off = model, n_chains, chain_offset

    def run(self, thermalization=None, sweeps=None):
        ids = np.arange(self.off, self.off + self.n, dtype=np.float64)
        out = dict(count=3)
        for k, shp in self.observable_shapes(self.model).items():
            out[k] = (1.0 + ids).reshape((-1,) + (1,) * len(shp)) * np.ones((self.n,) + shp)
        return out

    def close(self):
        pass


def _sharded_worker(rank, world, port, total, q):
    import torch.distributed as dist
    from montecarlo_jl_b200 import dqmc as dqmc_mod
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dqmc_mod.DQMC = _FakeDQMC
        sh = sharding.ShardedDQMC(mcb.HubbardModelAttractive(L=2, U=1.0), total, beta=1.0)
        res, n = sh.run(0, 1)
        q.put((rank, sh.count, n, {k: (np.array(v[0]), np.array(v[1])) for k, v in res.items()}))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [1, 3])
def test_sharded_run_with_an_empty_or_ragged_shard(total):
    """total_chains < world_size leaves a rank without chains: it must join the all-reduce with a payload of the
    same layout (mismatched sizes hang or corrupt)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sharded_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=120) for _ in procs], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [o[1] for o in out] == [total - total // 2, total // 2]
    ids = 1.0 + np.arange(total)
    for _, _, n, res in out:
        assert n == total
        assert set(res) == set(mcb.DQMC.observable_shapes(mcb.HubbardModelAttractive(L=2, U=1.0)))
        for k, (mean, err) in res.items():
            assert np.allclose(mean, ids.mean()) and mean.shape == mcb.DQMC.observable_shapes(
                mcb.HubbardModelAttractive(L=2, U=1.0))[k]
