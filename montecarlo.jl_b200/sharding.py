"""Multi-GPU layer: Markov chains / replicas are independent (`run!` only touches its own `mc`,
src/flavors/DQMC/DQMC.jl:369-515), so N GPUs run N disjoint shards of the global chain list with no
data-path collective. Philox streams are keyed by the GLOBAL chain id, so results do not depend on
the partition. The only exchange is the final reduction of the observable moments
(sum x, sum x^2, count) over ranks: one `all_reduce(SUM)` per run through `torch.distributed`
(NCCL over NVLink on the GPU box, gloo in the CPU tests).

One process per GPU, launched by torchrun; RANK / LOCAL_RANK / WORLD_SIZE come from the environment.
"""
import math
import os

import numpy as np


def _prefer_bundled_nccl():
    """If the NCCL wheel that torch links against is installed, point the library at it (MCB_NCCL_LIB): a process that
    loads the system libnccl.so.2 first and imports torch afterwards would resolve torch's NCCL symbols against it."""
    if os.environ.get("MCB_NCCL_LIB"):
        return
    import importlib.util
    spec = importlib.util.find_spec("nvidia")
    for root in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
        cand = os.path.join(root, "nccl", "lib", "libnccl.so.2")
        if os.path.exists(cand):
            os.environ["MCB_NCCL_LIB"] = cand
            return


class NcclComm:
    """Communicator of the C-ABI collective (mcb_comm_*, csrc/comm.cu): NCCL loaded by the library itself, no torch.
    Rank 0 calls NcclComm.unique_id() and passes the 128 bytes to the other ranks by any means."""

    @staticmethod
    def unique_id():
        import ctypes as C
        from . import _lib
        _prefer_bundled_nccl()
        buf = C.create_string_buffer(128)
        _lib.check(_lib.lib().mcb_comm_unique_id(buf))
        return buf.raw

    def __init__(self, unique_id, n_ranks, rank, device=0):
        import ctypes as C
        from . import _lib
        _prefer_bundled_nccl()
        self.L = _lib.lib()
        self.h = C.c_void_p()
        _lib.check(self.L.mcb_comm_create(C.c_char_p(bytes(unique_id)), n_ranks, rank, device, C.byref(self.h)))
        self.n_ranks, self.rank = n_ranks, rank

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.L.mcb_comm_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def dist_env():
    """(rank, local_rank, world_size) from the torchrun environment; (0, 0, 1) when run alone."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def shard(total, rank, world):
    """Static block partition of `total` global ids: returns (offset, count) of `rank`.
    The first `total % world` ranks hold one extra id (ragged totals are allowed, empty shards too)."""
    if total < 0 or world <= 0 or not 0 <= rank < world:
        raise ValueError("bad partition arguments")
    base, extra = divmod(total, world)
    count = base + (1 if rank < extra else 0)
    offset = rank * base + min(rank, extra)
    return offset, count


def local_moments(per_chain):
    """per-chain observable arrays (leading axis = chain) -> dict of [sum, sum of squares] stacked,
    plus the chain count. This is the payload of the collective."""
    out = {}
    n = None
    for key, val in per_chain.items():
        a = np.asarray(val, dtype=np.float64)
        if a.ndim == 0:
            continue  # not per chain (e.g. the measurement count)
        n = a.shape[0] if n is None else n
        if a.shape[0] != n:
            raise ValueError(f"observable {key!r} has {a.shape[0]} chains, expected {n}")
        out[key] = np.stack([a.sum(axis=0), (a * a).sum(axis=0)])
    return out, (0 if n is None else n)


def allreduce_moments(moments, count, group=None, device=None):
    """SUM-all-reduce of the moment arrays and of the chain count over the process group.
    Without an initialised process group (single process) it is the identity."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return moments, count
    keys = sorted(moments)
    flat = np.concatenate([moments[k].reshape(-1) for k in keys] + [np.array([float(count)])])
    t = torch.from_numpy(flat)
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    flat = t.cpu().numpy()
    out, pos = {}, 0
    for k in keys:
        size = moments[k].size
        out[k] = flat[pos:pos + size].reshape(moments[k].shape)
        pos += size
    return out, int(round(flat[pos]))


def mean_and_error_from_moments(moments, count):
    """mean over all chains and its standard error (chains are independent Markov chains)."""
    res = {}
    for k, m in moments.items():
        mean = m[0] / max(count, 1)
        if count > 1:
            var = np.maximum(m[1] / count - mean * mean, 0.0) * count / (count - 1)
            err = np.sqrt(var / count)
        else:
            err = np.zeros_like(mean)
        res[k] = (mean, err)
    return res


def reduce_observables(per_chain, group=None, device=None):
    """Global mean and standard error of per-chain observables held by all ranks."""
    mom, n = local_moments(per_chain)
    mom, n = allreduce_moments(mom, n, group=group, device=device)
    return mean_and_error_from_moments(mom, n), n


class ShardedDQMC:
    """`DQMC(model; ...)` for `total_chains` chains spread over the ranks of the job.
    Every rank owns the chains [offset, offset + count) of the global list on its local GPU."""

    def __init__(self, model, total_chains, seed=123, group=None, **kwargs):
        from .dqmc import DQMC
        self.rank, self.local_rank, self.world = dist_env()
        self.offset, self.count = shard(total_chains, self.rank, self.world)
        self.total_chains, self.group = total_chains, group
        self.mc = None
        self.model = model
        if self.count > 0:
            self.mc = DQMC(model, n_chains=self.count, seed=seed, chain_offset=self.offset,
                           device=self.local_rank, **kwargs)

    def run(self, thermalization=None, sweeps=None):
        if self.mc is not None:
            obs = self.mc.run(thermalization, sweeps)
            per_chain = {k: v for k, v in obs.items() if k != "count"}
        else:
            # a rank without chains (total_chains < world_size) sends zero moments of the SAME layout; an
            # all_reduce with mismatched payload sizes would hang or corrupt
            from .dqmc import DQMC
            per_chain = {k: np.zeros((0,) + shp) for k, shp in DQMC.observable_shapes(self.model).items()}
        device = None
        try:
            import torch
            if torch.cuda.is_available():
                device = torch.device("cuda", self.local_rank)
        except ImportError:
            pass
        return reduce_observables(per_chain, group=self.group, device=device)

    def close(self):
        if self.mc is not None:
            self.mc.close()


def ising_moments(E, M, beta, L):
    """Specific heat / susceptibility estimators of Ising/measurements.jl:38-44, 80-86 from per-sweep
    (E, |M|) samples of ONE replica: returns dict of scalars."""
    E = np.asarray(E, dtype=np.float64)
    M = np.abs(np.asarray(M, dtype=np.float64))
    n = float(L * L)
    e1, e2, m1, m2 = E.mean(), (E * E).mean(), M.mean(), (M * M).mean()
    return dict(E=e1, E2=e2, e=e1 / n, C=beta * beta * (e2 - e1 * e1) / n,
                M=m1, M2=m2, m=m1 / n, chi=beta * (m2 - m1 * m1) / n)


def _selfcheck():
    assert shard(10, 0, 4) == (0, 3) and shard(10, 3, 4) == (8, 2)
    assert math.isclose(sum(shard(1024, r, 8)[1] for r in range(8)), 1024)


_selfcheck()
