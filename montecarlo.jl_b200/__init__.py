"""B200-native DQMC / Ising sweep engine behind MonteCarlo.jl's flavor/model API.

Host-side mirror of the reference interface (Python stands in for the Julia host,
which is absent from this image); all compute goes through the C ABI of
csrc/libmcb200.so (include/mcb200.h). There is no CPU fallback.
"""
from .lattices import SquareLattice  # noqa: F401
from .models import HubbardModelAttractive, HubbardModelRepulsive, IsingModel, IsingTc  # noqa: F401
from .dqmc import DQMC, DQMCParameters, CombinedGreensIterator, compress, decompress, mean_and_error  # noqa: F401
from .mc import MC  # noqa: F401
from ._lib import McbError  # noqa: F401
from .sharding import ShardedDQMC, shard, reduce_observables  # noqa: F401
