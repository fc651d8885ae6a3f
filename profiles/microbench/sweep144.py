"""A few update() calls of config C's shape (N = 144, 128 chains); run under ncu for per-kernel times."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import _lib
if os.environ.get("MCB_LIB"):  # a differently built library (timing variants)
    _lib.LIB_PATH = os.environ["MCB_LIB"]

model = mcb.HubbardModelAttractive(L=12, U=4.0)
mc = mcb.DQMC(model, n_chains=128, seed=1, beta=2.0, delta_tau=0.1, safe_mult=10)
mc.build_stack()
mc.update(12)
mc.close()
