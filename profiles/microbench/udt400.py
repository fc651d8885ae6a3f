"""One batched UDT of config D's shape (N = 400, 32 chains) through the C ABI; run under ncu for per-kernel times."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from montecarlo_jl_b200 import _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 400
b = int(sys.argv[2]) if len(sys.argv) > 2 else 32
rng = np.random.default_rng(1)
X = rng.standard_normal((b, n, n)) @ np.diag(np.logspace(0, -8, n))
L = _lib.lib()
U, T = np.empty(b * n * n), np.empty(b * n * n)
D, piv = np.empty(b * n), np.empty(b * n, dtype=np.int32)
Xc = np.ascontiguousarray(X.transpose(0, 2, 1)).reshape(-1)
for _ in range(3):
    _lib.check(L.mcb_linalg_udt(n, b, Xc, U, D, T, piv, 1, 0))
Um = U.reshape(b, n, n).transpose(0, 2, 1)
print("orth err", np.abs(Um[0] @ Um[0].T - np.eye(n)).max())
