#!/usr/bin/env python
"""Kernel-level micro-benchmark driver for the batched linear-algebra entry points (run under
`ncu --metrics gpu__time_duration.sum` to read the kernel durations; the calls include H2D/D2H copies).
usage: python profiles/microbench_linalg.py udt 64 512 [general]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from montecarlo_jl_b200 import _lib  # noqa: E402

op, n, b = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
if os.environ.get("MCB_LIB"):  # a differently built library (kernel variants under test)
    _lib.LIB_PATH = os.environ["MCB_LIB"]
L = _lib.lib()
if len(sys.argv) > 4:
    L.mcb_linalg_force_general(1)
rng = np.random.default_rng(0)
if op == "vmul":   # batched n x n products, all three transpose modes (kernel times under ncu)
    A = rng.standard_normal(b * n * n)
    Bm = rng.standard_normal(b * n * n)
    Cm = np.empty(b * n * n)
    for _ in range(3):
        for mode in (0, 1, 2):
            _lib.check(L.mcb_linalg_vmul(n, b, A, Bm, Cm, mode, 0))
    print("done", op, n, b)
    sys.exit(0)
X = rng.standard_normal((b, n, n)) @ np.diag(np.logspace(0, -8, n))
buf = np.ascontiguousarray(X.transpose(0, 2, 1)).reshape(-1)
U, T = np.empty(b * n * n), np.empty(b * n * n)
D, piv = np.empty(b * n), np.empty(b * n, dtype=np.int32)
for _ in range(3):
    _lib.check(L.mcb_linalg_udt(n, b, buf, U, D, T, piv, 1, 0))
if op == "rdivp":
    A = rng.standard_normal((b, n, n)).reshape(-1)
    _lib.check(L.mcb_linalg_udt(n, b, buf, U, D, T, piv, 0, 0))
    for _ in range(3):
        out = A.copy()
        _lib.check(L.mcb_linalg_rdivp(n, b, out, T, piv, 0))
print("done", op, n, b)
