"""Extracts the literal golden vectors of the reference's statistical integration tests into JSON fixtures.

Source: /root/reference/test/integration_tests.jl
  :28-93   "DQMC: attractive HubbardModel Simulation"  HubbardModelAttractive(4, 2), DQMC(m, beta=1.0),
           run!(sweeps=1000, thermalization=10), measure_rate 10 (default)  -> attractive_4x4_beta1.json
  :96-175  "DQMC: repulsive HubbardModel Simulation"   HubbardModelRepulsive(2, 2), beta=1.0, K=3
                                                                            -> repulsive_2x2_beta1.json
The reference's runs are seeded with Julia's MersenneTwister, which cannot be reproduced here, so these are
*statistical* goldens: means with the reference's own std_error and its own tolerance (atol = 4 dtau^2 resp.
2 dtau^2). Run in the build container (reads /root/reference); the JSON files are committed.
"""
import json
import os
import re

SRC = "/root/reference/test/integration_tests.jl"
HERE = os.path.dirname(os.path.abspath(__file__))


def literal_arrays(block):
    """All `@test [ ... ] ≈ <expr>` literals of a testset: list of (expr, rows) with rows a list of lists."""
    out = []
    for m in re.finditer(r"@test \[\s*(.*?)\s*\]\s*(?:\[:\])?\s*≈\s*([^\n]*)", block, re.S):
        body, expr = m.group(1), m.group(2).strip()
        rows = []
        for row in body.split(";"):
            toks = [t for t in re.split(r"[\s,]+", row.strip()) if t]
            vals = []
            i = 0
            while i < len(toks):
                t = toks[i]
                if i + 2 < len(toks) and toks[i + 1] == "+" and toks[i + 2].endswith("im"):
                    vals.append(float(t))  # real part of a complex literal (imaginary parts are 0.0)
                    i += 3
                    continue
                vals.append(float(t))
                i += 1
            rows.append(vals)
        out.append((expr, rows))
    return out


def main():
    text = open(SRC).read()
    a0 = text.index('@testset "DQMC: attractive HubbardModel Simulation"')
    r0 = text.index('@testset "DQMC: repulsive HubbardModel Simulation"')
    r1 = text.index("# TODO", r0)
    for name, block, meta in (
        ("attractive_4x4_beta1", text[a0:r0],
         dict(model="HubbardModelAttractive(4, 2)", L=4, U=1.0, mu=0.0, t=1.0, beta=1.0, delta_tau=0.1,
              safe_mult=10, sweeps=1000, thermalization=10, measure_rate=10, K=5, atol=4 * 0.1 ** 2,
              source="test/integration_tests.jl:28-93")),
        ("repulsive_2x2_beta1", text[r0:r1],
         dict(model="HubbardModelRepulsive(2, 2)", L=2, U=1.0, t=1.0, beta=1.0, delta_tau=0.1,
              safe_mult=10, sweeps=1000, thermalization=10, measure_rate=10, K=3, atol=2 * 0.1 ** 2,
              source="test/integration_tests.jl:96-175")),
    ):
        rec = dict(meta)
        for expr, rows in literal_arrays(block):
            key = re.search(r"measured\[:(\w+)\]", expr)
            if key is None:
                key = "conf_mean"
            else:
                key = key.group(1) + ("_std_error" if "std_error" in expr else "_mean")
            flat = rows[0] if len(rows) == 1 else rows
            rec[key] = flat
        with open(os.path.join(HERE, name + ".json"), "w") as f:
            json.dump(rec, f, indent=1)
        print(name, {k: (len(v) if isinstance(v, list) else v) for k, v in rec.items()})


if __name__ == "__main__":
    main()
