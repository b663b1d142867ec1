"""Generates tests/golden/mom_adaptive.json: methodOfMoments sources of a few cells from the
reference-shaped nested adaptive integrals of oracle/mom_oracle.py (SciPy QUADPACK QAGI standing
in for GSL QAGIU; MOM.C:259-362).  Slow (seconds per cell), hence a committed fixture.

    python tests/golden/make_mom_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import mom_oracle as mo  # noqa: E402

CT = (1.0, 1.0, 0.0, 1.0)  # c1, c2, gamma, sigma: O(1) scales, where QAGIU's epsabs 1e-7 means something
SPECS = {
    "ct_ct": {"coalescence": "CoulaloglouTavlaridesC", "coal_c": CT,
              "breakup": "CoulaloglouTavlarides", "br_c": CT},
    "const_binary": {"coalescence": "constantCoalescence", "coal_c": (0.1,),
                     "breakup": "binaryBreakup", "br_c": ()},
    "ct_none": {"coalescence": "CoulaloglouTavlaridesC", "coal_c": CT, "breakup": "noBreakup", "br_c": ()},
}


def main():
    rng = np.random.default_rng(20240611)
    cases = []
    for name, spec in SPECS.items():
        for _ in range(3):
            m0 = rng.uniform(0.5, 3.0)
            k = rng.uniform(1.5, 4.0)
            theta = rng.uniform(0.15, 0.6)
            m = [m0, m0 * k * theta, m0 * k * (k + 1) * theta ** 2]
            eps, rho, nu = rng.uniform(0.5, 2.0), rng.uniform(0.8, 1.2), rng.uniform(0.5, 1.5)
            S = mo.mom_source_cell(spec, m, eps, rho, nu)
            cases.append({"spec": name, "m": m, "eps": eps, "rho": rho, "nu": nu, "S": S.tolist()})
            print(name, m, S, flush=True)
    out = {"generator": "tests/golden/make_mom_golden.py", "specs": {k: {kk: list(vv) if isinstance(vv, tuple) else vv
                                                                        for kk, vv in v.items()} for k, v in SPECS.items()},
           "cases": cases}
    json.dump(out, open(os.path.join(HERE, "mom_adaptive.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
