"""Golden vectors for MODE_SGBM_3WAY (mode 2) from the cv2 4.13.0 binary of this container.
Run here (needs cv2):  python tests/golden/make_golden_3way.py -> tests/golden/golden_3way.npz + golden_3way_full.json

cv2 runs this mode as 4 fixed row stripes ("the number of stripes is fixed, disregarding the number of threads"): the
script also checks that the output does not depend on cv2.setNumThreads."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import cases  # noqa: E402


def main():
    import cv2
    out = {}
    for i, c in enumerate(cases.sgbm_3way_cases()):
        L, R = cases.sgbm_inputs(c)
        out[f"w3_{i}"] = cv2.StereoSGBM_create(**c["kw"]).compute(L, R)
    np.savez_compressed(os.path.join(HERE, "golden_3way.npz"), **out)
    full = {"cv2": cv2.__version__}
    for name in ("c1", "c2", "c3"):
        L, R, kw = cases.full_inputs(name)
        kw = dict(kw, mode=2)
        hs = []
        for nt in (1, 3, 8):
            cv2.setNumThreads(nt)
            d = cv2.StereoSGBM_create(**kw).compute(L, R)
            hs.append(hashlib.sha256(d.tobytes()).hexdigest())
        assert len(set(hs)) == 1, f"{name}: MODE_SGBM_3WAY output depends on the thread count: {hs}"
        full[f"{name}_3way"] = hs[0]
        full[f"{name}_valid_frac"] = float((d >= 0).mean())
    full["thread_counts_checked"] = [1, 3, 8]
    with open(os.path.join(HERE, "golden_3way_full.json"), "w") as f:
        json.dump(full, f, indent=1)
    print("wrote", len(out), "cases, cv2", cv2.__version__)


if __name__ == "__main__":
    main()
