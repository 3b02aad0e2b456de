"""Golden vectors for MODE_HH4 (mode 3) from the cv2 4.13.0 binary of this container.
Run here (needs cv2):  python tests/golden/make_golden_hh4.py -> tests/golden/golden_hh4.npz (+ full-size hash)"""
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
    for i, c in enumerate(cases.sgbm_hh4_cases()):
        L, R = cases.sgbm_inputs(c)
        out[f"hh4_{i}"] = cv2.StereoSGBM_create(**c["kw"]).compute(L, R)
    np.savez_compressed(os.path.join(HERE, "golden_hh4.npz"), **out)
    L, R, kw = cases.full_inputs("c3")
    kw = dict(kw, mode=3)
    d = cv2.StereoSGBM_create(**kw).compute(L, R)
    with open(os.path.join(HERE, "golden_hh4_full.json"), "w") as f:
        json.dump({"cv2": cv2.__version__, "c3_hh4": hashlib.sha256(d.tobytes()).hexdigest(),
                   "valid_frac": float((d >= 0).mean())}, f, indent=1)
    print("wrote", len(out), "cases, cv2", cv2.__version__)


if __name__ == "__main__":
    main()
