"""CPU: host logic (parameter resolution, API surface, error mapping) and the C-ABI export list."""
import ctypes
import os
import re

import numpy as np
import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from stereo_vision_b200 import build, _lib
    build.build()
    return _lib.load()


def test_library_exports_every_symbol_in_header(lib):
    from stereo_vision_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "svb.h")).read()
    declared = set(re.findall(r"\b(svb_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    raw = ctypes.CDLL(_lib.SO_PATH)
    for name in declared:
        assert hasattr(raw, name), f"{name} not exported"
    assert b"sm_100a" in lib.svb_version()


def test_parameter_resolution_matches_survey_table_and_oracle(lib):
    import stereo_vision_b200 as sv
    rng = np.random.default_rng(0)
    for _ in range(200):
        kw = dict(minDisparity=int(rng.integers(-8, 9)), numDisparities=int(rng.integers(1, 300)),
                  blockSize=int(rng.integers(-2, 13)), P1=int(rng.integers(-5, 400)), P2=int(rng.integers(-5, 1600)),
                  disp12MaxDiff=int(rng.integers(-2, 5)), preFilterCap=int(rng.integers(-1, 70)),
                  uniquenessRatio=int(rng.integers(-2, 20)), speckleWindowSize=int(rng.integers(-1, 200)),
                  speckleRange=int(rng.integers(0, 4)), mode=int(rng.integers(0, 2)))
        assert sv.StereoSGBM_create(**kw).effective_params() == oracle.resolve(**kw)
    e = lambda **k: sv.StereoSGBM_create(**k).effective_params()
    assert e(blockSize=0)["blockSize"] == 5 and e(P1=0, P2=0)["P1"] == 2 and e(P1=0, P2=0)["P2"] == 5
    assert e(P1=10, P2=3)["P2"] == 11
    assert [e(preFilterCap=c)["preFilterCap"] for c in (0, 1, 14, 15, 16, 63)] == [15, 15, 15, 15, 17, 63]
    assert e(uniquenessRatio=-1)["uniquenessRatio"] == 10 and e(uniquenessRatio=0)["uniquenessRatio"] == 0
    assert e(disp12MaxDiff=0)["disp12MaxDiff"] == 1 and e(disp12MaxDiff=-1)["disp12MaxDiff"] == 1


def test_getters_setters_and_defaults():
    import stereo_vision_b200 as sv
    m = sv.StereoSGBM_create()
    assert (m.getMinDisparity(), m.getNumDisparities(), m.getBlockSize(), m.getP1(), m.getP2(), m.getDisp12MaxDiff(),
            m.getPreFilterCap(), m.getUniquenessRatio(), m.getSpeckleWindowSize(), m.getSpeckleRange(),
            m.getMode()) == (0, 16, 3, 0, 0, 0, 0, 0, 0, 0, 0)
    m.setNumDisparities(128); m.setMode(sv.STEREO_SGBM_MODE_HH); m.setBlockSize(7)
    assert m.getNumDisparities() == 128 and m.getMode() == 1 and m.getBlockSize() == 7
    assert (sv.INTER_LINEAR, sv.CV_32FC1, sv.STEREO_SGBM_MODE_SGBM, sv.STEREO_SGBM_MODE_HH) == (1, 5, 0, 1)


def test_geometry_and_contract(lib):
    import stereo_vision_b200 as sv
    g = sv.StereoSGBM_create(numDisparities=256, blockSize=5, P1=200, P2=800, preFilterCap=63).geometry(1920)
    assert (g["minX1"], g["maxX1"], g["width1"], g["D"], g["Dp"], g["invalid16"]) == (256, 1920, 1664, 256, 256, -16)
    assert g["max_cost"] == 25 * 189 and g["safe_packed"] == 1
    g = sv.StereoSGBM_create(minDisparity=-4, numDisparities=24, blockSize=3).geometry(200, 3)
    assert (g["minX1"], g["maxX1"], g["Dp"], g["invalid16"]) == (20, 196, 32, -80)
    with pytest.raises(ValueError):
        sv.StereoSGBM_create(numDisparities=64, blockSize=5).geometry(66)      # too narrow: cv2.error -2
    assert sv.StereoSGBM_create(numDisparities=64, mode=2).geometry(640)["width1"] == 576   # MODE_SGBM_3WAY is in the contract
    with pytest.raises(ValueError):
        sv.StereoSGBM_create(numDisparities=64, mode=9).geometry(640)
    with pytest.raises(NotImplementedError):                                   # int16 cost overflow (cv2 wraps)
        sv.StereoSGBM_create(numDisparities=64, blockSize=11, preFilterCap=63, P2=32 * 3 * 121).geometry(640, 3)


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a GPU")
    import stereo_vision_b200 as sv
    z = np.zeros((32, 128), np.uint8)
    with pytest.raises(sv.SvbError):
        sv.StereoSGBM_create(numDisparities=16).compute(z, z)
    with pytest.raises(sv.SvbError):
        sv.remap(z, z.astype(np.float32), z.astype(np.float32), sv.INTER_LINEAR)


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "stereo_vision_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "sv_oracle" not in src and "import cv2" not in src, f
