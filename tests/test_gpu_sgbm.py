"""GPU parity: StereoSGBM.compute through the C ABI vs the CPU oracle and the cv2 golden vectors (bit-exact)."""
import hashlib

import numpy as np
import pytest

import cases
import oracle
from stereo_vision_b200 import synth

pytestmark = pytest.mark.gpu


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_cost_volume_and_pass1_match_oracle(sv):
    for (H, W, D, bs, ch, minD, kind) in [(40, 150, 64, 5, 1, 0, "smooth"), (37, 120, 24, 3, 3, -4, "noise"),
                                           (33, 100, 16, 11, 1, 4, "binary"), (36, 140, 48, 7, 1, 0, "smooth")]:
        L, R = synth.stereo_pair(H, W, D, seed=3, kind=kind, channels=ch)
        kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=8 * ch * bs * bs, P2=32 * ch * bs * bs,
                  preFilterCap=63, uniquenessRatio=10, mode=1)
        raw_o, C_o, S_o = oracle.sgbm_stages(L, R, **kw)
        C_g, S_g, raw_g = sv.StereoSGBM_create(**kw).debug_stages(L, R)
        assert np.array_equal(C_g[:, :, :D], C_o), f"cost volume mismatch {H}x{W} D={D} bs={bs}"
        r = oracle.resolve(**kw)
        S4 = oracle.aggregate(C_o, r["P1"], r["P2"], 4)
        assert np.array_equal(S_g[:, :, :D], S4), f"pass-1 S mismatch {H}x{W} D={D}"
        assert np.array_equal(raw_g, raw_o), f"pre-median disparity mismatch {H}x{W} D={D}"


@pytest.mark.parametrize("latency_form", [True, False])
def test_small_sweep_matches_cv2_golden_and_oracle(sv, golden, latency_form):
    """96 cv2 golden cases through both schedules of the aggregation: the latency form (the two sweeps side by side +
    k_select_sum: what a single pair gets by default) and the back-to-back form with the fused selection."""
    sv.set_latency_form(latency_form)
    try:
        bad = []
        for i, c in enumerate(cases.sgbm_small_cases()):
            L, R = cases.sgbm_inputs(c)
            got = sv.StereoSGBM_create(**c["kw"]).compute(L, R)
            assert got.dtype == np.int16 and got.shape == L.shape[:2]
            if not np.array_equal(got, golden[f"sgbm_{i}"]):
                bad.append((i, c, int((got != golden[f"sgbm_{i}"]).sum())))
        assert not bad, f"{len(bad)} of 96 cases differ from cv2 golden (latency_form={latency_form}): {bad[:3]}"
    finally:
        sv.set_latency_form(True)


@pytest.mark.parametrize("latency_form", [True, False])
@pytest.mark.parametrize("name", ["c1", "c2", "c3", "c3_sgbm"])
def test_full_size_matches_cv2_golden_hash(sv, golden_full, name, latency_form):
    L, R, kw = cases.full_inputs(name)
    g = golden_full["full"][name]
    assert [_sha(L), _sha(R)] == g["in_sha256"], "synthetic inputs are not reproducible on this host"
    sv.set_latency_form(latency_form)
    try:
        got = sv.StereoSGBM_create(**kw).compute(L, R)
    finally:
        sv.set_latency_form(True)
    if _sha(got) != g["sha256"]:
        ref = oracle.sgbm_compute(L, R, **kw)
        raise AssertionError(f"{name} (latency_form={latency_form}): differs from cv2 golden; vs oracle: {(got != ref).sum()} px differ")


def test_full_size_c5_4k(sv, golden_full):
    L, R, kw = cases.full_inputs("c5")
    got = sv.StereoSGBM_create(**kw).compute(L, R)
    assert _sha(got) == golden_full["full"]["c5"]["sha256"]


def test_batch_equals_single_and_torch_zero_copy(sv):
    import torch
    D, H, W = 64, 96, 320
    kw = synth.sgbm_kwargs(D, 5, 1)
    pairs = [synth.stereo_pair(H, W, D, seed=1000 + i) for i in range(5)]
    Lb = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    Rb = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    m = sv.StereoSGBM_create(**kw)
    out = m.compute_batch(Lb, Rb)
    assert out.is_cuda and out.dtype == torch.int16
    for i, (L, R) in enumerate(pairs):
        assert np.array_equal(out[i].cpu().numpy(), oracle.sgbm_compute(L, R, **kw))
        assert np.array_equal(m.compute(L, R), out[i].cpu().numpy())


def test_properties_at_full_size(sv):
    """Size-independent properties at 1080p/D=256: determinism, invalid left band, value range."""
    L, R, kw = cases.full_inputs("c3")
    m = sv.StereoSGBM_create(**kw)
    a = m.compute(L, R)
    b = m.compute(L, R)
    assert np.array_equal(a, b)
    D = kw["numDisparities"]
    assert (a[:, :D] == -16).all()
    v = a[a != -16]
    assert v.min() >= 0 and v.max() <= (D - 1) * 16
    # the synthetic field is recovered on most valid pixels
    d_true = synth.disparity_field(L.shape[0], L.shape[1], D)
    ok = (np.abs(a / 16.0 - d_true) <= 1)[a != -16].mean()
    assert ok > 0.9


def test_error_behaviour(sv):
    m = sv.StereoSGBM_create(numDisparities=64, blockSize=5)
    z = np.zeros((32, 128), np.uint8)
    with pytest.raises(ValueError):
        m.compute(z, np.zeros((32, 127), np.uint8))
    with pytest.raises(ValueError):
        m.compute(z.astype(np.float32), z.astype(np.float32))
    with pytest.raises(ValueError):
        m.compute(z[:, :60], z[:, :60])          # too narrow (cv2.error -2)
    with pytest.raises(NotImplementedError):       # MODE_SGBM_3WAY on an image shorter than its stripe overlap
        sv.StereoSGBM_create(numDisparities=16, blockSize=9, mode=2).compute(z[:12], z[:12])
    with pytest.raises(ValueError):
        sv.StereoSGBM_create(numDisparities=16, mode=7).compute(z, z)
    with pytest.raises(ValueError):
        sv.StereoSGBM_create(numDisparities=0).compute(z, z)
    # non-contiguous views are accepted
    big = np.random.default_rng(0).integers(0, 255, (64, 256), dtype=np.uint8)
    v = big[::2, ::1]
    assert np.array_equal(m.compute(v, v), oracle.sgbm_compute(np.ascontiguousarray(v), np.ascontiguousarray(v),
                                                                numDisparities=64, blockSize=5))


@pytest.mark.parametrize("name,B", [("c3", 6), ("c5", 2)])
def test_full_size_batch_is_race_free(sv, golden_full, name, B):
    """A batch of identical full-size pairs keeps many bands of several frames in flight at once (cross-band hand-off,
    ticket scheduling, edge prefetch): every frame must still hash to the cv2 golden, twice in a row."""
    import torch
    L, R, kw = cases.full_inputs(name)
    Lb = torch.from_numpy(L).cuda()[None].repeat(B, 1, 1).contiguous()
    Rb = torch.from_numpy(R).cuda()[None].repeat(B, 1, 1).contiguous()
    m = sv.StereoSGBM_create(**kw)
    for _ in range(2):
        out = m.compute_batch(Lb, Rb).cpu().numpy()
        for b in range(B):
            assert _sha(out[b]) == golden_full["full"][name]["sha256"], f"{name}: frame {b} of the batch differs"
    sv.release_workspaces()


def test_mode_hh4_matches_cv2_golden(sv):
    """MODE_HH4 (mode 3): small sweep bit-exact vs cv2 golden vectors, c3 at full size vs the cv2 hash, batched."""
    import json
    import os
    import torch
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_hh4.npz"))
    bad = []
    for i, c in enumerate(cases.sgbm_hh4_cases()):
        L, R = cases.sgbm_inputs(c)
        got = sv.StereoSGBM_create(**c["kw"]).compute(L, R)
        if not np.array_equal(got, g[f"hh4_{i}"]):
            bad.append((i, c, int((got != g[f"hh4_{i}"]).sum())))
    assert not bad, f"{len(bad)} of 40 MODE_HH4 cases differ from cv2 golden: {bad[:3]}"
    with open(os.path.join(os.path.dirname(__file__), "golden", "golden_hh4_full.json")) as f:
        full = json.load(f)
    L, R, kw = cases.full_inputs("c3")
    m = sv.StereoSGBM_create(**dict(kw, mode=sv.STEREO_SGBM_MODE_HH4))
    assert _sha(m.compute(L, R)) == full["c3_hh4"]
    Lb = torch.from_numpy(np.stack([L] * 3)).cuda()
    Rb = torch.from_numpy(np.stack([R] * 3)).cuda()
    out = m.compute_batch(Lb, Rb).cpu().numpy()
    assert all(_sha(out[b]) == full["c3_hh4"] for b in range(3))


def test_host_pipeline_matches_compute_batch(sv):
    """HostPipeline (copies of neighbouring batches overlapped with the matcher) == compute_batch, batch by batch."""
    import torch
    H, W, D = 96, 320, 64
    kw = dict(minDisparity=0, numDisparities=D, blockSize=5, P1=200, P2=800, disp12MaxDiff=1, uniquenessRatio=10,
              speckleWindowSize=50, speckleRange=2, mode=1)
    m = sv.StereoSGBM_create(**kw)
    pipe = sv.HostPipeline(m)
    Q = np.array([[1, 0, 0, -W / 2], [0, 1, 0, -H / 2], [0, 0, 0, 500.0], [0, 0, 1 / 60.0, 0]])
    batches, outs, meds = [], [], []
    boxes = synth.boxes_for_frame(H, W, D, n=4, seed=1)
    boxes_b = np.concatenate([boxes, boxes]); bi = np.repeat(np.arange(2, dtype=np.int32), 4)
    for i in range(5):
        pairs = [synth.stereo_pair(H, W, D, seed=100 + 2 * i + j) for j in range(2)]
        L = torch.from_numpy(np.stack([p[0] for p in pairs])).pin_memory()
        R = torch.from_numpy(np.stack([p[1] for p in pairs])).pin_memory()
        batches.append((L, R))
        outs.append(torch.empty((2, H, W), dtype=torch.int16).pin_memory())
        meds.append(torch.empty(8, dtype=torch.float32).pin_memory())
    for (L, R), o, md in zip(batches, outs, meds):
        pipe.submit(L, R, o, Q=Q, boxes=boxes_b, box_image=bi, med_h=md)
    pipe.synchronize()
    for (L, R), o, md in zip(batches, outs, meds):
        ref = m.compute_batch(L.cuda(), R.cuda())
        assert torch.equal(o, ref.cpu())
        refm = sv.box_median_depth(ref, Q, boxes_b, box_image=bi).cpu().numpy()
        assert np.array_equal(md.numpy(), refm, equal_nan=True)


def test_geometry_fuzz_vs_oracle(sv):
    """Cost-tile / sweep geometry corners: every disparity-group count (D = 16 .. 512), widths around tile multiples,
    all block sizes, 1 and 3 channels, negative / positive minDisparity — GPU vs the CPU oracle, bit-exact."""
    rng = np.random.default_rng(77)
    cfgs = [(16, 1), (32, 3), (48, 5), (64, 7), (96, 9), (128, 11), (144, 5), (208, 3), (256, 5), (272, 5), (384, 3), (512, 5)]
    bad = []
    for D, bs in cfgs:
        for rep in range(2):
            ch = 3 if (rep == 1 and D <= 256) else 1
            minD = int(rng.choice([0, -16, 8]))
            mode = int(rng.choice([0, 1, 2, 3]))
            H = int(rng.integers(17, 40)) if mode != 2 else int(rng.integers(40, 75))   # (3WAY: stripes >= their overlap)
            W = D + abs(minD) + int(rng.choice([9, 23, 61, 64, 65, 127, 130]))
            kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=8 * ch * bs * bs, P2=32 * ch * bs * bs,
                      disp12MaxDiff=1, preFilterCap=31, uniquenessRatio=10, speckleWindowSize=20, speckleRange=2, mode=mode)
            if bs * bs * ch * (2 * 31 + 63) + kw["P2"] > 32767:
                continue
            L, R = synth.stereo_pair(H, W, D, seed=int(rng.integers(1 << 30)), channels=ch)
            got = sv.StereoSGBM_create(**kw).compute(L, R)
            ref = oracle.sgbm_compute(L, R, **kw)
            if not np.array_equal(got, ref):
                bad.append((D, bs, ch, minD, H, W, mode, int((got != ref).sum())))
    assert not bad, bad


def test_mode_sgbm_3way_matches_cv2_golden(sv):
    """MODE_SGBM_3WAY (mode 2): 48 small cases bit-exact vs cv2 golden vectors, c1/c2/c3 at full size vs the cv2 hashes,
    and a batch (the stripes of all frames run together)."""
    import json
    import os
    import torch
    gd = os.path.join(os.path.dirname(__file__), "golden")
    g = np.load(os.path.join(gd, "golden_3way.npz"))
    bad = []
    for i, c in enumerate(cases.sgbm_3way_cases()):
        L, R = cases.sgbm_inputs(c)
        got = sv.StereoSGBM_create(**c["kw"]).compute(L, R)
        if not np.array_equal(got, g[f"w3_{i}"]):
            bad.append((i, c, int((got != g[f"w3_{i}"]).sum())))
    assert not bad, f"{len(bad)} of 48 MODE_SGBM_3WAY cases differ from cv2 golden: {bad[:3]}"
    with open(os.path.join(gd, "golden_3way_full.json")) as f:
        full = json.load(f)
    for name in ("c1", "c2", "c3"):
        L, R, kw = cases.full_inputs(name)
        m = sv.StereoSGBM_create(**dict(kw, mode=sv.STEREO_SGBM_MODE_SGBM_3WAY))
        got = m.compute(L, R)
        if _sha(got) != full[f"{name}_3way"]:
            ref = oracle.sgbm_compute(L, R, **dict(kw, mode=2))
            raise AssertionError(f"{name} 3way: differs from cv2 golden; vs oracle: {(got != ref).sum()} px differ")
    Lb = torch.from_numpy(np.stack([L] * 3)).cuda()
    Rb = torch.from_numpy(np.stack([R] * 3)).cuda()
    out = m.compute_batch(Lb, Rb).cpu().numpy()
    assert all(_sha(out[b]) == full["c3_3way"] for b in range(3))


@pytest.fixture(scope="module")
def golden_r2():
    import json
    import os
    gd = os.path.join(os.path.dirname(__file__), "golden")
    with open(os.path.join(gd, "golden_r2.json")) as f:
        meta = json.load(f)
    return meta, np.load(os.path.join(gd, "golden_r2.npz"))


def test_c3_three_channel_and_negative_min_disparity_full_size(sv, golden_r2):
    """BASELINE-size holes of round 1: 3-channel input (6-plane cost path + saturating sweep) and minDisparity = -16."""
    meta, _ = golden_r2
    H, W, D = 1080, 1920, 256
    L3, R3 = synth.stereo_pair(H, W, D, seed=1000, channels=3)
    c = meta["c3_3ch"]
    assert [_sha(L3), _sha(R3)] == c["in_sha256"]
    assert _sha(sv.StereoSGBM_create(**c["kw"]).compute(L3, R3)) == c["sha256"]
    L, R = synth.stereo_pair(H, W, D, seed=1000)
    c = meta["c3_minD-16"]
    assert c["kw"]["minDisparity"] == -16
    assert _sha(sv.StereoSGBM_create(**c["kw"]).compute(L, R)) == c["sha256"]


def test_c3_batch_of_32_distinct_frames_matches_cv2_hashes(sv, golden_r2):
    """The bench regime: 32 DISTINCT 1080p pairs in one batch (72 bands x 32 frames >> resident CTAs), every output
    checked against cv2's hash of that frame."""
    import torch
    meta, _ = golden_r2
    c = meta["c3_batch"]
    s0, s1 = c["seeds"]
    H, W, D = 1080, 1920, 256
    pairs = [synth.stereo_pair(H, W, D, seed=s) for s in range(s0, s1)]
    Lb = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    Rb = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    out = sv.StereoSGBM_create(**c["kw"]).compute_batch(Lb, Rb).cpu().numpy()
    bad = [i for i in range(s1 - s0) if _sha(out[i]) != c["sha256"][i]]
    assert not bad, f"frames {bad} of the 32-frame batch differ from cv2"
    sv.release_workspaces()


def test_workspace_guards_stay_intact(sv):
    """Out-of-bounds WRITE check of our own (compute-sanitizer is closed on this pool): the caller-owned workspace and
    the output are embedded in larger buffers filled with a pattern; after compute the guard bytes must be untouched and
    the result must equal the oracle's.  Runs the single-pair latency form and a batch in the back-to-back form."""
    import ctypes as C
    import torch
    from stereo_vision_b200 import _lib
    lib, ctx = _lib.load(), _lib.context(0)
    GUARD = 1 << 20
    for (H, W, D, B, mode) in [(45, 400, 256, 1, 1), (45, 400, 256, 12, 1), (64, 300, 96, 3, 0), (70, 330, 128, 2, 2)]:
        kw = synth.sgbm_kwargs(D, 5, mode)
        pairs = [synth.stereo_pair(H, W, D, seed=50 + i) for i in range(B)]
        Lb = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
        Rb = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
        prm = _lib.SgbmParams(*[kw[k] for k in ("minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff",
                                                "preFilterCap", "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")])
        need = C.c_size_t()
        assert lib.svb_sgbm_workspace_bytes(C.byref(prm), H, W, 1, B, C.byref(need)) == 0
        ws = torch.full((need.value + 2 * GUARD,), 0xA5, dtype=torch.uint8, device="cuda")
        out = torch.full((B * H * W + 2 * GUARD,), 0x5A5A, dtype=torch.int16, device="cuda")
        rc = lib.svb_sgbm_compute(ctx, C.byref(prm), Lb.data_ptr(), Rb.data_ptr(), H, W, 1, W, H * W, B,
                                  out.data_ptr() + 2 * GUARD, W, H * W, ws.data_ptr() + GUARD, need.value,
                                  torch.cuda.current_stream().cuda_stream)
        assert rc == 0, lib.svb_last_error()
        torch.cuda.synchronize()
        assert bool((ws[:GUARD] == 0xA5).all()) and bool((ws[GUARD + need.value:] == 0xA5).all()), (H, W, D, B, mode)
        assert bool((out[:GUARD] == 0x5A5A).all()) and bool((out[GUARD + B * H * W:] == 0x5A5A).all()), (H, W, D, B, mode)
        got = out[GUARD:GUARD + B * H * W].reshape(B, H, W).cpu().numpy()
        for i, (l, r) in enumerate(pairs):
            assert np.array_equal(got[i], oracle.sgbm_compute(l, r, **kw)), (H, W, D, B, mode, i)


def test_two_devices_in_one_process(sv):
    """Several GPUs driven from one process ('8 threads with 8 contexts'): the per-device kernel
    attributes (dynamic shared memory opt-in) must be set on every device.  Skipped on single-GPU boxes."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    L, R, kw = cases.full_inputs("c1")
    ref = None
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            Ld, Rd = torch.from_numpy(L).to(f"cuda:{dev}"), torch.from_numpy(R).to(f"cuda:{dev}")
            got = sv.StereoSGBM_create(**kw).compute(Ld, Rd)
            assert got.device.index == dev
            got = got.cpu().numpy()
        ref = got if ref is None else ref
        assert np.array_equal(got, ref)
    D, H, W = 256, 60, 420
    kw = synth.sgbm_kwargs(D, 5, 1)
    l, r = synth.stereo_pair(H, W, D, seed=3)
    want = oracle.sgbm_compute(l, r, **kw)
    for dev in (1, 0):
        with torch.cuda.device(dev):
            assert np.array_equal(sv.StereoSGBM_create(**kw).compute(torch.from_numpy(l).to(f"cuda:{dev}"),
                                                                     torch.from_numpy(r).to(f"cuda:{dev}")).cpu().numpy(), want)
