"""svb_filter_speckles_s16_noalloc: the no-allocation variant of the speckle filter (static device scratch inside
libsvb.so; no counterpart in the reference).  Called through the C ABI
and compared bit for bit (integer work: tolerance 0) with the existing entry point svb_filter_speckles_s16, on two
calls in a row that share the static scratch.  Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NEWVAL, MAXSIZE, MAXDIFF = -16, 100, 32


def _disparity_like(H, W, seed):
    rng = np.random.default_rng(seed)
    a = np.kron(rng.integers(0, 30, ((H + 23) // 24, (W + 23) // 24)), np.ones((24, 24)))[:H, :W].astype(np.int16) * 16
    a += rng.integers(-6, 7, (H, W)).astype(np.int16)
    a[rng.random((H, W)) < 0.08] = NEWVAL
    a[H // 5:H // 2:3, W // 6:W // 2] = NEWVAL       # thin stripes: many small components next to a large one
    return a


def _existing(lib, ctx, t, H, W, B):
    import torch
    labels = torch.empty(2 * H * W * B, dtype=torch.int32, device=t.device)
    rc = lib.svb_filter_speckles_s16(ctx, t.data_ptr(), H, W, W, B, H * W, NEWVAL, MAXSIZE, MAXDIFF, labels.data_ptr(),
                                     labels.numel(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.svb_last_error()
    torch.cuda.synchronize()
    return t


def _noalloc(lib, ctx, t, H, W, B):
    import torch
    rc = lib.svb_filter_speckles_s16_noalloc(ctx, t.data_ptr(), H, W, W, B, H * W, NEWVAL, MAXSIZE, MAXDIFF,
                                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.svb_last_error()
    torch.cuda.synchronize()
    return t


def test_noalloc_speckle_matches_existing_entry_point_twice(sv):
    import torch
    from stereo_vision_b200 import _lib
    lib, ctx = _lib.load(), _lib.context(0)
    dev = torch.device("cuda", 0)
    # two calls in a row with different shapes: the second sees the first call's labels / sizes in the shared scratch
    cases = [(1080, 1920, 1, 11), (360, 500, 3, 12), (1080, 1920, 1, 13)]
    for H, W, B, seed in cases:
        a = np.stack([_disparity_like(H, W, seed + 100 * b) for b in range(B)])
        want = _existing(lib, ctx, torch.from_numpy(a.copy()).to(dev), H, W, B).cpu().numpy()
        inp = torch.from_numpy(a.copy()).to(dev)
        before = torch.cuda.memory_allocated(dev)
        got = _noalloc(lib, ctx, inp, H, W, B).cpu().numpy()
        assert torch.cuda.memory_allocated(dev) == before          # nothing taken from torch's allocator either
        assert np.array_equal(got, want), (H, W, B)
        assert (got != a).any()                                    # the filter did remove something
        # and once more on the filtered map: idempotent, like the existing entry point
        again = _noalloc(lib, ctx, torch.from_numpy(got.copy()).to(dev), H, W, B).cpu().numpy()
        assert np.array_equal(again, _existing(lib, ctx, torch.from_numpy(got.copy()).to(dev), H, W, B).cpu().numpy())


def test_noalloc_speckle_rejects_sizes_beyond_the_static_scratch(sv):
    import torch
    from stereo_vision_b200 import _lib
    lib, ctx = _lib.load(), _lib.context(0)
    H, W = 2160, 3840                                              # 8.3 Mpx > SVB_NOALLOC_SPECKLE_MAX_PIXELS (2^21)
    t = torch.zeros((H, W), dtype=torch.int16, device="cuda:0")
    rc = lib.svb_filter_speckles_s16_noalloc(ctx, t.data_ptr(), H, W, W, 1, H * W, NEWVAL, MAXSIZE, MAXDIFF, None)
    assert rc == -4                                                # SVB_ERR_WORKSPACE
    msg = lib.svb_last_error().decode()
    assert "exceeds the static scratch" in msg and "SVB_NOALLOC_SPECKLE_MAX_PIXELS" in msg
    # 2 x 1080p in one call is over the limit too (the limit is on B*H*W)
    t2 = torch.zeros((2, 1080, 1920), dtype=torch.int16, device="cuda:0")
    assert lib.svb_filter_speckles_s16_noalloc(ctx, t2.data_ptr(), 1080, 1920, 1920, 2, 1080 * 1920, NEWVAL, MAXSIZE,
                                               MAXDIFF, None) != 0
