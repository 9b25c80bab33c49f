"""rs_redo_map_static, the entry point for hosts that cannot let this library allocate device memory.

Some users of the library have no CUDA heap to draw from at the moment they need the redo map of the passes loop:
a host application whose own pool has reserved the whole card, a frame loop that cannot afford the device-wide stall
a cudaMalloc causes, or multi-GPU code between rs_signal and rs_wait_counter, where a cudaMalloc waits for peers that
are spinning.  For them the library keeps the intermediate noise map in a fixed 16 MiB array of its own
(RS_STATIC_MAX_PIXELS pixels) instead of a context buffer.

The test drives the new entry point through the C ABI (ctypes) and compares it, bit for bit, with what the existing
entry points give on the same pixels: rs_noise_map followed by the threshold, and the redo count rs_render_passes
reports.  It does so on two calls in a row with different inputs, because the array is shared and keeps the previous
call's contents: a second call that saw stale data would show up here.

This file is named to be collected last: `pytest -x` stops at the first failure, and a failure here must not hide
the parity tests.  It passes on its own in a fresh process (the first call then does the per-device address look-up)
and after the rest of the suite (the look-up may or may not have happened; the array holds whatever came before)."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_scene

pytestmark = pytest.mark.gpu

RS_ERR_UNSUPPORTED = 5
RS_STATIC_MAX_PIXELS = 4194304


def _redo_static(renderer, d_pixels, w, h, quirk, thr):
    """One rs_redo_map_static call on torch-owned buffers (inputs and outputs are the caller's; the call itself allocates nothing)."""
    import torch
    d_mask = torch.full((h, w), 7, dtype=torch.uint8, device="cuda")          # poisoned: every byte must be rewritten
    d_count = torch.full((1,), 123456789, dtype=torch.int64, device="cuda")   # poisoned: the call has to zero it itself
    torch.cuda.synchronize()
    rc = renderer._lib.rs_redo_map_static(renderer._ctx, d_pixels.data_ptr(), d_mask.data_ptr(), d_count.data_ptr(), w, h, quirk,
                                          C.c_float(thr), None)
    assert rc == 0, renderer._lib.rs_last_error(renderer._ctx).decode()
    torch.cuda.synchronize()
    return d_mask.cpu().numpy(), int(d_count.item())


def _redo_existing(renderer, d_pixels, w, h, quirk, thr):
    """The same result from the entry point that was there before: rs_noise_map into a caller buffer, threshold on the host."""
    import torch
    d_noise = torch.empty((h, w), dtype=torch.float32, device="cuda")
    renderer.noise_map(d_pixels.data_ptr(), d_noise.data_ptr(), w, h, quirk)
    torch.cuda.synchronize()
    mask = (d_noise.cpu().numpy() >= np.float32(thr)).astype(np.uint8)
    return mask, int(mask.sum())


def test_redo_map_static_equals_noise_map_plus_threshold_twice_in_a_row(renderer):
    import torch
    rng = np.random.default_rng(11)
    # two calls in a row; the second frame is SMALLER and different, so stale noise of the first call would be seen
    cases = [(211, 97, 1, 0.35), (64, 40, 0, 0.2), (211, 97, 0, 0.35)]
    for w, h, quirk, thr in cases:
        px = rng.random((h, w, 4), dtype=np.float32)
        px[h // 3:h // 2] = px[h // 3, 0]  # a flat band: noise exactly 0 there, below any threshold
        d_px = torch.from_numpy(px).cuda()
        exp_mask, exp_count = _redo_existing(renderer, d_px, w, h, quirk, thr)
        for _ in range(2):  # and the very same call twice: identical answers, nothing accumulates across calls
            mask, count = _redo_static(renderer, d_px, w, h, quirk, thr)
            assert np.array_equal(mask, exp_mask)
            assert count == exp_count
        assert 0 < exp_count < w * h


def test_redo_map_static_matches_the_count_of_rs_render_passes(renderer):
    """One pass of rs_render_passes reports how many pixels its own (allocated) noise map puts over the threshold; the
    static entry point, fed the image that pass returned, has to arrive at the same number."""
    import torch
    from raysnail_b200 import make_params
    from raysnail_b200._abi import RS_REDO_NOISE, RS_REDO_NOISE_FIXED
    scene = load_scene("rtow_final")
    renderer.upload(scene)
    w, h = 96, 60
    cam = scene.camera(w, h)
    params = make_params(samples=4, seed=5)
    for mode, quirk in ((RS_REDO_NOISE, 1), (RS_REDO_NOISE_FIXED, 0)):
        img, counts = renderer.render_passes(cam, params, 1, mode, 0.01)
        d_px = torch.from_numpy(np.ascontiguousarray(img, dtype=np.float32)).cuda()
        mask, count = _redo_static(renderer, d_px, w, h, quirk, 0.01)
        assert count == int(counts[0]) == int(mask.sum())


def test_redo_map_static_refuses_frames_beyond_its_fixed_scratch(renderer):
    import torch
    d = torch.zeros(16, dtype=torch.float32, device="cuda")  # never touched: the size check comes first
    w, h = 2049, 2048
    assert w * h > RS_STATIC_MAX_PIXELS
    rc = renderer._lib.rs_redo_map_static(renderer._ctx, d.data_ptr(), d.data_ptr(), d.data_ptr(), w, h, 1, C.c_float(0.01), None)
    assert rc == RS_ERR_UNSUPPORTED
    msg = renderer._lib.rs_last_error(renderer._ctx).decode()
    assert "RS_STATIC_MAX_PIXELS" in msg and str(w * h) in msg
    # the largest admitted frame is accepted (2048 x 2048)
    w = h = 2048
    px = torch.rand((h, w, 4), dtype=torch.float32, device="cuda")
    exp_mask, exp_count = _redo_existing(renderer, px, w, h, 0, 0.5)
    mask, count = _redo_static(renderer, px, w, h, 0, 0.5)
    assert np.array_equal(mask, exp_mask) and count == exp_count
