"""sphb_bbox_noalloc: the one entry point that allocates nothing (its scratch is a fixed __device__ array
inside libsphb200.so), for callers that may not allocate GPU memory when they need the box -- a render or
real-time thread that cannot afford the device-wide synchronisation of cudaMalloc/cudaFree, or a plug-in
living under its host application's memory budget.

Checked through the C ABI against the engine's own sphb_get_bbox (exact equality: min/max involve no
rounding, the same bar as test_grid_default_scene_bbox_mode), on two calls in a row that share the scratch:
a large cloud first, a small one second, so that anything the first call left behind would show.
This file sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

import sph_b200
from sph_b200 import capi, scenes

pytestmark = pytest.mark.gpu


def noalloc_bbox(lib, eng, out):
    """Box of the engine's current (device-resident) positions through the allocation-free entry point."""
    posm = eng.device_state()[0]
    rc = lib.sphb_bbox_noalloc(C.c_void_p(posm), eng.count(), C.c_void_p(eng.stream()), C.c_void_p(out.data_ptr()))
    assert rc == 0, lib.sphb_last_error().decode()
    eng.synchronize()           # same stream as the engine's own kernels: the result is complete now
    return out.cpu().numpy()


def test_bbox_noalloc_matches_engine_bbox_twice_in_a_row():
    import torch
    torch.cuda.set_device(0)    # the entry point works on the CURRENT device
    lib = capi.load_library()
    out = torch.full((6,), float("nan"), device="cuda:0", dtype=torch.float32)   # the caller's output buffer

    big, walls = scenes.dam_break(32, 64, 32, stressed=True)       # 65,536 points -> 64 partial results
    small, _ = scenes.dam_break(7, 9, 11, stressed=True)           # 693 points -> 1 partial result
    small["pos"] += np.float32(3.0)                                 # a box that lies inside none of the big one's partials
    small["pos"][5, 1] = np.nan                                     # NaN coordinates are skipped, as in sphb_get_bbox

    with sph_b200.Engine() as e1, sph_b200.Engine() as e2:
        e1.set_walls(walls)
        e1.upload(big)
        e1.step(2)                                                  # positions as the engine holds them (sorted order)
        e2.upload(small)
        first = noalloc_bbox(lib, e1, out)                          # call 1: fills 64 partial slots of the shared scratch
        second = noalloc_bbox(lib, e2, out)                         # call 2 right after it: must not see call 1's leftovers
        assert np.array_equal(first, e1.get_bbox())
        assert np.array_equal(second, e2.get_bbox())
        assert not np.array_equal(first, second)
        third = noalloc_bbox(lib, e1, out)                          # and back again
        assert np.array_equal(third, first)

        # beyond the fixed scratch: refused with a clear message, nothing launched
        rc = lib.sphb_bbox_noalloc(C.c_void_p(e1.device_state()[0]), capi.NOALLOC_MAX_POINTS + 1, None, C.c_void_p(out.data_ptr()))
        assert rc == 4 and b"exceed the fixed scratch" in lib.sphb_last_error()
        rc = lib.sphb_bbox_noalloc(C.c_void_p(e1.device_state()[0]), 0, None, C.c_void_p(out.data_ptr()))
        assert rc == 1
