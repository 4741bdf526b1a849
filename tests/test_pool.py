"""StitchPool (GPU): one process, one worker thread + ctx per GPU, a shared job queue, no collective.

SURVEY.md 4 "Multi-GPU" pin: every job's output equals the single-GPU output -- bytes for host outputs, the device
digest for canvases that stay on their GPU.  On a one-GPU box the pool is built with the same device listed twice (two
workers, two contexts in one process), which is also the regression test for the per-device kernel attributes
(a second context never got the shared-memory opt-in)."""
import ctypes as C

import numpy as np
import pytest

from tests import util as U


def _devices():
    import torch
    n = torch.cuda.device_count()
    return list(range(n)) if n >= 2 else [0, 0]


def test_digest_twin_matches_definition():
    a = np.arange(1000, dtype=np.uint8)
    assert U.digest(a) != U.digest(a[::-1].copy()) and U.digest(a[:999]) != U.digest(a)
    assert U.digest(b"") == U.digest(np.zeros(0, np.uint8))


@pytest.mark.gpu
def test_pool_outputs_equal_single_gpu_outputs():
    from imagestitching_b200 import StitchContext, StitchPool
    jobs = []
    rng = np.random.default_rng(5)
    for k in range(12):
        n = int(rng.integers(1, 5))
        dims = [(int(rng.integers(300, 900)), int(rng.integers(200, 700))) for _ in range(n)]
        bitmaps = [U.synth_image(w, h, 100 * k + i, 0) for i, (w, h) in enumerate(dims)]
        jobs.append((U.page_images(dims, bitmaps), "vertical" if k % 2 else "horizontal",
                     dict(mode=("min", "max", "original")[k % 3], gap=k % 4, platform="ios", deviceMaxCanvasSize=900,
                          output=("rgba", "yuv420", "png", "jpeg")[k % 4])))
    ctx = StitchContext(0)
    want = [ctx.run(imgs, d, **kw) for imgs, d, kw in jobs]
    ctx.close()
    pool = StitchPool(_devices(), lanes_per_device=2)
    assert pool.device_count == 2 or pool.device_count == len(_devices())
    handles = [pool.submit(imgs, d, **kw) for imgs, d, kw in jobs]       # host buffers: any GPU
    outs = [pool.wait(h) for h in handles]
    ran = [pool.jobs_run(i) for i in range(pool.device_count)]
    pool.close()
    assert sum(ran) == len(jobs) and {o.device for o in outs} <= set(range(len(ran)))
    for got, ref in zip(outs, want):
        assert (got.width, got.height) == (ref.width, ref.height)
        for f in ("rgba", "y", "cb", "cr"):
            a, b = getattr(got, f), getattr(ref, f)
            assert (a is None) == (b is None) and (a is None or np.array_equal(a, b)), f
        assert got.png == ref.png and got.jpeg == ref.jpeg


@pytest.mark.gpu
def test_pool_device_resident_jobs_and_digests():
    """Device-resident bitmaps and canvases on EVERY worker's GPU: per-job digests are identical on all of them and equal
    the CPU digest of the canvas a single context produces."""
    from imagestitching_b200 import DeviceBitmap, StitchContext, StitchPool
    from imagestitching_b200 import _lib as L
    from imagestitching_b200.stitch import _profile_struct
    dims = [(808, 604), (604, 806), (808, 604)]   # 4*w is a multiple of 16: packed device bitmaps take the same units as staged ones
    bitmaps = [U.synth_image(w, h, 31 + i, 0) for i, (w, h) in enumerate(dims)]
    ctx = StitchContext(0)
    ref = ctx.run(U.page_images(dims, bitmaps), "vertical", platform="ios", deviceMaxCanvasSize=700)
    ctx.close()
    want = U.digest(ref.rgba)
    pool = StitchPool(_devices(), lanes_per_device=2)
    lib = pool.lib
    nd = pool.device_count
    prof = _profile_struct("ios", 700, None)
    helper = pool._helper
    keep, digests, tickets = [], [], []
    res = [L.StitchResult() for _ in range(3 * nd)]
    for d in range(nd):
        h = pool.ctx_handle(d)
        dev = []
        for a in bitmaps:
            p = C.c_void_p()
            assert lib.stitch_device_alloc(h, a.nbytes, C.byref(p)) == 0
            assert lib.stitch_memcpy_h2d(h, p, a.ctypes.data, a.nbytes) == 0
            dev.append(DeviceBitmap(p.value, a.shape[1], a.shape[0], a.shape[1] * 4, keepalive=p))
        for r in range(3):
            out = C.c_void_p()
            assert lib.stitch_device_alloc(h, ref.rgba.nbytes, C.byref(out)) == 0
            job, k2 = helper.build_job(U.page_images(dims, dev), "vertical", "min", 0, prof, "rgba", True, out.value, 0,
                                       ref.rgba.nbytes, 0)
            t = C.c_uint64()
            assert lib.stitch_pool_submit(pool.handle, C.byref(job), C.byref(res[3 * d + r]), d, C.byref(t)) == 0
            dg, t2 = C.c_uint64(0), C.c_uint64()
            assert lib.stitch_pool_digest(pool.handle, d, out, ref.rgba.nbytes, C.byref(dg), C.byref(t2)) == 0
            keep += [job, k2, out, dev]
            digests.append(dg)
            tickets += [t.value, t2.value]
    # a device-memory job that does not say where it lives is refused
    bad = C.c_uint64()
    assert lib.stitch_pool_submit(pool.handle, C.byref(keep[0]), C.byref(res[0]), -1, C.byref(bad)) == L.E_BAD_ARG
    for t in tickets:
        assert lib.stitch_pool_wait(pool.handle, t, None) == 0, lib.stitch_pool_last_error(pool.handle)
    assert [pool.jobs_run(i) for i in range(nd)] == [6] * nd        # 3 jobs + 3 digests on each worker
    pool.close()
    assert [int(d.value) for d in digests] == [want] * (3 * nd)
