"""Progressive JPEG inputs (SOF2; P:4 SUPPORTED_IMAGE_TYPES 'jpg'/'jpeg', loadImageFrom utils/canvas.js:27-121).

CPU: the oracle decoder (oracle/jpeg_dec_ref.c, T.81 Annex G) is pinned byte-identical to libjpeg-turbo as shipped in
     PIL; the sequential scan decoder a GPU thread runs for refinement scans (imagestitching_b200/csrc/jpeg_prog_core.h)
     is driven on the host by tests/native/jprog_host.cpp and held to the oracle's coefficients; the header parser's scan
     list is checked.
GPU: stitch_decode_jpeg and STITCH_MEM_HOST_JPEG job inputs are byte-identical to the oracle decoder, through the
     self-synchronising path for first scans and through the one-thread path for everything (STITCH_JPROG_SEQ=1).
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import jpeg_c as J
from tests import util as U
from tests.test_jpeg_decode import _corrupt, _jpeg, _pil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ZZ = [0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
      35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
      62, 63]
SIZES = [(1, 1), (7, 5), (16, 16), (17, 33), (123, 77), (250, 131), (640, 481)]


@pytest.mark.parametrize("size", SIZES, ids=[f"{w}x{h}" for w, h in SIZES])
@pytest.mark.parametrize("subsampling", [0, 1, 2])
def test_oracle_progressive_equals_libjpeg(size, subsampling):
    for q in (30, 90, 100):
        for kw in ({}, {"restart_marker_blocks": 3}, {"restart_marker_rows": 1}):
            data = _jpeg(*size, seed=q, quality=q, subsampling=subsampling, progressive=True, **kw)
            assert b"\xff\xc2" in data
            got = J.decode(data)
            assert (got[..., 3] == 255).all()
            assert np.array_equal(got[..., :3], _pil(data)), f"{size} ss={subsampling} q={q} {kw}"


def test_oracle_progressive_gray_and_large():
    for size in SIZES:
        data = _jpeg(*size, gray=True, quality=85, progressive=True)
        assert np.array_equal(J.decode(data)[..., :3], _pil(data))
    data = _jpeg(1600, 1200, seed=2, quality=92, subsampling=2, progressive=True)
    assert np.array_equal(J.decode(data)[..., :3], _pil(data))


# ---------------------------------------------------------------- CPU drive of the device code
def _host_lib():
    out = os.path.join(ROOT, "tests", "_build", "libjprog_host.so")
    src = os.path.join(ROOT, "tests", "native", "jprog_host.cpp")
    inc = os.path.join(ROOT, "imagestitching_b200", "csrc")
    deps = [src] + [os.path.join(inc, f) for f in ("jpeg_prog_core.h", "jpeg_parse.h")]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I" + inc, "-o", out, src])
    lib = ctypes.CDLL(out)
    lib.jprog_host_info.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p]
    lib.jprog_host_decode.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
    return lib


def _buffer_order(planes, hs, vs):
    """Oracle coefficient planes (natural order, [by, bx, 64]) -> the GPU buffer's order: interleaved MCUs, zig-zag."""
    y = planes[0][..., ZZ]
    mcuy, mcux = y.shape[0] // vs, y.shape[1] // hs
    yb = y.reshape(mcuy, vs, mcux, hs, 64).transpose(0, 2, 1, 3, 4).reshape(mcuy * mcux, vs * hs, 64)
    if len(planes) == 1:
        return yb.reshape(-1, 64)
    cb, cr = (p[..., ZZ].reshape(mcuy * mcux, 1, 64) for p in planes[1:])
    return np.concatenate([yb, cb, cr], 1).reshape(-1, 64)


def _host_coefficients(lib, data):
    info = np.zeros(8, np.int32)
    assert lib.jprog_host_info(data, len(data), info.ctypes.data) == 0
    coef = np.empty((info[6], 64), np.int16)
    desc = np.zeros((info[5], 8), np.int32)
    assert lib.jprog_host_decode(data, len(data), coef.ctypes.data, desc.ctypes.data) == 0
    return coef, info, desc


def test_sequential_scan_decoder_equals_oracle_coefficients():
    lib = _host_lib()
    for size in SIZES + [(800, 600)]:
        for ss in (0, 1, 2):
            for q, kw in ((30, {}), (90, {"restart_marker_blocks": 3}), (100, {"restart_marker_rows": 1})):
                data = _jpeg(*size, seed=q, quality=q, subsampling=ss, progressive=True, **kw)
                coef, info, _ = _host_coefficients(lib, data)
                want = _buffer_order(J.coefficients(data), int(info[3]), int(info[4]))
                assert np.array_equal(coef, want), f"{size} ss={ss} q={q} {kw}"
        data = _jpeg(*size, gray=True, quality=85, progressive=True)
        coef, info, _ = _host_coefficients(lib, data)
        assert np.array_equal(coef, _buffer_order(J.coefficients(data), 1, 1))


def test_block_walk_equals_block_map():
    lib = _host_lib()
    for size in SIZES + [(333, 211)]:
        for kw in ({"subsampling": 0}, {"subsampling": 1}, {"subsampling": 2}, {"gray": True}):
            data = _jpeg(*size, quality=80, progressive=True, **kw)
            assert lib.jprog_host_walk_check(data, len(data)) == 0, (size, kw)


def test_parser_lists_the_scans_of_libjpegs_script():
    """jpeg_simple_progression for YCbCr: DC (all components, Al 1), Y 1-5 Al 2, Cr 1-63 Al 1, Cb 1-63 Al 1, Y 6-63 Al 2,
    Y 1-63 refine to 1, DC refine, Cr / Cb refine, Y refine to 0."""
    lib = _host_lib()
    data = _jpeg(250, 131, quality=90, subsampling=2, progressive=True)
    _, info, desc = _host_coefficients(lib, data)
    assert info[5] == 10
    # columns: mode, ns, comp, ss, se, ah, al, blocks
    assert desc[0].tolist()[:7] == [0, 3, 0, 0, 0, 0, 1] and desc[0][7] == 16 * 9 * 6
    assert desc[1].tolist()[:7] == [1, 1, 0, 1, 5, 0, 2] and desc[1][7] == 32 * 17          # luma: ceil(250/8) x ceil(131/8)
    assert desc[2].tolist()[:7] == [1, 1, 2, 1, 63, 0, 1] and desc[2][7] == 16 * 9          # chroma: ceil(125/8) x ceil(66/8)
    assert [d[0] for d in desc.tolist()] == [0, 1, 1, 1, 1, 3, 2, 3, 3, 3]
    assert desc[9].tolist()[:7] == [3, 1, 0, 1, 63, 1, 0]


# ---------------------------------------------------------------- scan scripts PIL never writes
def _script_files(sizes=((123, 77), (64, 48), (17, 9)), restarts=(0, 5)):
    """(label, file) for every script of tests/jprog_encode.py x sampling x restart interval, re-coded from PIL baselines."""
    from tests import jprog_encode as E
    for size in sizes:
        for ss, gray in ((0, False), (1, False), (2, False), (None, True)):
            base = _jpeg(*size, seed=5, quality=88, gray=gray, **({} if gray else {"subsampling": ss}))
            assert E.component_ids_from_one(base)
            for name, make in E.SCRIPTS.items():
                for ri in restarts:
                    yield f"{size} ss={ss} {name} ri={ri}", base, E.encode_progressive(base, make(1 if gray else 3), ri)


def test_oracle_and_scan_decoder_on_other_scripts():
    """DC scans per component, spectral selection alone, three refinement passes, restart intervals in AC scans: the
    re-coded file must give PIL the pixels of its source (the test encoder is sound), the oracle PIL's pixels, and both
    CPU-driven forms of the GPU's scan decoder the oracle's coefficients."""
    lib = _host_lib()
    lib.jprog_host_set_mask_path.argtypes = [ctypes.c_int]
    n = 0
    for label, base, f in _script_files():
        assert np.array_equal(_pil(f), _pil(base)), label
        assert np.array_equal(J.decode(f)[..., :3], _pil(f)), label
        for by_mask in (0, 1, 2):
            lib.jprog_host_set_mask_path(by_mask)
            coef, info, _ = _host_coefficients(lib, f)
            assert np.array_equal(coef, _buffer_order(J.coefficients(f), int(info[3]), int(info[4]))), (label, by_mask)
        n += 1
    lib.jprog_host_set_mask_path(0)
    assert n == 96


def test_mask_path_of_the_scan_decoder_equals_oracle_coefficients():
    """pj_refine_parse / pj_refine_parse_z + pj_refine_apply (what the warp kernels run) on PIL's files."""
    lib = _host_lib()
    lib.jprog_host_set_mask_path.argtypes = [ctypes.c_int]
    try:
        for size in SIZES + [(800, 600)]:
            for ss in (0, 1, 2):
                for q, kw in ((20, {}), (90, {"restart_marker_blocks": 3}), (100, {})):
                    data = _jpeg(*size, seed=q, quality=q, subsampling=ss, progressive=True, **kw)
                    want = _buffer_order(J.coefficients(data), 2 if ss else 1, 2 if ss == 2 else 1)
                    for path in (1, 2):
                        lib.jprog_host_set_mask_path(path)
                        coef, info, _ = _host_coefficients(lib, data)
                        assert np.array_equal(coef, want), (size, ss, q, path)
    finally:
        lib.jprog_host_set_mask_path(0)


# ---------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("seq", [0, 1], ids=["parallel-first-scans", "one-thread-per-scan"])
def test_gpu_other_scripts_equal_oracle(seq, monkeypatch):
    from imagestitching_b200 import StitchContext
    monkeypatch.setenv("STITCH_JPROG_SEQ", str(seq))
    ctx = StitchContext(0)
    for label, _, f in _script_files(sizes=((123, 77), (333, 211)), restarts=(0, 7)):
        assert np.array_equal(ctx.decode_jpeg(f), J.decode(f)), label
    ctx.close()


GPU_FILES = [
    dict(w=1, h=1, quality=90, subsampling=2), dict(w=7, h=5, quality=75, subsampling=0),
    dict(w=123, h=77, quality=90, subsampling=2), dict(w=250, h=131, quality=50, subsampling=1),
    dict(w=640, h=481, quality=100, subsampling=2), dict(w=640, h=481, quality=30, subsampling=0),
    dict(w=333, h=211, quality=92, subsampling=2, restart_marker_blocks=3),
    dict(w=333, h=211, quality=92, subsampling=0, restart_marker_rows=1),
    dict(w=100, h=37, quality=85, gray=True), dict(w=801, h=603, quality=95, gray=True),
    dict(w=2000, h=1500, quality=90, subsampling=2),
    dict(w=1500, h=2000, quality=98, subsampling=1),
]


@pytest.mark.gpu
@pytest.mark.parametrize("seq", [0, 1], ids=["parallel-first-scans", "one-thread-per-scan"])
@pytest.mark.parametrize("spec", GPU_FILES, ids=[f"{s['w']}x{s['h']}-{i}" for i, s in enumerate(GPU_FILES)])
def test_gpu_progressive_equals_oracle(spec, seq, monkeypatch):
    from imagestitching_b200 import StitchContext
    monkeypatch.setenv("STITCH_JPROG_SEQ", str(seq))
    spec = dict(spec)
    data = _jpeg(spec.pop("w"), spec.pop("h"), seed=3, progressive=True, **spec)
    assert b"\xff\xc2" in data
    want = J.decode(data)
    ctx = StitchContext(0)
    got = ctx.decode_jpeg(data)
    ctx.close()
    assert got.shape == want.shape
    assert np.array_equal(got, want), f"{np.count_nonzero(got != want)} bytes differ"


@pytest.mark.gpu
@pytest.mark.parametrize("seq", [0, 1])
def test_gpu_progressive_launch_per_level(seq, monkeypatch):
    """STITCH_JPROG_LEVELS=1: the one-warp scans in a launch per dependency level instead of one launch chained through
    progress words (the form a file with order-incompatible dependencies falls back to)."""
    from imagestitching_b200 import StitchContext
    monkeypatch.setenv("STITCH_JPROG_LEVELS", "1")
    monkeypatch.setenv("STITCH_JPROG_SEQ", str(seq))
    ctx = StitchContext(0)
    for spec in (GPU_FILES[2], GPU_FILES[6], GPU_FILES[8], GPU_FILES[10]):
        spec = dict(spec)
        data = _jpeg(spec.pop("w"), spec.pop("h"), seed=3, progressive=True, **spec)
        assert np.array_equal(ctx.decode_jpeg(data), J.decode(data)), spec
    for label, _, f in _script_files(sizes=((123, 77),), restarts=(0, 7)):
        assert np.array_equal(ctx.decode_jpeg(f), J.decode(f)), label
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sub", [128, 256, 512])
def test_gpu_progressive_every_subsequence_size(sub, monkeypatch):
    from imagestitching_b200 import StitchContext
    monkeypatch.setenv("STITCH_JDEC_SUB", str(sub))
    ctx = StitchContext(0)
    for i, (w, h) in enumerate([(1210, 908), (333, 200), (64, 48), (805, 1203)]):
        f = _jpeg(w, h, seed=60 + i, quality=(92, 75, 98, 85)[i], subsampling=(2, 1, 0, 2)[i], progressive=True,
                  **({"restart_marker_blocks": 3} if i == 3 else {}))
        assert np.array_equal(ctx.decode_jpeg(f), J.decode(f)), (sub, w, h)
    ctx.close()


@pytest.mark.gpu
def test_gpu_progressive_randomized():
    from imagestitching_b200 import StitchContext
    rng = np.random.default_rng(4242)
    ctx = StitchContext(0)
    for _ in range(16):
        w, h = int(rng.integers(1, 900)), int(rng.integers(1, 700))
        kw = dict(quality=int(rng.integers(20, 101)), subsampling=int(rng.integers(0, 3)))
        r = rng.random()
        if r < 0.3:
            kw["restart_marker_blocks"] = int(rng.integers(1, 40))
        elif r < 0.45:
            kw["restart_marker_rows"] = int(rng.integers(1, 4))
        gray = rng.random() < 0.15
        if gray:
            kw.pop("subsampling")
        data = _jpeg(w, h, seed=int(rng.integers(0, 1 << 30)), gray=gray, progressive=True, **kw)
        assert np.array_equal(ctx.decode_jpeg(data), J.decode(data)), f"{w}x{h} {kw} gray={gray}"
    ctx.close()


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_gpu_progressive_survives_corruption():
    """Damaged files: every kernel terminates; a bitmap of the right size comes back, or -- when the damage hit the
    header of a later scan or a table definition -- the page's decode error."""
    from imagestitching_b200 import StitchContext, StitchError
    from imagestitching_b200 import _lib as L
    data = _jpeg(400, 300, seed=6, quality=90, subsampling=2, progressive=True)
    ctx = StitchContext(0)
    decoded = 0
    for seed in range(12):
        try:
            assert ctx.decode_jpeg(_corrupt(data, seed, n_flips=6)).shape == (300, 400, 4)
            decoded += 1
        except StitchError as e:
            assert e.status == L.E_BAD_IMAGE
    assert decoded >= 3
    assert np.array_equal(ctx.decode_jpeg(data), J.decode(data))
    ctx.close()


@pytest.mark.gpu
def test_gpu_job_mixes_progressive_baseline_and_bitmaps():
    """A job whose inputs are a progressive file, a baseline file and a decoded bitmap equals the job over the oracle's
    bitmaps, synchronously (stitch_run) and submitted (stitch_submit / stitch_wait, deferred decoder rounds)."""
    from imagestitching_b200 import StitchContext
    dims = [(640, 481), (500, 375), (333, 211)]
    files = [_jpeg(*dims[0], seed=11, quality=90, subsampling=2, progressive=True),
             _jpeg(*dims[1], seed=12, quality=85, subsampling=2),
             _jpeg(*dims[2], seed=13, quality=95, subsampling=0, progressive=True)]
    bitmaps = [J.decode(f) for f in files]
    ctx = StitchContext(0)
    want = ctx.run(U.page_images(dims, bitmaps), "vertical", platform="ios")
    imgs = U.page_images(dims, bitmaps)
    for i in (0, 1):
        imgs[i] = dict(imgs[i], file=files[i])
        imgs[i].pop("bitmap")
    got = ctx.run(imgs, "vertical", platform="ios")
    assert np.array_equal(got.rgba, want.rgba)
    bare = ctx.run([{"file": f} for f in files], "vertical", platform="ios")     # dims from the headers (P:729-734)
    assert np.array_equal(bare.rgba, want.rgba)
    ctx.set_lanes(2)
    tickets = [ctx.submit(imgs if k % 2 else [{"file": f} for f in files], "vertical", platform="ios") for k in range(4)]
    for t in tickets:
        assert np.array_equal(ctx.wait(t).rgba, want.rgba)
    ctx.close()


@pytest.mark.gpu
def test_gpu_progressive_full_size_photo():
    from imagestitching_b200 import StitchContext
    data = _jpeg(4032, 3024, seed=9, quality=90, subsampling=2, progressive=True)
    want = J.decode(data)
    ctx = StitchContext(0)
    got = ctx.decode_jpeg(data)
    ctx.close()
    assert np.array_equal(got, want)
