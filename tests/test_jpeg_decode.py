"""JPEG decode (utils/canvas.js:27-121 loadImageFrom, index.js:1441-1520).

CPU: the oracle decoder (oracle/jpeg_dec_ref.c) is pinned byte-identical to libjpeg as shipped in PIL.
GPU: stitch_decode_jpeg and STITCH_MEM_HOST_JPEG job inputs are byte-identical to the oracle decoder.
"""
import io

import numpy as np
import pytest
from PIL import Image, ImageFile

from oracle import jpeg_c as J
from tests import util as U

ImageFile.MAXBLOCK = 1 << 24


def _picture(w, h, seed=0, noise=20):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
    a = a + rng.integers(-noise, noise + 1, a.shape)
    return Image.fromarray(np.clip(a, 0, 255).astype(np.uint8))


def _jpeg(w, h, seed=0, gray=False, **kw):
    im = _picture(w, h, seed)
    if gray:
        im = im.convert("L")
    b = io.BytesIO()
    im.save(b, "JPEG", **kw)
    return b.getvalue()


def _pil(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))


SIZES = [(1, 1), (7, 5), (16, 16), (17, 33), (123, 77), (250, 131), (640, 481)]


@pytest.mark.parametrize("size", SIZES, ids=[f"{w}x{h}" for w, h in SIZES])
@pytest.mark.parametrize("subsampling", [0, 1, 2])
def test_oracle_decoder_equals_libjpeg(size, subsampling):
    for q, kw in ((50, {}), (90, {"optimize": True}), (100, {})):
        data = _jpeg(*size, seed=q, quality=q, subsampling=subsampling, **kw)
        got = J.decode(data)
        assert (got[..., 3] == 255).all()
        assert np.array_equal(got[..., :3], _pil(data)), f"{size} ss={subsampling} q={q}"


def test_oracle_decoder_restart_markers_gray_and_own_files():
    for kw in ({"restart_marker_blocks": 3}, {"restart_marker_rows": 1}, {"restart_marker_blocks": 1}):
        for ss in (0, 2):
            data = _jpeg(333, 211, quality=92, subsampling=ss, **kw)
            assert b"\xff\xdd" in data
            assert np.array_equal(J.decode(data)[..., :3], _pil(data))
    data = _jpeg(100, 37, gray=True, quality=85)
    assert np.array_equal(J.decode(data)[..., :3], _pil(data))
    rng = np.random.default_rng(5)
    y, cb, cr = (rng.integers(0, 256, s, dtype=np.uint8) for s in ((200, 333), (100, 167), (100, 167)))
    data = J.encode_i420(y, cb, cr, 90, 8)                       # the product's own export format
    assert np.array_equal(J.decode(data)[..., :3], _pil(data))


def _as_arithmetic(data: bytes) -> bytes:
    """The same file with its SOF marker turned into SOF9 (arithmetic coding): what no decoder here accepts."""
    sof = data.index(b"\xff\xc0")
    return data[:sof + 1] + b"\xc9" + data[sof + 2:]


def test_oracle_decoder_rejects_arithmetic_coding():
    with pytest.raises(ValueError):
        J.decode(_as_arithmetic(_jpeg(64, 64, quality=80)))


def _corrupt(data: bytes, seed: int, n_flips: int = 12, scan_only: bool = True) -> bytes:
    rng = np.random.default_rng(seed)
    b = bytearray(data)
    lo = data.index(b"\xff\xda") + 14 if scan_only else 2
    for pos in rng.integers(lo, len(b) - 2, n_flips):
        b[pos] = int(rng.integers(0, 256))
    return bytes(b)


def test_oracle_decoder_survives_corruption():
    """Damaged entropy data must give SOME bitmap of the right size (libjpeg behaves the same way); damaged
    headers must give an error or a bitmap, never a crash."""
    data = _jpeg(250, 131, quality=85, subsampling=2, restart_marker_blocks=4)
    for seed in range(40):
        out = J.decode(_corrupt(data, seed))
        assert out.shape == (131, 250, 4)
    for seed in range(40):
        try:
            J.decode(_corrupt(data, 1000 + seed, n_flips=3, scan_only=False))
        except ValueError:
            pass


# ---------------------------------------------------------------- GPU
GPU_FILES = [
    dict(w=1, h=1, quality=90, subsampling=2), dict(w=7, h=5, quality=75, subsampling=0),
    dict(w=123, h=77, quality=90, subsampling=2), dict(w=250, h=131, quality=50, subsampling=1),
    dict(w=640, h=481, quality=100, subsampling=2), dict(w=640, h=481, quality=95, subsampling=0, optimize=True),
    dict(w=333, h=211, quality=92, subsampling=2, restart_marker_blocks=3),
    dict(w=333, h=211, quality=92, subsampling=0, restart_marker_rows=1),
    dict(w=100, h=37, quality=85, gray=True),
    dict(w=2000, h=1500, quality=90, subsampling=2),             # ~ 0.5 MB of entropy data: thousands of subsequences
    dict(w=1500, h=2000, quality=98, subsampling=1, optimize=True),
]


@pytest.mark.gpu
@pytest.mark.parametrize("spec", GPU_FILES, ids=[f"{s['w']}x{s['h']}-{i}" for i, s in enumerate(GPU_FILES)])
def test_gpu_decoder_equals_oracle(spec):
    from imagestitching_b200 import StitchContext
    spec = dict(spec)
    data = _jpeg(spec.pop("w"), spec.pop("h"), seed=3, **spec)
    want = J.decode(data)
    ctx = StitchContext(0)
    got = ctx.decode_jpeg(data)
    ctx.close()
    assert got.shape == want.shape
    assert np.array_equal(got, want), f"{np.count_nonzero(got != want)} bytes differ"


def _with_fill_bytes(data: bytes) -> bytes:
    """Insert 0xFF fill bytes in front of every RSTn marker (legal: T.81 B.1.1.2)."""
    sos = data.index(b"\xff\xda")
    head, scan = data[:sos], bytearray(data[sos:])
    out = bytearray()
    i = 0
    while i < len(scan):
        if scan[i] == 0xFF and i + 1 < len(scan) and 0xD0 <= scan[i + 1] <= 0xD7:
            out += b"\xff\xff"                         # two fill bytes, then the marker itself
        out.append(scan[i])
        i += 1
    return head + bytes(out)


def test_oracle_decoder_fill_bytes():
    data = _with_fill_bytes(_jpeg(200, 120, quality=90, subsampling=2, restart_marker_blocks=2))
    assert b"\xff\xff\xff\xd0" in data
    assert np.array_equal(J.decode(data)[..., :3], _pil(data))


@pytest.mark.gpu
def test_gpu_decoder_fill_bytes():
    from imagestitching_b200 import StitchContext
    data = _with_fill_bytes(_jpeg(640, 481, seed=4, quality=90, subsampling=2, restart_marker_blocks=5))
    ctx = StitchContext(0)
    got = ctx.decode_jpeg(data)
    ctx.close()
    assert np.array_equal(got, J.decode(data))


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(4))
def test_gpu_decoder_randomized(seed):
    """Random sizes, qualities, samplings, table optimisation and restart intervals: byte-identical to the oracle."""
    from imagestitching_b200 import StitchContext
    rng = np.random.default_rng(900 + seed)
    ctx = StitchContext(0)
    for _ in range(10):
        w, h = int(rng.integers(1, 900)), int(rng.integers(1, 700))
        kw = dict(quality=int(rng.integers(25, 101)), subsampling=int(rng.integers(0, 3)))
        if rng.random() < 0.5:
            kw["optimize"] = True
        r = rng.random()
        if r < 0.3:
            kw["restart_marker_blocks"] = int(rng.integers(1, 40))
        elif r < 0.45:
            kw["restart_marker_rows"] = int(rng.integers(1, 4))
        gray = rng.random() < 0.15
        if gray:
            kw.pop("subsampling")
        data = _jpeg(w, h, seed=int(rng.integers(0, 1 << 30)), gray=gray, **kw)
        want = J.decode(data)
        got = ctx.decode_jpeg(data)
        assert np.array_equal(got, want), f"{w}x{h} {kw} gray={gray}: {np.count_nonzero(got != want)} bytes differ"
    ctx.close()


@pytest.mark.gpu
def test_gpu_decoder_full_size_photo():
    """A 12 MP file as the page's inputs are (BASELINE configs[1] shape)."""
    from imagestitching_b200 import StitchContext
    data = _jpeg(4032, 3024, seed=9, quality=90, subsampling=2)
    want = J.decode(data)
    ctx = StitchContext(0)
    got = ctx.decode_jpeg(data)
    ctx.close()
    assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_gpu_decoder_survives_corruption():
    """Damaged scans: the decoder terminates and returns a bitmap of the right size (the synchronisation loop is
    bounded, block indices are range-checked).  Blocks are placed by COUNT, so damage that changes the number of
    blocks of a restart interval displaces what follows (libjpeg would re-align at the next RSTn); the picture in
    front of the damage is intact."""
    from imagestitching_b200 import StitchContext
    data = _jpeg(640, 481, seed=6, quality=90, subsampling=2, restart_marker_rows=1)
    want = J.decode(data)
    ctx = StitchContext(0)
    for seed in range(6):
        bad = _corrupt(data, seed, n_flips=4)
        first = next(i for i, (x, y) in enumerate(zip(data, bad)) if x != y)
        got = ctx.decode_jpeg(bad)
        assert got.shape == want.shape
        # rows wholly coded before the first damaged byte: a generous lower bound from the byte position
        safe_rows = int(481 * (first - data.index(b"\xff\xda")) / len(data)) // 16 * 16 - 32
        if safe_rows > 0:
            assert np.array_equal(got[:safe_rows], want[:safe_rows])
    plain = _jpeg(333, 211, seed=7, quality=90, subsampling=0)
    for seed in range(4):
        assert ctx.decode_jpeg(_corrupt(plain, 50 + seed, n_flips=2)).shape == (211, 333, 4)
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sub", [128, 256, 512])
def test_gpu_decoder_every_subsequence_size(sub, monkeypatch):
    """The batch picks 128 / 256 / 512 raw bytes per subsequence by its size (jdec_pick_sub); every size must give the
    oracle's bytes -- forced here on small files, restart markers and all three samplings included."""
    from imagestitching_b200 import StitchContext
    monkeypatch.setenv("STITCH_JDEC_SUB", str(sub))
    ctx = StitchContext(0)
    for i, (w, h) in enumerate([(1210, 908), (333, 200), (64, 48), (805, 1203)]):
        f = _jpeg(w, h, seed=60 + i, quality=(92, 75, 98, 85)[i], subsampling=(2, 1, 0, 2)[i],
                  **({"restart_marker_blocks": 3} if i == 3 else {}))
        assert np.array_equal(ctx.decode_jpeg(f)[:, :, :3], J.decode(f)[:, :, :3]), (sub, w, h)
    ctx.close()


@pytest.mark.gpu
def test_gpu_decoder_errors():
    from imagestitching_b200 import StitchContext, StitchError
    from imagestitching_b200 import _lib as L
    ctx = StitchContext(0)
    with pytest.raises(StitchError) as e:
        ctx.decode_jpeg(_as_arithmetic(_jpeg(64, 64, quality=80)))
    assert e.value.status == L.E_BAD_IMAGE and "解码异常" in e.value.message
    # three components that are NOT YCbCr (libjpeg decodes these as RGB): refused, never decoded to wrong colours
    base = _jpeg(64, 48, quality=85, subsampling=0)
    app0 = base.index(b"\xff\xe0")          # drop the JFIF APP0 segment: with it libjpeg (and the parser) say YCbCr
    assert base[app0 + 4:app0 + 9] == b"JFIF\0"
    base = base[:app0] + base[app0 + 2 + ((base[app0 + 2] << 8) | base[app0 + 3]):]
    sof = base.index(b"\xff\xc0")
    sos = base.index(b"\xff\xda")
    rgb_ids = bytearray(base)
    for c, ch in enumerate(b"RGB"):            # component ids 'R','G','B' in the frame and the scan header
        rgb_ids[sof + 4 + 6 + 3 * c] = ch
        rgb_ids[sos + 4 + 1 + 2 * c] = ch
    adobe = base[:2] + b"\xff\xee\x00\x0eAdobe\x00\x64\x00\x00\x00\x00\x00" + base[2:]     # APP14, transform = 0
    swapped = bytearray(base)                  # scan names the components in another order than the frame
    swapped[sos + 4 + 1], swapped[sos + 4 + 3] = base[sos + 4 + 3], base[sos + 4 + 1]
    for bad in (bytes(rgb_ids), adobe, bytes(swapped)):
        with pytest.raises(StitchError) as e:
            ctx.decode_jpeg(bad)
        assert e.value.status == L.E_BAD_IMAGE, bad[:4]
    assert ctx.decode_jpeg(base).shape == (48, 64, 4)
    ctx.close()


@pytest.mark.gpu
def test_job_with_jpeg_files_equals_job_with_decoded_bitmaps():
    """index.js:1441-1554 end to end: files in -> decode on the GPU -> stitch; same canvas as stitching the
    oracle-decoded bitmaps."""
    from imagestitching_b200 import StitchContext, prepare_image
    dims = [(806, 604), (604, 806), (1210, 908)]
    files = [_jpeg(w, h, seed=20 + i, quality=92, subsampling=(2, 0, 1)[i]) for i, (w, h) in enumerate(dims)]
    recs = [prepare_image(f) for f in files]
    ctx = StitchContext(0)
    a = ctx.run([dict(r, file=f) for r, f in zip(recs, files)], "vertical", mode="min", gap=3, platform="ios")
    b = ctx.run([dict(r, bitmap=J.decode(f)) for r, f in zip(recs, files)], "vertical", mode="min", gap=3, platform="ios")
    ctx.close()
    assert a.width == b.width and a.height == b.height
    assert np.array_equal(a.rgba, b.rgba)
