"""GIF inputs (P:4 SUPPORTED_IMAGE_TYPES 'gif'; utils/canvas.js:27-121 loadImageFrom): the first frame.

CPU: the oracle (oracle/gif_ref.py: classic table-of-strings LZW) is pinned to PIL's GifImagePlugin, frame 0.
GPU: stitch_decode_gif / stitch_decode_image and GIF job inputs equal the oracle.
"""
import io

import numpy as np
import pytest
from PIL import Image

from imagestitching_b200 import _lib as L
from oracle import gif_ref as G


def _picture(w, h, seed=0, noise=8):
    y, x = np.mgrid[0:h, 0:w]
    a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
    return np.clip(a + np.random.default_rng(seed).integers(-noise, noise + 1, a.shape), 0, 255).astype(np.uint8)


def _gif(im, **kw):
    b = io.BytesIO()
    im.save(b, "GIF", **kw)
    return b.getvalue()


def _pil_rgba(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGBA"))


def files():
    rng = np.random.default_rng(1)
    out = []
    for (w, h) in ((1, 1), (7, 5), (64, 48), (123, 77), (250, 131)):
        # PIL interlaces by default (pictures of 16 px and more): interlace=0 asks for rows in order
        for kw in ({"interlace": 0}, {"interlace": 1}, {"transparency": 3, "interlace": 0}, {"interlace": 1, "transparency": 0}):
            for colours in (2, 16, 256):
                out.append((f"{w}x{h}_{colours}_{kw}", _gif(Image.fromarray(_picture(w, h, colours)).quantize(colours), **kw)))
    g = np.zeros((300, 400), np.uint8)        # a flat graphic: long strings, the dictionary fills up and is cleared
    g[50:200, 80:300] = 7
    g[::17] = 3
    im = Image.fromarray(g, "P")
    im.putpalette([int(v) for v in rng.integers(0, 256, 768)])
    out.append(("graphic", _gif(im)))
    out.append(("graphic_rows_in_order", _gif(im, interlace=0)))
    out.append(("gray", _gif(Image.fromarray(_picture(200, 150)[..., 0], "L"))))
    out.append(("photo_800x600", _gif(Image.fromarray(_picture(800, 600, 5)).quantize(256), interlace=0)))
    # strings longer than the decoder's shared-memory window of recent output, and sources behind it
    big = np.zeros((900, 1600), np.uint8)
    big[100:800, 200:1400] = 1
    big[::97, ::3] = 2
    im = Image.fromarray(big, "P")
    im.putpalette([int(v) for v in rng.integers(0, 256, 768)])
    out.append(("flat_1600x900", _gif(im)))
    # two frames: only the first is painted
    a, b = Image.fromarray(_picture(90, 60, 1)).quantize(64), Image.fromarray(_picture(90, 60, 2)).quantize(64)
    buf = io.BytesIO()
    a.save(buf, "GIF", save_all=True, append_images=[b], duration=100)
    out.append(("animated", buf.getvalue()))
    return out


def test_oracle_equals_pil():
    for name, data in files():
        assert np.array_equal(G.decode(data), _pil_rgba(data)), name


def _lzw_as_copies(stream: bytes, min_code: int, n_out: int) -> np.ndarray:
    """The GPU kernel's form of LZW (gif_decode.cuh G1), restated: the dictionary holds (start, length) into the output
    already written, a 256-byte prelude in front of the output holds the roots, a code is a copy."""
    clear, eoi = 1 << min_code, (1 << min_code) + 1
    out = bytearray(range(256))
    start, length = list(range(clear)) + [0, 0], [1] * clear + [0, 0]
    size, nxt, prev, prev_pos = min_code + 1, clear + 2, -1, 0
    acc = cnt = 0
    it = iter(stream)
    total = 256 + n_out
    while len(out) < total:
        while cnt < size:
            b = next(it, None)
            if b is None:
                return np.frombuffer(bytes(out[256:]).ljust(n_out, b"\0"), np.uint8)
            acc |= b << cnt
            cnt += 8
        code = acc & ((1 << size) - 1)
        acc >>= size
        cnt -= size
        if code == clear:
            start, length = start[:clear + 2], length[:clear + 2]
            size, nxt, prev = min_code + 1, clear + 2, -1
            continue
        if code == eoi:
            break
        pos = len(out)
        if code < nxt and (code < clear or code > eoi):
            src, n = start[code], length[code]
        else:
            assert code == nxt and prev >= 0
            src, n = prev_pos, length[prev] + 1
        for j in range(n):                       # the lanes' copy: the last byte of the second case is the copy's own first
            s = src + j
            out.append(out[s if s < pos else src])
        if prev >= 0 and nxt < 4096:
            start.append(prev_pos)
            length.append(length[prev] + 1)
            nxt += 1
            if nxt == (1 << size) and size < 12:
                size += 1
        prev, prev_pos = code, pos
    return np.frombuffer(bytes(out[256:total]).ljust(n_out, b"\0"), np.uint8)


def test_lzw_as_copies_equals_table_of_strings():
    """The kernel's dictionary-as-substrings form against the oracle's classic one, on the streams of real files."""
    for name, data in files():
        if len(data) < 200000:
            assert np.array_equal(G.decode(data, lzw_fn=_lzw_as_copies), G.decode(data)), name


def test_oracle_rejects():
    good = _gif(Image.fromarray(_picture(20, 10)).quantize(16))
    with pytest.raises(G.GifError):
        G.decode(b"BM" + bytes(40))
    lsd = bytearray(good)
    lsd[6] = 30                                   # the logical screen is wider than the frame
    with pytest.raises(G.GifError):
        G.decode(bytes(lsd))


@pytest.fixture()
def ctx():
    from imagestitching_b200 import StitchContext
    c = StitchContext(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_gpu_gif_equals_oracle(ctx):
    for name, data in files():
        want = G.decode(data)
        assert np.array_equal(ctx.decode_gif(data), want), name
        assert np.array_equal(ctx.decode_image(data), want), name


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_gpu_gif_errors_and_damage(ctx):
    from imagestitching_b200 import StitchError
    good = _gif(Image.fromarray(_picture(120, 80)).quantize(64))
    lsd = bytearray(good)
    lsd[6] = 200
    with pytest.raises(StitchError) as e:
        ctx.decode_gif(bytes(lsd))
    assert e.value.status == L.E_BAD_IMAGE
    rng = np.random.default_rng(4)
    start = good.index(b"\x2c") + 11
    for seed in range(8):                          # damaged LZW data: a bitmap, or the decode error -- never a hang
        bad = bytearray(good)
        for pos in rng.integers(start + 2, len(bad) - 4, 5):
            if pos % 256 != (start % 256):         # keep clear of most sub-block length bytes
                bad[pos] ^= int(rng.integers(1, 256))
        try:
            assert ctx.decode_gif(bytes(bad)).shape == (80, 120, 4)
        except StitchError as e2:
            assert e2.status == L.E_BAD_IMAGE
    assert np.array_equal(ctx.decode_gif(good), G.decode(good))
    # a job with a damaged GIF fails; the next job on the same context -- no GIF in it -- is not blamed for it
    from tests import util as U
    worst = bytearray(good)
    worst[start + 3: start + 40] = bytes([0xFF]) * 37
    bmpish = G.decode(good)
    try:
        ctx.run([{"file": bytes(worst)}, {"file": good}], "vertical", platform="ios")
    except StitchError as e4:
        assert e4.status == L.E_BAD_IMAGE
    ok = ctx.run(U.page_images([(120, 80)] * 2, [bmpish, bmpish]), "vertical", platform="ios")
    assert ok.rgba.shape[1] > 0
    cut = good[: len(good) // 2]                   # cut short: what the data delivers, index 0 behind it (or refused)
    try:
        assert ctx.decode_gif(cut).shape == (80, 120, 4)
    except StitchError as e3:
        assert e3.status == L.E_BAD_IMAGE


@pytest.mark.gpu
def test_gpu_job_with_gif_files(ctx):
    from tests import util as U
    dims = [(123, 77), (250, 131)]
    fs = [_gif(Image.fromarray(_picture(*dims[0], 3)).quantize(256), transparency=5), _gif(Image.fromarray(_picture(*dims[1], 4)).quantize(32), interlace=1)]
    bitmaps = [G.decode(f) for f in fs]
    want = ctx.run(U.page_images(dims, bitmaps), "vertical", platform="ios")
    got = ctx.run([{"file": f} for f in fs], "vertical", platform="ios")
    assert np.array_equal(got.rgba, want.rgba)
    ctx.set_lanes(2)
    tickets = [ctx.submit([{"file": f} for f in fs], "vertical", platform="ios") for _ in range(3)]
    for t in tickets:
        assert np.array_equal(ctx.wait(t).rgba, want.rgba)
