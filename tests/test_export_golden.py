"""Frozen bytes of the export files: tests/golden/export_golden.json (made by `python -m oracle.gen_export_golden`).

CPU: the sequential writers / the decoder of the test tree still produce exactly these files.  GPU: the PNG file of an
identity job (canvas == bitmap) is the frozen file too -- the GPU JPEG encoder is held to the same writer by the
byte-identity tests of tests/test_jpeg.py on the planes a job composes."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import gen_export_golden as G
from oracle import jpeg_c, png_c
from tests import util as U

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "export_golden.json")))
BIG = 26 * 1024 * 1024


def test_golden_cases_are_the_generator_cases():
    assert [(c["w"], c["h"], c["quality"], c["restart_interval"], c["seed"]) for c in GOLD["jpeg"]] == G.JPEG_CASES
    assert [(c["w"], c["h"], c["seed"]) for c in GOLD["png"]] == G.PNG_CASES


@pytest.mark.parametrize("c", GOLD["jpeg"], ids=lambda c: f"{c['w']}x{c['h']}q{c['quality']}ri{c['restart_interval']}")
def test_jpeg_writer_and_decoder_bytes(c):
    y, cb, cr = G.planes(c["w"], c["h"], c["seed"])
    f = jpeg_c.encode_i420(y, cb, cr, c["quality"], c["restart_interval"])
    assert len(f) == c["bytes"] and hashlib.sha256(f).hexdigest() == c["sha256"]
    dec = np.ascontiguousarray(jpeg_c.decode(f))
    assert dec.shape[:2] == (c["h"], c["w"]) and hashlib.sha256(dec.tobytes()).hexdigest() == c["decoded_sha256"]


@pytest.mark.parametrize("c", GOLD["png"], ids=lambda c: f"{c['w']}x{c['h']}")
def test_png_writer_bytes(c):
    f = png_c.encode_rgba(G.rgba(c["w"], c["h"], c["seed"]))
    assert len(f) == c["bytes"] and hashlib.sha256(f).hexdigest() == c["sha256"]


@pytest.fixture(scope="module")
def ctx():
    from imagestitching_b200 import StitchContext
    c = StitchContext(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("c", [c for c in GOLD["png"] if c["w"] >= 5], ids=lambda c: f"{c['w']}x{c['h']}")
def test_gpu_png_of_identity_job_is_the_frozen_file(ctx, c):
    bm = G.rgba(c["w"], c["h"], c["seed"])
    dims = [(c["w"], c["h"])]
    got = ctx.run(U.page_images(dims, [bm], None, [BIG]), "vertical", platform="ios", output="rgba+png")
    assert (got.rgba == bm).all()
    assert len(got.png) == c["bytes"] and hashlib.sha256(got.png).hexdigest() == c["sha256"]
