"""Frozen decoder outputs (tests/golden/decode_golden.json, written by oracle/gen_decode_golden.py): SHA-256 of the
bitmaps for 52 files built from seeded data by deterministic writers -- progressive JPEG under four scan scripts, Adam7
PNG, BMP of every depth, GIF.  CPU: the oracle decoders still return them.  GPU: stitch_decode_image returns them."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import gen_decode_golden as GEN

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _golden():
    with open(os.path.join(ROOT, "tests", "golden", "decode_golden.json")) as fh:
        return {e["name"]: e for e in json.load(fh)}


def _sha(px):
    return hashlib.sha256(np.ascontiguousarray(px).tobytes()).hexdigest()


def test_oracle_decoders_return_the_frozen_bitmaps():
    gold = _golden()
    files = GEN.files()
    assert sorted(n for n, _, _ in files) == sorted(gold)
    for name, kind, f in files:
        px = GEN.DECODERS[kind](f)
        assert (px.shape[1], px.shape[0]) == (gold[name]["width"], gold[name]["height"]), name
        assert _sha(px) == gold[name]["decoded_sha256"], name


@pytest.mark.gpu
def test_gpu_decoders_return_the_frozen_bitmaps():
    from imagestitching_b200 import StitchContext
    gold = _golden()
    ctx = StitchContext(0)
    for name, kind, f in GEN.files():
        assert _sha(ctx.decode_image(f)) == gold[name]["decoded_sha256"], name
    ctx.close()
