"""Layout parity (no GPU): golden vectors vs the Python restatement vs the C++ twin in libstitch.so.

The golden vectors are outputs of the reference's own lines, executed by oracle/jsmini.py (oracle/gen_golden.py);
tests/test_jsmini.py re-runs them where /root/reference exists.  This file needs only the committed JSON.

Bar: bit-exact doubles for canvas size, scaleDown, superSample and every rect (BASELINE north_star).
"""
import json
import math
import os

import pytest
from hypothesis import given, settings, strategies as st

from imagestitching_b200 import compute_layout
from oracle import layout_ref as LR
from oracle.gen_golden import cases, run_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "layout_golden.json"), encoding="utf-8"))


def _twin(c):
    imgs = [{"naturalWidth": w, "naturalHeight": h, "width": w, "height": h, "orientation": 1, "fileSize": fs}
            for (w, h), fs in zip(c["dims"], c["fileSizes"])]
    return compute_layout(imgs, c["direction"], c["mode"], c["gap"], c["platform"], c["size"] or None,
                          c["pixels"] or None)


def _same(a, b, what):
    for k in ("bigTask", "maxSide", "maxPixels", "preW", "preH", "scaleDown", "targetW", "targetH", "superSample",
              "canvasW", "canvasH", "rects", "progress"):
        assert a[k] == b[k], f"{what}: {k}: {a[k]!r} != {b[k]!r}"


@pytest.mark.parametrize("entry", GOLDEN, ids=[g["case"]["name"] for g in GOLDEN])
def test_golden_vs_restatement_and_twin(entry):
    c = entry["case"]
    c = dict(c, dims=[tuple(d) for d in c["dims"]])
    want = entry["layout"]
    _same(run_case(c), want, "restatement")
    _same(_twin(c), want, "C++ twin")


def test_golden_file_is_current():
    assert [g["case"]["name"] for g in GOLDEN] == [c["name"] for c in cases()]


def test_hand_checked_values():
    """SURVEY.md App. B rows, hand-derived from index.js:1251-1386."""
    by = {g["case"]["name"]: g["layout"] for g in GOLDEN}
    a = by["cfg1_ios"]
    assert (a["canvasW"], a["canvasH"], a["superSample"], a["scaleDown"]) == (2376.0, 8448.0, 2.2, 1.0)
    a = by["cfg1_android"]
    assert (a["canvasW"], a["canvasH"], a["superSample"]) == (1152.0, 4096.0, 1.0666666666666667)
    a = by["cfg2_ios"]
    assert (a["preW"], a["preH"], a["scaleDown"]) == (4032.0, 27216.0, 0.4514991181657848)
    assert (a["canvasW"], a["canvasH"], a["bigTask"]) == (1820.0, 12288.0, True)
    assert a["rects"] == [[0.0, 1365.0 * k, 1820.0, 1365.0] for k in range(9)]
    a = by["cfg2_android"]
    assert (a["canvasW"], a["canvasH"], a["scaleDown"]) == (606.0, 4096.0, 0.1504997060552616)
    a = by["cfg3_max_ios"]
    assert (a["canvasW"], a["canvasH"]) == (12288.0, 1139.0)
    assert [r[0] for r in a["rects"]] == [0.0, 1519.0, 2373.0, 4398.0, 5039.0, 6558.0, 7412.0, 8551.0, 10260.0]
    a = by["mixed_original_gap7_ios"]
    assert a["rects"][7] == [-1.0, 7661.605471039883, 2401.0, 1600.0]
    a = by["guard_break_h_max"]          # App. E quirk: over-limit canvas survives, ss reset to 1
    assert (a["targetW"], a["targetH"], a["superSample"], a["canvasW"], a["canvasH"]) == (5323.0, 4173.0, 1.0, 5409.0, 4241.0)
    assert a["canvasW"] * a["canvasH"] > a["maxPixels"]
    assert by["bigtask_by_bytes"]["superSample"] == 1.0 and by["not_big_by_bytes"]["superSample"] == 2.2


def test_js_round_semantics():
    """Math.round: ties toward +inf; not Python round (banker's) and not C round (away from zero)."""
    assert LR.js_round(0.5) == 1 and LR.js_round(1.5) == 2 and LR.js_round(2.5) == 3
    assert LR.js_round(-0.5) == 0 and LR.js_round(-1.5) == -1
    assert LR.js_round(0.49999999999999994) == 0
    assert LR.js_round(1364.6875) == 1365


def test_progress_sequences():
    assert LR.progress_sequence(9) == [1, 4, 6, 9, 12, 14, 17, 20, 22, 25, 30, 37, 43, 50, 57, 63, 70, 77, 83, 90, 96, 100]
    assert LR.progress_sequence(2) == [1, 13, 25, 30, 60, 90, 96, 100]
    for n in range(1, 10):
        imgs = [{"naturalWidth": 10, "naturalHeight": 10}] * n
        assert compute_layout(imgs, "vertical")["progress"] == LR.progress_sequence(n)


dims_st = st.lists(st.tuples(st.integers(1, 12000), st.integers(1, 12000)), min_size=1, max_size=9)


@settings(max_examples=400, deadline=None)
@given(dims=dims_st, direction=st.sampled_from(["vertical", "horizontal"]),
       mode=st.sampled_from(["min", "max", "original", "other"]), gap=st.integers(0, 20),
       platform=st.sampled_from(["ios", "android", "other"]),
       size=st.sampled_from([0, 0, 2048, 4096, 8192, 16384]), pixels=st.sampled_from([0, 0, 4096 * 4096, 16384 * 1400]),
       big=st.booleans())
def test_twin_equals_restatement_random(dims, direction, mode, gap, platform, size, pixels, big):
    fs = [26 * 1024 * 1024 if big else 0] + [0] * (len(dims) - 1)
    c = {"dims": dims, "direction": direction, "mode": mode, "gap": gap, "platform": platform, "size": size,
         "pixels": pixels, "fileSizes": fs}
    want = run_case(c)
    got = _twin(c)
    _same(got, want, f"{c}")
    # soft invariants
    assert want["targetW"] <= want["maxSide"] and want["targetH"] <= want["maxSide"]
    assert want["targetW"] * want["targetH"] <= want["maxPixels"]
    assert all(r[2] >= 0 and r[3] >= 0 for r in want["rects"])
    assert math.isfinite(want["superSample"]) and want["superSample"] >= 1


def test_missing_dims_fall_back(monkeypatch=None):
    """naturalWidth || width || 1 (index.js:1243-1244)."""
    a = compute_layout([{"width": 640, "height": 480}, {"naturalWidth": 320, "naturalHeight": 200}], "vertical")
    assert a["natural"] == [[640, 480], [320, 200]]
    b = compute_layout([{}], "vertical", platform="other", deviceMaxCanvasSize=4096, deviceMaxCanvasPixels=1)
    assert b["natural"] == [[1, 1]] and b["canvasW"] == 1.0
    want = LR.compute_layout([LR.ImageRec()], "vertical", "min", 0, LR.Profile("other", 4096, 1))
    assert (b["canvasW"], b["canvasH"], b["superSample"]) == (want.canvasW, want.canvasH, want.superSample)


def test_errors_without_gpu():
    from imagestitching_b200 import StitchError
    from imagestitching_b200 import _lib as L
    with pytest.raises(StitchError) as e:
        compute_layout([], "vertical")
    assert e.value.status == L.E_EMPTY
    with pytest.raises(StitchError) as e:
        compute_layout([{"naturalWidth": 5, "naturalHeight": 5}] * 10, "vertical")
    assert e.value.status == L.E_TOO_MANY
