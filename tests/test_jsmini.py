"""The layout oracle pinned by the reference's OWN source lines, executed (no GPU).

oracle/jsmini.py is a JavaScript-subset interpreter; it runs the verbatim text of pages/index/index.js:1193,
1211-1213, 1243-1248, 1251-1386, 1423-1433, 1522-1557, 1581, 1611-1615 and utils/canvas.js:153-202, sliced by line
number out of /root/reference.  Held against those runs here: the golden vectors, the Python restatement
(oracle/layout_ref.py), the C++ twin in libstitch.so (stitch_layout) and the orientation table (placement_for).
Bar: bit-exact doubles.  Skipped where the reference tree is absent (the GPU box).
"""
import json
import math
import os

import pytest
from hypothesis import given, settings, strategies as st

from imagestitching_b200 import compute_layout
from oracle import jsmini as J
from oracle import layout_ref as LR
from oracle.gen_golden import cases, reference_layout_dict, run_case_reference

pytestmark = pytest.mark.skipif(not J.reference_available(), reason="/root/reference is not on this machine")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "layout_golden.json"), encoding="utf-8"))
KEYS = ("bigTask", "maxSide", "maxPixels", "preW", "preH", "scaleDown", "targetW", "targetH", "superSample", "canvasW",
        "canvasH", "rects", "progress")


# ---- the interpreter itself: JavaScript semantics the layout lines rely on ----
def _js(src, **scope):
    it = J.Interp()
    env = J.Env(it.globals)
    for k, v in scope.items():
        env.declare(k, v)
    env.declare("out", {})
    it.exec_top(J.Parser(src).program(), env, J.UNDEF)
    return env.get("out")


def test_math_round_is_ecma262():
    o = _js("out.a = Math.round(2.5); out.b = Math.round(-2.5); out.c = Math.round(0.49999999999999994);"
            "out.d = Math.round(-0.2); out.e = Math.round(1364.9999999999998); out.f = Math.floor(-1.5);")
    assert (o["a"], o["b"], o["c"], o["e"], o["f"]) == (3.0, -2.0, 0.0, 1365.0, -2.0)
    assert o["d"] == 0.0


def test_logical_operators_return_operands():
    o = _js("const i = { width: 7 }; out.a = i.naturalWidth || i.width || 1; out.b = 0 || 0; out.c = 5 && 0;"
            "out.d = (undefined || 0) || 12288; out.e = !!'x'; out.f = !NaN0;", NaN0=math.nan)
    assert (o["a"], o["b"], o["c"], o["d"], o["e"], o["f"]) == (7.0, 0.0, 0.0, 12288.0, True, True)


def test_array_methods_arrows_and_apply():
    o = _js("const xs = [3, 0, 9, 4]; const pos = xs.map(n => n * 2).filter(n => n > 0);"
            "out.min = Math.min.apply(null, pos); out.max = Math.max.apply(null, pos);"
            "out.sum = xs.reduce((s, v, idx) => { const w = v + (idx ? 0.5 : 0); return s + w; }, 0);"
            "out.empty = Math.min.apply(null, []); out.len = pos.length;")
    assert (o["min"], o["max"], o["sum"], o["empty"], o["len"]) == (6.0, 18.0, 17.5, math.inf, 3.0)


def test_control_flow():
    o = _js("let s = 1; let g = 0; while (s < 100 && g < 20) { s *= 3; if (s > 50) { s = -1; break; } g++; } out.s = s; out.g = g;"
            "let n = 0; for (let i = 0; i < 5; i++) { if (i === 2) continue; n += i; } out.n = n;"
            "function f(o) { switch (o) { case 2: return 'two'; case 3: out.fall = 1; default: return 'other'; } } out.f2 = f(2);"
            " out.f3 = f(3); out.f9 = f('3'); try { throw new Error('boom'); } catch (e) { out.msg = e.message; }"
            "out.t = `a${1 + 1}b${'x'}`; out.fx = (1234.5678).toFixed(1); out.ty = typeof nothing;")
    assert (o["s"], o["g"], o["n"], o["f2"], o["f3"], o["f9"], o["msg"], o["t"], o["fx"], o["ty"]) == \
        (-1.0, 3.0, 8.0, "two", "other", "other", "boom", "a2bx", "1234.6", "undefined")
    assert o["fall"] == 1.0


def test_number_semantics():
    o = _js("out.a = 1 / 0; out.b = Number.isFinite(1 / 0); out.c = 0.1 + 0.2; out.d = 4032 * (1820 / 4032);"
            "out.e = Math.sqrt(22937600 / (4032 * 27216)); out.f = 1 + '2'; out.g = 7 === 7.0; out.h = '7' === 7;")
    assert o["a"] == math.inf and o["b"] is False and o["c"] == 0.1 + 0.2 and o["d"] == 4032 * (1820 / 4032)
    assert o["e"] == math.sqrt(22937600 / (4032 * 27216)) and o["f"] == "12" and o["g"] is True and o["h"] is False


# ---- the program really is the reference's text ----
def test_program_is_sliced_verbatim_from_the_reference():
    ref = open(J.INDEX_JS, encoding="utf-8").read().split("\n") + open(J.CANVAS_JS, encoding="utf-8").read().split("\n")
    ref = set(ref)
    prog = J._page_program().split("\n")
    glue = [l for l in prog if l not in ref]
    assert len(prog) - len(glue) >= 190, len(prog) - len(glue)      # ~200 reference lines are executed
    assert len(glue) <= 20, glue                                     # the rest is loop headers + the hand-back
    for l in glue:
        assert ("GLUE" not in l) and ("Math." not in l) and ("cursor" not in l), l   # glue computes nothing
    draw = J.ref_lines(J.CANVAS_JS, 153, 202).split("\n")
    assert draw[0].startswith("function drawWithOrientation(ctx, img, sx, sy, sw, sh, dx, dy, dw, dh, orientation)")
    assert all(l in ref for l in draw)


# ---- golden vectors == fresh runs of the reference's lines == restatement == C++ twin ----
def _twin(recs, direction, mode, gap, platform, size, pixels):
    return compute_layout(recs, direction, mode, gap, platform, size or None, pixels or None)


@pytest.mark.parametrize("entry", GOLDEN, ids=[g["case"]["name"] for g in GOLDEN])
def test_golden_equals_reference_run(entry):
    c = dict(entry["case"], dims=[tuple(d) for d in entry["case"]["dims"]])
    assert run_case_reference(c) == entry["layout"]
    assert "jsmini" in entry.get("source", "")


def test_golden_covers_the_case_list():
    assert [g["case"]["name"] for g in GOLDEN] == [c["name"] for c in cases()]


dim = st.integers(min_value=1, max_value=12000)
maybe = st.one_of(st.just(0), dim)


@settings(max_examples=250, deadline=None)
@given(
    recs=st.lists(st.tuples(maybe, maybe, maybe, maybe, st.sampled_from([0, 1, 3, 6, 8, 9]),
                            st.sampled_from([0, 0, 0, 3 << 20, 13 << 20])), min_size=1, max_size=9),
    direction=st.sampled_from(["vertical", "horizontal"]),
    mode=st.sampled_from(["min", "max", "original", "weird"]),
    gap=st.one_of(st.integers(min_value=0, max_value=20), st.sampled_from([0.5, 7.25])),
    platform=st.sampled_from(["ios", "android", "devtools"]),
    stored=st.sampled_from([(0, 0), (8192, 0), (0, 3000 * 3000), (16384, 16384 * 16384), (2048, 2048 * 1024)]),
)
def test_restatement_and_twin_equal_the_reference_lines(recs, direction, mode, gap, platform, stored):
    page = [{"naturalWidth": nw, "naturalHeight": nh, "width": w, "height": h, "orientation": o, "fileSize": fs}
            for (nw, nh, w, h, o, fs) in recs]
    want = reference_layout_dict(page, direction, mode, gap, platform, *stored)
    prof = LR.Profile(platform, *stored).with_fallback()
    mine = LR.compute_layout([LR.ImageRec(nw, nh, w, h, o, fs) for (nw, nh, w, h, o, fs) in recs], direction, mode, gap,
                             prof).as_dict()
    twin = _twin(page, direction, mode, gap, platform, *stored)
    for k in KEYS:
        assert mine[k] == want[k], ("restatement", k, mine[k], want[k])
        assert twin[k] == want[k], ("C++ twin", k, twin[k], want[k])


def test_the_reference_never_reaches_its_dimension_throw_after_normalisation():
    """P:1243-1244 force natural dims >= 1, so P:1254 ('图片尺寸不可用') cannot fire on this path: the run succeeds
    with 1x1 images -- exactly what the restatement and the twin do with records that carry no size at all."""
    r = reference_layout_dict([{"orientation": 1}], "vertical", "min", 0, "ios")
    assert r["natural"] == [[1, 1]] and r["preW"] == 1.0 and r["preH"] == 1.0


# ---- drawWithOrientation (utils/canvas.js:153-202), executed, against the closed-form table ----
@pytest.mark.parametrize("ss", [1.0, 2.2, 1.0666666666666667])
@pytest.mark.parametrize("orientation", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, None, "up", "right", True, 2.5])
def test_placement_table_equals_the_executed_ctm(orientation, ss):
    rect = (10.0, 20.5, 30.0, 50.25)
    bw, bh = 8.0, 6.0
    o = J.UNDEF if orientation is None else orientation
    d = J.run_reference_draw(rect, o, ss, bw, bh)
    a, b, c, dd, e, f = d["ctm"]
    sx, sy, sw, sh, lx, ly, lw, lh = d["args"]
    assert (sx, sy, sw, sh) == (0.0, 0.0, bw, bh)

    def ref_dev(u, v):   # bitmap point (u, v in [0, 1]) -> local dest rect -> device, through the recorded CTM
        x, y = lx + u * lw, ly + v * lh
        return (a * x + c * y + e, b * x + dd * y + f)

    p = LR.placement_for(rect, orientation, ss)

    def mine(u, v):
        uu, vv = (1 - u if p.flip_u else u), (1 - v if p.flip_v else v)
        if p.transpose:
            return (p.X0 + vv * p.W, p.Y0 + uu * p.H)
        return (p.X0 + uu * p.W, p.Y0 + vv * p.H)

    for (u, v) in ((0, 0), (1, 0), (0, 1), (1, 1), (0.25, 0.75)):
        gx, gy = ref_dev(u, v)
        mx, my = mine(u, v)
        assert abs(gx - mx) < 1e-9 and abs(gy - my) < 1e-9, (orientation, ss, u, v, (gx, gy), (mx, my))


def test_page_applies_white_fill_and_supersample_ctm():
    """P:1423-1428: one white fillRect over the whole canvas under the identity CTM, then ctx.scale(ss, ss) for every draw."""
    r = J.run_reference_layout([{"naturalWidth": 1080, "naturalHeight": 1920}] * 2, "vertical", 0, platform="ios",
                               device_max_size=12288, device_max_pixels=12288 * 12288)
    assert r["fills"] == [{"ctm": [1.0, 0.0, 0.0, 1.0, 0.0, 0.0], "args": [0.0, 0.0, 2376.0, 8448.0], "fillStyle": "#ffffff"}]
    assert [d["ctm"] for d in r["draws"]] == [[2.2, 0.0, 0.0, 2.2, 0.0, 0.0]] * 2
    assert [d["args"][4:] for d in r["draws"]] == [[0.0, 0.0, 1080.0, 1920.0], [0.0, 1920.0, 1080.0, 1920.0]]
