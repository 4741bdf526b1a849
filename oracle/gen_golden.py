"""Regenerates tests/golden/layout_golden.json by EXECUTING the reference's own layout lines (oracle/jsmini.py).

    python -m oracle.gen_golden        (in the build container: needs /root/reference)

The reference ships no golden vectors and this image has no JavaScript engine, so the vectors are
produced by running the verbatim text of pages/index/index.js:1193, 1211-1213, 1243-1248, 1251-1386, 1423-1433,
1522-1557, 1581, 1611-1615 and utils/canvas.js:153-202 under jsmini's tree-walking interpreter (`run_case_reference`)
for the BASELINE configs and the quirk cases of SURVEY.md App. B / App. E.  The restatement (layout_ref.py,
`run_case`) and the C++ twin are then held against the file by tests/test_layout.py, and against fresh jsmini
runs by tests/test_jsmini.py.  Doubles are stored with repr() precision (json round-trips them exactly).
"""
import json
import os

from . import jsmini
from .layout_ref import HORIZONTAL, VERTICAL, ImageRec, Profile, compute_layout, preset

MIXED = [(4032, 3024), (3024, 4032), (1920, 1080), (1080, 1920), (4000, 3000), (3000, 4000), (2048, 2048),
         (6000, 4000), (1280, 720)]


def cases():
    out = []

    def add(name, dims, direction, mode, gap, platform, size=0, pixels=0, file_sizes=None):
        out.append({"name": name, "dims": dims, "direction": direction, "mode": mode, "gap": gap,
                    "platform": platform, "size": size, "pixels": pixels, "fileSizes": file_sizes or [0] * len(dims)})

    for plat in ("ios", "android", "other"):
        add(f"cfg1_{plat}", [(1080, 1920)] * 2, VERTICAL, "min", 0, plat)
        add(f"cfg2_{plat}", [(4032, 3024)] * 9, VERTICAL, "min", 0, plat)
        add(f"portrait9_{plat}", [(3024, 4032)] * 9, VERTICAL, "min", 0, plat)
        add(f"cfg3_min_{plat}", MIXED, HORIZONTAL, "min", 0, plat)
        add(f"cfg3_max_{plat}", MIXED, HORIZONTAL, "max", 0, plat)
        add(f"mixed_original_gap7_{plat}", MIXED, VERTICAL, "original", 7, plat)
        add(f"mixed_v_max_gap20_{plat}", MIXED, VERTICAL, "max", 20, plat)
        add(f"mixed_h_original_gap3_{plat}", MIXED, HORIZONTAL, "original", 3, plat)
    # guard-loop break quirk (App. E): canvas keeps the over-limit rounded size while superSample resets to 1
    add("guard_break_h_max", [(5323, 4173)], HORIZONTAL, "max", 0, "ios")
    add("guard_break_v_original", [(3707, 567), (632, 5419)], VERTICAL, "original", 0, "ios")
    # big task by file size only
    add("bigtask_by_bytes", [(1080, 1920)] * 2, VERTICAL, "min", 0, "ios", file_sizes=[13 * 1024 * 1024] * 2)
    add("not_big_by_bytes", [(1080, 1920)] * 2, VERTICAL, "min", 0, "ios", file_sizes=[12 * 1024 * 1024] * 2)
    # unknown mode string: base size takes the 'original' branch, placement the min/max branch (index.js:1280 vs 1527)
    add("mode_other", MIXED[:4], VERTICAL, "weird", 5, "ios")
    # stored device limits
    add("stored_limits_ios", MIXED, VERTICAL, "min", 0, "ios", 8192, 16384 * 1400)
    add("stored_limits_android", MIXED, HORIZONTAL, "max", 4, "android", 8192, 8192 * 8192)
    add("tiny", [(1, 1)], VERTICAL, "min", 0, "other")
    add("strips", [(1, 300), (200, 1)], HORIZONTAL, "max", 1, "ios")
    return out


def run_case(c):
    recs = [ImageRec(w, h, w, h, 1, fs) for (w, h), fs in zip(c["dims"], c["fileSizes"])]
    prof = Profile(c["platform"], c["size"], c["pixels"]).with_fallback()
    return compute_layout(recs, c["direction"], c["mode"], c["gap"], prof).as_dict()


def reference_layout_dict(recs, direction, mode, gap, platform, size=0, pixels=0):
    """One run of the reference's lines (jsmini) in the shape of layout_ref.Layout.as_dict().
    recs: page records (dicts).  size / pixels: this.deviceMaxCanvasSize / deviceMaxCanvasPixels as the page holds
    them when onStitch runs, i.e. AFTER onLoad's fallback profile (index.js:126-156, 342)."""
    prof = Profile(platform, size, pixels).with_fallback()
    r = jsmini.run_reference_layout(recs, direction, gap, vertical_mode=mode, horizontal_mode=mode, platform=platform,
                                    device_max_size=prof.deviceMaxCanvasSize, device_max_pixels=prof.deviceMaxCanvasPixels)
    return {
        "n": len(recs), "direction": direction, "mode": r["mode"], "bigTask": bool(r["bigTask"]),
        "maxSide": r["MAX_CANVAS_SIDE"], "maxPixels": r["MAX_TOTAL_PIXELS"], "preW": r["preW"], "preH": r["preH"],
        "scaleDown": r["scaleDown"], "targetW": r["targetW"], "targetH": r["targetH"], "superSample": r["superSample"],
        "canvasW": r["canvasOutW"], "canvasH": r["canvasOutH"], "rects": [list(c["rect"]) for c in r["calls"]],
        "natural": [[int(im["naturalWidth"]), int(im["naturalHeight"])] for im in r["images"]],
        "progress": [int(p) for p in r["progress"]],
    }


def run_case_reference(c):
    recs = [{"naturalWidth": w, "naturalHeight": h, "width": w, "height": h, "orientation": 1, "fileSize": fs}
            for (w, h), fs in zip(c["dims"], c["fileSizes"])]
    return reference_layout_dict(recs, c["direction"], c["mode"], c["gap"], c["platform"], c["size"], c["pixels"])


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    path = os.path.join(here, "..", "tests", "golden", "layout_golden.json")
    if not jsmini.reference_available():
        raise SystemExit("the reference sources are not here (/root/reference): golden vectors come from running them")
    data = [{"case": c, "layout": run_case_reference(c), "source": "reference lines executed by oracle/jsmini.py"}
            for c in cases()]
    for d in data:   # the restatement must agree before anything is written
        mine = run_case(d["case"])
        assert mine == d["layout"], (d["case"]["name"], mine, d["layout"])
    with open(path, "w", encoding="utf-8") as f:
        json.dump(data, f, indent=1)
    print(f"wrote {len(data)} cases to {os.path.normpath(path)}")


if __name__ == "__main__":
    main()
