"""Layout oracle: literal restatement of the reference's stitch layout math.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Reference (all under /root/reference/miniprogram-stitch/miniprogram/):
  pages/index/index.js:1211-1213   bigTask flag
  pages/index/index.js:1243-1244   dimension normalisation (a1)
  pages/index/index.js:1251-1320   base canvas size + integerise (a2, a3)
  pages/index/index.js:126-156     fallback limit profile (a4)
  pages/index/index.js:1324-1336   effective limits (a4)
  pages/index/index.js:1337-1357   clamp (a5)
  pages/index/index.js:1360-1386   supersample + guard loop (a6)
  pages/index/index.js:1432-1433,1522-1554  placement (a8)
  pages/index/index.js:1247,1358,1556-1557,1581,1611  progress values (a11)

Python floats are IEEE-754 doubles, as JS numbers are; every operation below is
written in the same association order as the JS source.  ``js_round`` follows
ECMA-262 Math.round (ties toward +inf), which is neither Python's ``round`` nor
C's ``round``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

VERTICAL, HORIZONTAL = "vertical", "horizontal"
PLATFORMS = ("ios", "android", "other")


class StitchError(Exception):
    """Mirrors the reference's thrown Error objects (message text preserved)."""


def js_round(x: float) -> float:
    """ECMA-262 Math.round: nearest integer, ties toward +infinity."""
    f = math.floor(x)
    return float(f + 1) if (x - f) >= 0.5 else float(f)


def _truthy(*vals):
    """JS `a || b || c` for numbers: first non-zero, non-None value, else last."""
    for v in vals[:-1]:
        if v:
            return v
    return vals[-1]


@dataclass
class ImageRec:
    """The subset of the prepared-image record (index.js:724-739) the layout reads."""
    naturalWidth: int = 0
    naturalHeight: int = 0
    width: int = 0
    height: int = 0
    orientation: object = 1
    fileSize: int = 0


@dataclass
class Profile:
    """Device canvas-limit profile (index.js:126-156 + wx.getSystemInfoSync().platform)."""
    platform: str = "ios"
    deviceMaxCanvasSize: int = 0       # 0 == unset (JS falsy)
    deviceMaxCanvasPixels: int = 0     # 0 == unset

    def with_fallback(self) -> "Profile":
        """_applyCanvasLimitFromStorageOrFallback, fallback branch (index.js:128-138)."""
        size = self.deviceMaxCanvasSize or (12288 if self.platform == "ios" else 4096)
        pixels = self.deviceMaxCanvasPixels
        if not pixels:
            side = size
            cap = side * min(side, 4096) if self.platform == "android" else side * min(side, 12288)
            pixels = max(cap, 4096 * 2048)
        return Profile(self.platform, size, pixels)


def preset(platform: str) -> Profile:
    return Profile(platform).with_fallback()


@dataclass
class Layout:
    n: int = 0
    direction: str = VERTICAL
    mode: str = "min"
    bigTask: bool = False
    maxSide: float = 0.0
    maxPixels: float = 0.0
    preW: float = 0.0          # after integerise, before clamp
    preH: float = 0.0
    scaleDown: float = 1.0
    targetW: float = 0.0       # == outW/outH after clamp
    targetH: float = 0.0
    superSample: float = 1.0
    canvasW: float = 0.0
    canvasH: float = 0.0
    rects: List[Tuple[float, float, float, float]] = field(default_factory=list)  # user space (dx,dy,dw,dh)
    natural: List[Tuple[int, int]] = field(default_factory=list)
    progress: List[int] = field(default_factory=list)

    def as_dict(self):
        return {
            "n": self.n, "direction": self.direction, "mode": self.mode, "bigTask": self.bigTask,
            "maxSide": self.maxSide, "maxPixels": self.maxPixels, "preW": self.preW, "preH": self.preH,
            "scaleDown": self.scaleDown, "targetW": self.targetW, "targetH": self.targetH,
            "superSample": self.superSample, "canvasW": self.canvasW, "canvasH": self.canvasH,
            "rects": [list(r) for r in self.rects], "natural": [list(x) for x in self.natural],
            "progress": list(self.progress),
        }


def normalise_dims(img: ImageRec) -> Tuple[int, int]:
    """index.js:1243-1244 (also 1236-1237): naturalW/H = max(1, naturalWidth||width||1)."""
    nw = max(1, _truthy(img.naturalWidth, img.width, 1))
    nh = max(1, _truthy(img.naturalHeight, img.height, 1))
    return nw, nh


def effective_limits(p: Profile) -> Tuple[float, float]:
    """index.js:1325-1336."""
    side = p.deviceMaxCanvasSize or (4096 if p.platform == "android" else 12288)
    if p.platform == "ios":
        ios_limit = 16384 * 1400
        pix = min(p.deviceMaxCanvasPixels or ios_limit, ios_limit)
    elif p.platform == "android":
        lim = side * min(side, 8192)
        pix = min(p.deviceMaxCanvasPixels or lim, lim)
    else:
        lim = side * side
        pix = min(p.deviceMaxCanvasPixels or lim, lim)
    return float(side), float(pix)


def progress_sequence(n: int) -> List[int]:
    """Values passed to setData({stitchProgress}) on a successful run, in order.

    index.js:1193 (1), :1247 (probe loop), :1358 (30), :1556-1557 (draw loop),
    :1581 (96), :1611 (100).
    """
    seq = [1]
    for done in range(1, n + 1):
        seq.append(int(min(25, js_round(1 + (done / n) * 24))))
    seq.append(30)
    for idx in range(n):
        p2 = 30 + js_round(((idx + 1) / n) * 60)
        seq.append(int(min(90, p2)))
    seq.append(96)
    seq.append(100)
    return seq


def compute_layout(images: Sequence[ImageRec], direction: str, mode: Optional[str], gap,
                   profile: Profile) -> Layout:
    """Restates index.js:1211-1213, 1243-1386, 1432-1433, 1522-1554."""
    n = len(images)
    if n == 0:
        raise StitchError("empty")          # index.js:1189 returns silently; callers map this to a no-op
    L = Layout(n=n, direction=direction)

    total_bytes = 0
    for it in images:                        # index.js:1211
        total_bytes = total_bytes + (it.fileSize if it.fileSize else 0)
    big = (n >= 7) or (total_bytes >= 25 * 1024 * 1024)    # :1212
    L.bigTask = big

    nat = [normalise_dims(i) for i in images]               # :1243-1244
    L.natural = nat
    allW = [w for (w, _) in nat if w > 0]                   # :1252
    allH = [h for (_, h) in nat if h > 0]                   # :1253
    if not allW or not allH:
        raise StitchError("图片尺寸不可用")                   # :1254

    gapPx = gap or 0                                        # :1256
    mode = mode or "min"                                    # :1257
    L.mode = mode

    outW = 0.0
    outH = 0.0
    if direction == VERTICAL:                               # :1261
        if mode == "min" or mode == "max":
            outW = float(min(allW) if mode == "min" else max(allW))     # :1264 / :1273
            s = 0.0
            for idx, (iw, ih) in enumerate(nat):
                drawH = ih * (outW / iw)                    # :1268 divide first, then multiply
                s = s + drawH + (gapPx if idx else 0)       # :1269 left-assoc
            outH = s
        else:                                               # any other string -> 'original' (:1280)
            outW = float(max(allW))
            s = 0.0
            for idx, (iw, ih) in enumerate(nat):
                s = s + ih + (gapPx if idx else 0)          # :1285
            outH = s
    else:
        if mode == "min" or mode == "max":
            outH = float(min(allH) if mode == "min" else max(allH))     # :1291 / :1300
            s = 0.0
            for idx, (iw, ih) in enumerate(nat):
                drawW = iw * (outH / ih)                    # :1295
                s = s + drawW + (gapPx if idx else 0)       # :1296
            outW = s
        else:
            outH = float(max(allH))
            s = 0.0
            for idx, (iw, ih) in enumerate(nat):
                s = s + iw + (gapPx if idx else 0)          # :1312
            outW = s

    outW = max(1.0, js_round(outW))                         # :1318
    outH = max(1.0, js_round(outH))                         # :1319
    if not outW or not outH or math.isnan(outW) or math.isnan(outH):
        raise StitchError("输出尺寸计算失败")                 # :1320
    L.preW, L.preH = outW, outH

    side, pix = effective_limits(profile)                   # :1325-1336
    L.maxSide, L.maxPixels = side, pix

    scaleDown = 1.0                                         # :1337
    if outW > side or outH > side:                          # :1340
        scaleDown = min(side / outW, side / outH)
    totalPixels = outW * outH                               # :1345
    if totalPixels > pix:
        pixelScale = math.sqrt(pix / totalPixels)           # :1347
        scaleDown = min(scaleDown, pixelScale)
    if scaleDown < 1:                                       # :1351
        outW = max(1.0, float(math.floor(outW * scaleDown)))
        outH = max(1.0, float(math.floor(outH * scaleDown)))
    L.scaleDown = scaleDown

    targetW, targetH = outW, outH                           # :1360-1361
    L.targetW, L.targetH = targetW, targetH
    basePixels = targetW * targetH
    max_ss = 1.0 if big else (2.2 if profile.platform == "ios" else 2.6)     # :1363
    ss = 1.0
    if basePixels > 0 and basePixels < pix:                 # :1365
        ratio = math.sqrt(pix / basePixels)
        if ratio > 1.01:
            sideCap = min(side / targetW, side / targetH)
            ss = min(max_ss, ratio, sideCap)                # :1369
    if not math.isfinite(ss) or ss < 1:                     # :1372
        ss = 1.0
    cw = max(1.0, js_round(targetW * ss))                   # :1373
    ch = max(1.0, js_round(targetH * ss))
    guard = 0
    while cw * ch > pix and guard < 20:                     # :1377
        ss *= 0.96
        if ss < 1:
            ss = 1.0
            break                                           # leaves cw/ch from the previous iteration
        cw = max(1.0, float(math.floor(targetW * ss)))
        ch = max(1.0, float(math.floor(targetH * ss)))
        guard += 1
    L.superSample, L.canvasW, L.canvasH = ss, cw, ch

    scaledGap = gapPx * scaleDown                           # :1432
    cx = 0.0
    cy = 0.0
    rects = []
    for (nw, nh) in nat:                                    # :1522-1554
        if direction == VERTICAL:
            if mode == "original":
                dw = js_round(nw * scaleDown)
                dh = js_round(nh * scaleDown)
                dx = float(math.floor((outW - dw) / 2))
                rects.append((dx, cy, dw, dh))
                cy += dh + scaledGap
            else:
                drawH = js_round(nh * (outW / nw))
                rects.append((0.0, cy, outW, drawH))
                cy += drawH + scaledGap
        else:
            if mode == "original":
                dw = js_round(nw * scaleDown)
                dh = js_round(nh * scaleDown)
                dy = float(math.floor((outH - dh) / 2))
                rects.append((cx, dy, dw, dh))
                cx += dw + scaledGap
            else:
                drawW = js_round(nw * (outH / nh))
                rects.append((cx, 0.0, drawW, outH))
                cx += drawW + scaledGap
    L.rects = rects
    L.progress = progress_sequence(n)
    return L


# ---------------------------------------------------------------------------
# drawWithOrientation (utils/canvas.js:153-202) as a placement in DEVICE space.
# ---------------------------------------------------------------------------

@dataclass
class Placement:
    """Axis-aligned device-space placement of one drawImage call.

    The bitmap's x axis runs along device x (transpose False) or device y
    (transpose True); flip_u / flip_v mirror the bitmap's x / y axis.
    (X0,Y0,W,H) is the device rect the whole bitmap is stretched over.
    """
    X0: float
    Y0: float
    W: float
    H: float
    transpose: bool = False
    flip_u: bool = False
    flip_v: bool = False


def placement_for(rect, orientation, ss: float) -> Placement:
    """Closed form of the CTMs built in utils/canvas.js:160-200,
    composed with ctx.scale(ss, ss) (index.js:1426-1428)."""
    dx, dy, dw, dh = rect
    o = orientation if isinstance(orientation, int) and not isinstance(orientation, bool) else 1
    if o == 2:
        p = Placement(dx, dy, dw, dh, False, True, False)
    elif o == 3:
        p = Placement(dx, dy, dw, dh, False, True, True)
    elif o == 4:
        p = Placement(dx, dy, dw, dh, False, False, True)
    elif o == 5:
        p = Placement(dx, dy, dw, dh, True, False, False)
    elif o == 6:
        p = Placement(dx, dy, dw, dh, True, False, True)
    elif o == 7:
        # reference quirk: translate(dx+dw, dy) lacks +dh, so the image lands ABOVE the rect
        p = Placement(dx, dy - dh, dw, dh, True, True, True)
    elif o == 8:
        p = Placement(dx, dy, dw, dh, True, True, False)
    else:
        p = Placement(dx, dy, dw, dh, False, False, False)
    p.X0 *= ss
    p.Y0 *= ss
    p.W *= ss
    p.H *= ss
    return p


# ---------------------------------------------------------------------------
# Preview of the exported file on the page's <canvas> (index.js:1593-1603).
# ---------------------------------------------------------------------------

@dataclass
class PreviewLayout:
    canvasW: int = 0          # device pixels of the restored preview canvas (utils/canvas.js:19-20)
    canvasH: int = 0
    scaleFit: float = 0.0     # index.js:1600
    pvW: int = 0              # index.js:1601-1602 (CSS px)
    pvH: int = 0
    dx: float = 0.0           # index.js:1603 (CSS px, may end in .5)
    dy: float = 0.0
    dpr: float = 1.0


def preview_layout(export_w: int, export_h: int, view_w: float, view_h: float, dpr) -> PreviewLayout:
    """createHighResCanvas(node, size.width, size.height) (utils/canvas.js:17-24: canvas = floor(size*dpr),
    ctx.scale(dpr, dpr)) followed by the fit-scale draw of index.js:1599-1603."""
    dpr = float(dpr) if dpr else 1.0                                   # `info.pixelRatio || 1`
    pl = PreviewLayout(dpr=dpr)
    pl.canvasW = int(math.floor(view_w * dpr))
    pl.canvasH = int(math.floor(view_h * dpr))
    pl.scaleFit = min(view_w / export_w, view_h / export_h)
    pl.pvW = int(js_round(export_w * pl.scaleFit))
    pl.pvH = int(js_round(export_h * pl.scaleFit))
    pl.dx = (view_w - pl.pvW) / 2
    pl.dy = (view_h - pl.pvH) / 2
    return pl


def preview_placement(pl: PreviewLayout) -> Placement:
    return Placement(pl.dx * pl.dpr, pl.dy * pl.dpr, pl.pvW * pl.dpr, pl.pvH * pl.dpr)
