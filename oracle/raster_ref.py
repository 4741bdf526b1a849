"""Pixel oracle: the stated reference filter "AB1" in fp64 numpy (readable version).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED for pixels:
the reference's resampler is the closed WeChat Canvas2D ``ctx.drawImage``
(utils/canvas.js:156-199 with imageSmoothingQuality='high',
pages/index/index.js:1416-1421); there is nothing in the reference to pin it
against, so this file is the DEFINITION the CUDA kernels are held to
(SURVEY.md App. C), cross-checked against cv2 in tests/test_oracle_raster.py.

What it restates:
  white background            pages/index/index.js:1423-1424
  ctx.scale(ss, ss)           pages/index/index.js:1426-1428
  per-image draw, list order  pages/index/index.js:1439-1554
  orientation CTMs            utils/canvas.js:153-202 (via layout_ref.placement_for)
  export region = full canvas utils/canvas.js:206-209, pages/index/index.js:1577-1579

Filter (per axis, separable): device pixel j covers [j, j+1); its footprint on the
device rect [d0, d0+dlen) is [a,b) with coverage c=b-a.  scale s=dlen/slen.
  s <  1 : exact area average of the source over [(a-d0)/s, (b-d0)/s)
  s >= 1 : bilinear at u=((a+b)/2-d0)/s-0.5, half-pixel centres, clamp-to-edge
Composite: source-over on opaque white in 8-bit gamma space, painter's order,
premultiplied colour; result rounded half-up to u8, alpha = 255.
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np
import scipy.sparse as sp

from .layout_ref import Layout, Placement, placement_for


def axis_weights(n_out: int, d0: float, dlen: float, slen: int, flip: bool):
    """Return (W, cov): W is a CSR [n_out x slen] of filter weights (rows sum to 1
    where cov>0), cov[j] is the coverage of device pixel j by [d0, d0+dlen)."""
    rows: List[int] = []
    cols: List[int] = []
    vals: List[float] = []
    cov = np.zeros(n_out, dtype=np.float64)
    if dlen <= 0 or slen <= 0:
        return sp.csr_matrix((n_out, max(slen, 1))), cov
    s = dlen / slen
    d1 = d0 + dlen
    j_lo = max(0, int(math.floor(d0)))
    j_hi = min(n_out, int(math.ceil(d1)))
    for j in range(j_lo, j_hi):
        a = max(float(j), d0)
        b = min(float(j + 1), d1)
        c = b - a
        if c <= 0:
            continue
        cov[j] = c
        if s < 1:
            s0 = min(max((a - d0) / s, 0.0), float(slen))
            s1 = min(max((b - d0) / s, 0.0), float(slen))
            i0 = min(int(math.floor(s0)), slen - 1)
            i1 = max(i0 + 1, min(int(math.ceil(s1)), slen))
            tot = s1 - s0
            if tot <= 0:
                taps = [(i0, 1.0)]
            else:
                taps = []
                for i in range(i0, i1):
                    w = min(s1, i + 1.0) - max(s0, float(i))
                    if w > 0:
                        taps.append((i, w / tot))
        else:
            u = ((a + b) / 2 - d0) / s - 0.5
            f = math.floor(u)
            t = u - f
            ia = min(max(int(f), 0), slen - 1)
            ib = min(max(int(f) + 1, 0), slen - 1)
            taps = [(ia, 1.0 - t), (ib, t)]
        for i, w in taps:
            rows.append(j)
            cols.append(slen - 1 - i if flip else i)
            vals.append(w)
    W = sp.csr_matrix((vals, (rows, cols)), shape=(n_out, slen))   # duplicate (j,i) entries are summed
    return W, cov


def draw_image(canvas: np.ndarray, bmp: np.ndarray, p: Placement) -> None:
    """Source-over one bitmap [h,w,4] u8 onto canvas [H,W,3] float64 (in place)."""
    H, W = canvas.shape[:2]
    src = bmp.astype(np.float64)
    if p.transpose:
        src = src.transpose(1, 0, 2)        # device x now runs along bitmap y (v)
        flip_x, flip_y = p.flip_v, p.flip_u
    else:
        flip_x, flip_y = p.flip_u, p.flip_v
    sh, sw = src.shape[:2]
    Wx, cx = axis_weights(W, p.X0, p.W, sw, flip_x)
    Wy, cy = axis_weights(H, p.Y0, p.H, sh, flip_y)
    xs = np.nonzero(cx)[0]
    ys = np.nonzero(cy)[0]
    if xs.size == 0 or ys.size == 0:
        return
    x0, x1 = xs[0], xs[-1] + 1
    y0, y1 = ys[0], ys[-1] + 1
    Wx = Wx[x0:x1]
    Wy = Wy[y0:y1]
    alpha = src[:, :, 3] / 255.0
    C = np.outer(cy[y0:y1], cx[x0:x1])
    A = (Wx @ (Wy @ alpha).T).T
    one_minus = 1.0 - C * A
    for ch in range(3):
        prem = src[:, :, ch] * alpha
        F = (Wx @ (Wy @ prem).T).T
        region = canvas[y0:y1, x0:x1, ch]
        canvas[y0:y1, x0:x1, ch] = C * F + one_minus * region


def render_float(layout: Layout, bitmaps: Sequence[np.ndarray], orientations: Sequence[object]) -> np.ndarray:
    """Un-rounded canvas [canvasH, canvasW, 3] float64."""
    cw, ch = int(layout.canvasW), int(layout.canvasH)
    canvas = np.full((ch, cw, 3), 255.0, dtype=np.float64)
    ss = layout.superSample if layout.superSample != 1 else 1.0      # index.js:1426
    for rect, bmp, o in zip(layout.rects, bitmaps, orientations):
        draw_image(canvas, bmp, placement_for(rect, o, ss))
    return canvas


def quantise(x: np.ndarray) -> np.ndarray:
    return np.clip(np.floor(x + 0.5), 0, 255).astype(np.uint8)


def render_rgba(layout: Layout, bitmaps, orientations) -> np.ndarray:
    f = render_float(layout, bitmaps, orientations)
    out = np.empty(f.shape[:2] + (4,), dtype=np.uint8)
    out[:, :, :3] = quantise(f)
    out[:, :, 3] = 255
    return out


def yuv420_from_float(rgb: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """JFIF full-range BT.601 from un-rounded RGB; 2x2 box chroma, edge-replicated
    for odd sizes; planes Y[h,w], Cb/Cr[ceil(h/2), ceil(w/2)] (I420)."""
    R, G, B = rgb[:, :, 0], rgb[:, :, 1], rgb[:, :, 2]
    Y = 0.299 * R + 0.587 * G + 0.114 * B
    Cb = 128.0 - 0.168736 * R - 0.331264 * G + 0.5 * B
    Cr = 128.0 + 0.5 * R - 0.418688 * G - 0.081312 * B
    h, w = Y.shape
    ph, pw = h % 2, w % 2

    def pool(P):
        P = np.pad(P, ((0, ph), (0, pw)), mode="edge")
        return (P[0::2, 0::2] + P[0::2, 1::2] + P[1::2, 0::2] + P[1::2, 1::2]) * 0.25

    return quantise(Y), quantise(pool(Cb)), quantise(pool(Cr))


def render_yuv420(layout: Layout, bitmaps, orientations):
    return yuv420_from_float(render_float(layout, bitmaps, orientations))


def psnr(a: np.ndarray, b: np.ndarray) -> float:
    d = a.astype(np.float64) - b.astype(np.float64)
    mse = float(np.mean(d * d))
    return float("inf") if mse == 0 else 10.0 * math.log10(255.0 * 255.0 / mse)


def render_preview(bmp: np.ndarray, pl) -> np.ndarray:
    """Preview draw (index.js:1597-1603): clearRect -> transparent canvas, one drawImage of the exported
    bitmap into the centred fit rect under ctx.scale(dpr, dpr).  Same filter as draw_image; the canvas keeps
    premultiplied colour + alpha, the result is STRAIGHT-alpha RGBA8 (colour = premultiplied / alpha before
    rounding, alpha = round(255 * coverage * filtered alpha), colour 0 where alpha rounds to 0)."""
    from .layout_ref import preview_placement
    p = preview_placement(pl)
    cw, ch = int(pl.canvasW), int(pl.canvasH)
    out = np.zeros((ch, cw, 4), np.uint8)
    src = np.asarray(bmp).astype(np.float64)
    sh, sw = src.shape[:2]
    Wx, cx = axis_weights(cw, p.X0, p.W, sw, False)
    Wy, cy = axis_weights(ch, p.Y0, p.H, sh, False)
    xs, ys = np.nonzero(cx)[0], np.nonzero(cy)[0]
    if xs.size == 0 or ys.size == 0:
        return out
    x0, x1, y0, y1 = xs[0], xs[-1] + 1, ys[0], ys[-1] + 1
    Wx, Wy = Wx[x0:x1], Wy[y0:y1]
    alpha = src[:, :, 3] / 255.0
    C = np.outer(cy[y0:y1], cx[x0:x1])
    A = C * (Wx @ (Wy @ alpha).T).T                    # canvas alpha in 0..1
    a8 = quantise(A * 255.0)
    safe = np.where(A > 0, A, 1.0)
    for c in range(3):
        F = C * (Wx @ (Wy @ (src[:, :, c] * alpha)).T).T
        out[y0:y1, x0:x1, c] = np.where(a8 > 0, quantise(F / safe), 0)
    out[y0:y1, x0:x1, 3] = a8
    return out
