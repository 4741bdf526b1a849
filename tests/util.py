"""Shared helpers for the parity tests: seeded inputs, oracle calls, comparison."""
from __future__ import annotations

import numpy as np

from oracle import layout_ref as LR
from oracle import raster_c as RC
from oracle import raster_ref as RR

MIXED = [(4032, 3024), (3024, 4032), (1920, 1080), (1080, 1920), (4000, 3000), (3000, 4000), (2048, 2048),
         (6000, 4000), (1280, 720)]

MASK64 = (1 << 64) - 1


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def synth_image(w: int, h: int, seed: int, pattern: int = 0) -> np.ndarray:
    """CPU twin of the on-device generator (imagestitching_b200/csrc/synth.cuh synth_pixel)."""
    with np.errstate(over="ignore"):
        ys, xs = np.meshgrid(np.arange(h, dtype=np.uint64), np.arange(w, dtype=np.uint64), indexing="ij")
        hsh = splitmix64(np.uint64(seed) ^ (ys * np.uint64(w) + xs))
    out = np.empty((h, w, 4), np.uint8)
    xi, yi = xs.astype(np.int64), ys.astype(np.int64)
    if pattern in (1, 2):
        for k in range(4):
            out[:, :, k] = ((hsh >> np.uint64(8 * k)) & np.uint64(0xFF)).astype(np.uint8)
        if pattern == 1:
            out[:, :, 3] = 255
    elif pattern == 3:
        v = np.where(((xi ^ yi) & 1) == 1, 255, 0).astype(np.uint8)
        out[:, :, 0] = out[:, :, 1] = out[:, :, 2] = v
        out[:, :, 3] = 255
    elif pattern == 4:
        v = (xi * 255) // (w - 1 if w > 1 else 1)
        out[:, :, 0] = v
        out[:, :, 1] = 255 - v
        out[:, :, 2] = yi & 255
        out[:, :, 3] = 255
    else:
        for k in range(3):
            t = (xi * 3 + yi * 5 + 40 * k) & 511
            tri = np.where(t < 256, t, 511 - t)
            noise = ((hsh >> np.uint64(8 * k)) & np.uint64(63)).astype(np.int64) - 32
            d = (tri - 128) * 3
            q = np.where(d >= 0, d // 4, -((-d) // 4))          # C division truncates toward zero
            out[:, :, k] = np.clip(128 + q + noise, 0, 255)
        out[:, :, 3] = 255
    return out


def digest(buf) -> int:
    """CPU twin of stitch_device_digest (synth.cuh::digest_kernel) for a bytes-like object."""
    b = np.frombuffer(bytes(buf) if not isinstance(buf, np.ndarray) else np.ascontiguousarray(buf).tobytes(), np.uint8)
    n, nw = b.size, b.size // 4
    K = np.uint64(0xD6E8FEB86659FD93)
    with np.errstate(over="ignore"):
        w = b[: nw * 4].view("<u4").astype(np.uint64)
        acc = np.add.reduce(splitmix64(w ^ (np.arange(nw, dtype=np.uint64) * K)), dtype=np.uint64) if nw else np.uint64(0)
        v = np.uint64(0)
        for k in range(nw * 4, n):
            v |= np.uint64(int(b[k])) << np.uint64(8 * (k & 3))
        acc = acc + splitmix64(v ^ (np.uint64(nw) * K) ^ (np.uint64(n) << np.uint64(32)))
    return int(acc)


def recs(dims, orientations=None, file_sizes=None):
    out = []
    for i, (w, h) in enumerate(dims):
        out.append(LR.ImageRec(w, h, w, h, orientations[i] if orientations else 1, file_sizes[i] if file_sizes else 0))
    return out


def page_images(dims, bitmaps, orientations=None, file_sizes=None):
    out = []
    for i, (w, h) in enumerate(dims):
        out.append({"naturalWidth": w, "naturalHeight": h, "width": w, "height": h,
                    "orientation": orientations[i] if orientations else 1,
                    "fileSize": file_sizes[i] if file_sizes else 0, "bitmap": bitmaps[i]})
    return out


def profile_of(platform="ios", size=0, pixels=0) -> LR.Profile:
    """Same construction as imagestitching_b200.stitch._profile_struct (page state after onLoad)."""
    return LR.Profile(platform, size or 0, pixels or 0).with_fallback()


def oracle_render(dims, bitmaps, direction, mode, gap, profile, orientations=None, yuv=False, use_numpy=False,
                  file_sizes=None):
    orientations = orientations or [1] * len(dims)
    L = LR.compute_layout(recs(dims, orientations, file_sizes), direction, mode, gap, profile)
    if use_numpy:
        out = {"rgba": RR.render_rgba(L, bitmaps, orientations)}
        if yuv:
            out["y"], out["cb"], out["cr"] = RR.render_yuv420(L, bitmaps, orientations)
        return L, out
    return L, RC.render(L, bitmaps, orientations, want_rgba=True, want_yuv=yuv)


def check_pixels(got: np.ndarray, want: np.ndarray, what: str = "", min_psnr: float = 50.0):
    """The parity bar: |diff| <= 1 LSB per channel and PSNR >= 50 dB vs the fp64 oracle."""
    assert got.shape == want.shape, f"{what}: shape {got.shape} vs {want.shape}"
    d = np.abs(got.astype(np.int16) - want.astype(np.int16))
    mx = int(d.max()) if d.size else 0
    p = RR.psnr(got, want)
    assert mx <= 1, f"{what}: max |diff| {mx} > 1 LSB at {np.unravel_index(int(d.argmax()), d.shape)}"
    assert p >= min_psnr, f"{what}: PSNR {p:.2f} dB < {min_psnr}"
    return mx, p
