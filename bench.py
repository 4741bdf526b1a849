#!/usr/bin/env python
"""bench.py -- output MP/s of the 9x12MP vertical stitch (BASELINE.json configs[1]) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2|cfg2_yuv|cfg2_jpeg|cfg2_png|cfg2_android|cfg5|cfg1|cfg3|cfg3_min]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one stitch job (layout + fused resample/compose of all 9 images into the
canvas) with the decoded RGBA bitmaps already resident in HBM.  `value` = output
megapixels of all ranks / device time (CUDA events on the launch stream, max over
ranks).  `e2e` = the same job through the public C ABI with pinned HOST buffers
(H2D of the 9 bitmaps + D2H of the canvas inside the timed region).
`roofline` = algorithmic bytes (sum of source bytes + canvas bytes) per
launch / average launch duration vs the measured HBM copy peak.
`cpu_baseline` = the C restatement of the same path (oracle/raster_ref.c, fp32, OpenMP)
timed on this box's host cores on a bounded sample.  --impl reference times that CPU
path alone: the reference itself (a WeChat mini-program) cannot run outside WeChat.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os

# the library's job lanes are streams of their own; more of them than hardware queues (default 8) makes kernels of one
# job wait behind another job's (stitch_ctx_create sets the same default when it is the first to touch CUDA; here torch is)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MIXED = [(4032, 3024), (3024, 4032), (1920, 1080), (1080, 1920), (4000, 3000), (3000, 4000), (2048, 2048),
         (6000, 4000), (1280, 720)]
WORKLOADS = {
    # name: (dims, direction, mode, platform, output, description)
    "cfg2": ([(4032, 3024)] * 9, "vertical", "min", "ios", "rgba",
             "9x 4032x3024 RGBA vertical stitch, iOS canvas clamp -> 1820x12288 (BASELINE configs[1])"),
    "cfg2_yuv": ([(4032, 3024)] * 9, "vertical", "min", "ios", "yuv420",
                 "9x 4032x3024 vertical stitch with fused RGBA->YUV420 epilogue (BASELINE configs[3])"),
    "cfg2_jpeg": ([(4032, 3024)] * 9, "vertical", "min", "ios", "jpeg",
                  "9x 4032x3024 vertical stitch exported as a baseline JPEG file (quality 90, GPU encoder; SURVEY 8f rank 1)"),
    "cfg2_png": ([(4032, 3024)] * 9, "vertical", "min", "ios", "png",
                 "9x 4032x3024 vertical stitch exported as the PNG file the page asks for (GPU encoder; index.js:1579)"),
    "cfg2_android": ([(4032, 3024)] * 9, "vertical", "min", "android", "rgba",
                     "9x 4032x3024 RGBA vertical stitch, Android/other canvas clamp -> 606x4096 (6.65x downscale; SURVEY 8d)"),
    "cfg3_min": (MIXED, "horizontal", "min", "ios", "rgba",
                 "9 mixed-aspect images horizontal 'min' -> 7765x720 (all downscale; SURVEY 8d)"),
    "cfg5": ([(4032, 3024)] * 9, "vertical", "min", "ios", "rgba",
             "batch of 10k independent 9x 4032x3024 stitch jobs sharded over the ranks, 32 resident job slots per GPU "
             "(BASELINE configs[4]; sustained load: see clocks)"),
    "cfg1": ([(1080, 1920)] * 2, "vertical", "min", "ios", "rgba",
             "2x 1080x1920 vertical stitch, 2.2x supersampled canvas 2376x8448 (BASELINE configs[0])"),
    "cfg3": (MIXED, "horizontal", "max", "ios", "rgba",
             "9 mixed-aspect images horizontal 'max' -> 12288x1139 (BASELINE configs[2])"),
}
SEED = 0x5717C4
PREWARM_JOBS = 100   # untimed jobs in front of the warm-up steps (about 10 ms of GPU work)


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def bind_near_gpu(cuda_index: int) -> dict:
    """Host placement for the e2e leg: run this rank's submitting thread on the cores of the GPU's NUMA node and
    prefer that node for new pages, so the pinned bitmaps / canvases (stitch_host_alloc) sit next to the GPU's
    PCIe root.  With one rank per GPU every rank otherwise allocates wherever the launcher put it, and at N > 1
    the uploads cross the socket link.  Best effort: returns what was done, never raises."""
    info = {"node": None, "cpus": None, "mempolicy": None}
    try:
        import torch
        pr = torch.cuda.get_device_properties(cuda_index)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return info
        info["node"] = node
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        sel = cpus & os.sched_getaffinity(0)
        if sel:
            os.sched_setaffinity(0, sel)
            info["cpus"] = len(sel)
        # set_mempolicy(MPOL_PREFERRED, {node}): x86-64 syscall 238
        libc = C.CDLL(None, use_errno=True)
        mask = (C.c_ulong * 16)()
        mask[node // 64] = 1 << (node % 64)
        rc = libc.syscall(C.c_long(238), C.c_int(1), mask, C.c_ulong(16 * 64))
        info["mempolicy"] = "preferred" if rc == 0 else f"errno {C.get_errno()}"
    except Exception as e:   # no sysfs entry, no permission: stay where we are
        info["error"] = str(e)[:80]
    return info


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period: float = 0.01):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.nv = None

    def sample(self):
        if not self.ok:
            return
        nv = self.nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            try:
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            names = {
                0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting",
                0x10: "sync_boost",
            }
            for bit, name in names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self._stop_evt.is_set():
            self.sample()
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        self.sample()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_reference_run(workload: str, steps: int, warmup: int, threads: int | None = None):
    """Times the C restatement (fp32, OpenMP) of the path on the host cores: one full job per step."""
    import numpy as np

    from oracle import layout_ref as LR
    from oracle import raster_c as RC
    from tests import util as U

    dims, direction, mode, platform, output, _ = WORKLOADS[workload]
    cores = threads or len(os.sched_getaffinity(0)) or os.cpu_count() or 1
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm is meant to use every host core
    os.environ["OMP_NUM_THREADS"] = str(cores)
    bitmaps = []
    cache = {}
    for i, (w, h) in enumerate(dims):
        key = (w, h)
        if key not in cache:                      # content does not affect CPU timing; reuse per shape
            cache[key] = U.synth_image(w, h, SEED + i, 0)
        bitmaps.append(cache[key])
    lay = LR.compute_layout(U.recs(dims), direction, mode, 0, U.profile_of(platform))
    mp = lay.canvasW * lay.canvasH / 1e6
    want_yuv = output != "rgba"

    def one():
        lay2 = LR.compute_layout(U.recs(dims), direction, mode, 0, U.profile_of(platform))
        RC.render(lay2, bitmaps, [1] * len(dims), want_rgba=not want_yuv or output == "both", want_yuv=want_yuv,
                  precision="f32")

    RC.set_threads(cores)
    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t0
    return mp * steps / dt, dt / steps * 1e3, cores, mp


def cpu_sanity_baselines(workload: str):
    """beside the all-cores port, (a) the same port on ONE thread and (b) a library sanity baseline,
    cv2.resize (INTER_AREA when shrinking, INTER_LINEAR when enlarging) + numpy paste on all cores -- one job each."""
    out = {}
    try:
        v, ms, _, _ = cpu_reference_run(workload, 1, 0, threads=1)
        out["one_thread"] = {"value": v, "unit": "MP/s", "ms_per_step": ms, "cores": 1}
    except Exception as e:
        out["one_thread"] = {"value": None, "note": f"failed: {e}"}
    try:
        import cv2
        import numpy as np

        from oracle import layout_ref as LR
        from tests import util as U
        dims, direction, mode, platform, output, _ = WORKLOADS[workload]
        cores = len(os.sched_getaffinity(0)) or 1
        cv2.setNumThreads(cores)
        bitmaps = {}
        for i, (w, h) in enumerate(dims):
            if (w, h) not in bitmaps:
                bitmaps[(w, h)] = U.synth_image(w, h, SEED + i, 0)
        lay = LR.compute_layout(U.recs(dims), direction, mode, 0, U.profile_of(platform))
        ss = lay.superSample if lay.superSample != 1 else 1.0
        cw, ch = int(lay.canvasW), int(lay.canvasH)

        def one():
            canvas = np.full((ch, cw, 4), 255, np.uint8)
            for (dx, dy, dw, dh), (w, h) in zip(lay.rects, dims):
                x0, y0 = int(round(dx * ss)), int(round(dy * ss))
                tw, th = max(1, int(round(dw * ss))), max(1, int(round(dh * ss)))
                interp = cv2.INTER_AREA if (tw < w or th < h) else cv2.INTER_LINEAR
                r = cv2.resize(bitmaps[(w, h)], (tw, th), interpolation=interp)
                xa, ya, xb, yb = max(0, x0), max(0, y0), min(cw, x0 + tw), min(ch, y0 + th)
                if xb > xa and yb > ya:
                    canvas[ya:yb, xa:xb] = r[ya - y0:yb - y0, xa - x0:xb - x0]
            return canvas

        one()
        t0 = time.perf_counter()
        for _ in range(3):
            one()
        dt = (time.perf_counter() - t0) / 3
        out["cv2"] = {"value": cw * ch / 1e6 / dt, "unit": "MP/s", "ms_per_step": dt * 1e3, "cores": cores,
                      "what": "cv2.resize (INTER_AREA / INTER_LINEAR) + numpy paste, integer-rounded rects: a sanity figure, not the oracle"}
    except Exception as e:
        out["cv2"] = {"value": None, "note": f"failed: {e}"}
    return out


def synth_screenshot_png(w: int, h: int, seed: int = 0) -> bytes:
    """A phone-screenshot-like RGBA PNG file: a coloured title bar, white ground, blocks of text-like speckle, a gradient
    card.  Encoded by PIL (libpng filter heuristics, zlib level 6) -- the encoder real screenshots come from."""
    import io

    import numpy as np
    from PIL import Image

    rng = np.random.default_rng(seed)
    a = np.full((h, w, 4), 255, np.uint8)
    a[: h // 10] = (32, 120, 240, 255)
    for _ in range(60):
        y0 = int(rng.integers(h // 10, h - 40))
        x0 = int(rng.integers(0, w - 400))
        ww = int(rng.integers(100, 400))
        a[y0:y0 + 28, x0:x0 + ww, :3] = rng.integers(0, 2, (28, ww, 1)) * 220
    a[h // 2: h // 2 + h // 8, w // 10: w - w // 10, 0] = np.linspace(0, 255, w - 2 * (w // 10)).astype(np.uint8)
    b = io.BytesIO()
    Image.fromarray(a, "RGBA").save(b, "PNG")
    return b.getvalue()


def synth_progressive_jpeg(w: int, h: int, seed: int = 0) -> bytes:
    """A progressive JPEG file (SOF2, libjpeg's standard 10-scan script, quality 90, 4:2:0) of a smooth picture with mild
    noise, written by PIL / libjpeg-turbo -- the encoder such files come from."""
    import io

    import numpy as np
    from PIL import Image, ImageFile

    ImageFile.MAXBLOCK = 1 << 26
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
    a = np.clip(a + rng.integers(-4, 5, a.shape), 0, 255).astype(np.uint8)
    b = io.BytesIO()
    Image.fromarray(a).save(b, "JPEG", quality=90, subsampling=2, progressive=True)
    return b.getvalue()


def cpu_png_files_run(files, direction, mode, platform, steps: int):
    """The PNG-in / PNG-out job on the host cores: oracle PNG decoder (one thread per file), C raster (OpenMP), oracle PNG
    writer.  Returns (MP/s, ms per job, cores)."""
    from concurrent.futures import ThreadPoolExecutor

    from oracle import layout_ref as LR
    from oracle import png_c as PC
    from oracle import png_dec_c as PD
    from oracle import raster_c as RC
    from tests import util as U

    cores = len(os.sched_getaffinity(0)) or os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    pool = ThreadPoolExecutor(max_workers=min(cores, len(files)))

    def one():
        bitmaps = list(pool.map(PD.decode, files))
        dims = [(b.shape[1], b.shape[0]) for b in bitmaps]
        lay = LR.compute_layout(U.recs(dims), direction, mode, 0, U.profile_of(platform))
        out = RC.render(lay, bitmaps, [1] * len(dims), want_rgba=True, want_yuv=False, precision="f32")
        return lay, len(PC.encode_rgba(out["rgba"]))

    lay, _ = one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = (time.perf_counter() - t0) / steps
    pool.shutdown()
    mp = lay.canvasW * lay.canvasH / 1e6
    return mp / dt, dt * 1e3, cores


def cpu_files_run(files, direction, mode, platform, steps: int, out_kind: str = "png"):
    """The file-in / file-out job on the host cores: oracle JPEG decoder (one thread per file), C raster (OpenMP),
    oracle PNG / JPEG writer (sequential).  Returns (MP/s, ms per job, cores)."""
    from concurrent.futures import ThreadPoolExecutor

    from oracle import jpeg_c as JC
    from oracle import layout_ref as LR
    from oracle import png_c as PC
    from oracle import raster_c as RC
    from tests import util as U

    cores = len(os.sched_getaffinity(0)) or os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    pool = ThreadPoolExecutor(max_workers=min(cores, len(files)))

    def one():
        bitmaps = list(pool.map(JC.decode, files))            # ctypes releases the GIL
        dims = [(b.shape[1], b.shape[0]) for b in bitmaps]
        lay = LR.compute_layout(U.recs(dims), direction, mode, 0, U.profile_of(platform))
        out = RC.render(lay, bitmaps, [1] * len(dims), want_rgba=out_kind == "png", want_yuv=out_kind == "jpeg", precision="f32")
        if out_kind == "png":
            return lay, len(PC.encode_rgba(out["rgba"]))
        return lay, len(JC.encode_i420(out["y"], out["cb"], out["cr"], 90, 8))

    lay, _ = one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = (time.perf_counter() - t0) / steps
    pool.shutdown()
    mp = lay.canvasW * lay.canvasH / 1e6
    return mp / dt, dt * 1e3, cores


def pool_sweep(args):
    """BASELINE configs[4] through the PRODUCT's multi-GPU entry: ONE process, stitch_pool_* (one worker thread + ctx per
    GPU, include/stitch.h), 10 000 jobs with device-resident bitmaps, static assignment gpu = job_id % G.
    Wall clock around submit-all + wait-all; digests of every GPU's K slot canvases must agree."""
    from imagestitching_b200 import DeviceBitmap, StitchPool, compute_layout
    from imagestitching_b200 import _lib as L
    from imagestitching_b200.stitch import _profile_struct

    dims, direction, mode, platform, output, desc = WORKLOADS[args.workload]
    G = args.pool
    pool = StitchPool(list(range(G)), lanes_per_device=max(1, args.lanes))
    lib = pool.lib
    K = args.slots or 32
    n_jobs = args.steps or 10000
    profile = _profile_struct(platform)
    recs = [{"naturalWidth": w, "naturalHeight": h, "width": w, "height": h, "orientation": 1, "fileSize": 0} for (w, h) in dims]
    lay = compute_layout(recs, direction, mode, 0, platform)
    cw, ch = int(lay["canvasW"]), int(lay["canvasH"])
    mp, rgba_bytes = cw * ch / 1e6, cw * ch * 4
    src_bytes = sum(w * h * 4 for (w, h) in dims)
    jobs, keep = [], []   # jobs[g][s]
    for g in range(G):
        h = pool.ctx_handle(g)
        row = []
        for s_ in range(K):
            imgs = []
            for i, (w, hh) in enumerate(dims):
                pitch = (w * 4 + 15) & ~15
                p = C.c_void_p()
                assert lib.stitch_device_alloc(h, hh * pitch, C.byref(p)) == 0
                assert lib.stitch_generate_image(h, p, w, hh, pitch, SEED + 16 * s_ + i, 0) == 0
                imgs.append(dict(recs[i], bitmap=DeviceBitmap(p.value, w, hh, pitch, keepalive=p)))
            out = C.c_void_p()
            assert lib.stitch_device_alloc(h, rgba_bytes, C.byref(out)) == 0
            job, k2 = pool._helper.build_job(imgs, direction, mode, 0, profile, "rgba", True, out.value, 0, rgba_bytes, 0)
            row.append((job, out))
            keep.append((imgs, k2))
        assert lib.stitch_memcpy_h2d(h, row[0][1], (C.c_char * 16)(), 16) == 0   # drains the generator launches
        jobs.append(row)
    results = [[L.StitchResult() for _ in range(K)] for _ in range(G)]

    def sweep(n):
        t = C.c_uint64()
        for j in range(n):
            g, s_ = j % G, (j // G) % K
            st = lib.stitch_pool_submit(pool.handle, C.byref(jobs[g][s_][0]), C.byref(results[g][s_]), g, C.byref(t))
            assert st == 0, st
        assert lib.stitch_pool_wait_all(pool.handle) == 0, lib.stitch_pool_last_error(pool.handle)

    sweep(100 * G)   # warm-up: lanes' buffers, tensor-map caches, clocks
    sampler = ClockSampler(0)
    sampler.start()
    t0 = time.perf_counter()
    sweep(n_jobs)
    dt = time.perf_counter() - t0
    sampler.sample()
    sampler.stop()
    digs = []
    for g in range(G):
        row = []
        for s_ in range(K):
            d, t = C.c_uint64(0), C.c_uint64()
            assert lib.stitch_pool_digest(pool.handle, g, jobs[g][s_][1], rgba_bytes, C.byref(d), C.byref(t)) == 0
            assert lib.stitch_pool_wait(pool.handle, t.value, None) == 0
            row.append(int(d.value))
        digs.append(row)
    ran = [pool.jobs_run(g) for g in range(G)]
    pool.close()
    peak, peak_src = _peaks()
    per_gpu_us = dt / (n_jobs / G) * 1e6
    achieved = (src_bytes + rgba_bytes) / (per_gpu_us * 1e-6) / 1e9
    line = {
        "metric": "output MP/s, 9x12MP vertical stitch", "value": mp * n_jobs / dt, "unit": "MP/s", "n_gpus": G, "steps": n_jobs,
        "warmup": 100 * G, "ms_per_step": dt / n_jobs * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u8 in/out, i32 vertical / f32 horizontal accumulate", "data": "synthetic", "mode": "pool",
        "config": {"workload": desc, "name": args.workload, "canvas": [cw, ch],
                   "parallelism": f"1 process, StitchPool over {G} GPU(s): a worker thread + ctx per GPU, {max(1, args.lanes)} lanes each, "
                                  f"job j -> GPU j % {G}, slot (j // {G}) % {K}; no collective",
                   "timing": "host wall clock around submit-all + wait-all (the pool has no cross-device event)"},
        "us_per_job_per_gpu": per_gpu_us, "gpu_launches": n_jobs,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "peak_source": peak_src, "note": "per GPU, wall clock: includes host submission"},
        "clocks": sampler.summary(),
        "digests": {"slots": K, "gpus": G, "equal_across_gpus": all(r == digs[0] for r in digs), "distinct": len(set(digs[0])),
                    "slot0": f"{digs[0][0]:016x}"},
        "jobs_run_per_gpu": ran,
    }
    assert line["digests"]["equal_across_gpus"], "canvases differ between GPUs"
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="default 300; cfg5: 10000 // ranks (the batch is sharded)")
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--slots", type=int, default=None,
                    help="distinct resident job slots rotated between steps (default 3; cfg5: 32 = 14 GB of bitmaps per GPU)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = auto (min(steps, 40))")
    ap.add_argument("--cpu-steps", type=int, default=10, help="jobs the CPU port is timed on (cpu_baseline)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-numa", action="store_true", help="do not bind the rank to the GPU's NUMA node")
    ap.add_argument("--lanes", type=int, default=2,
                    help="job lanes of the StitchCtx the device-resident steps rotate over (1 = one job at a time)")
    ap.add_argument("--e2e-lanes", type=int, default=3, help="job lanes for the host-buffer (e2e) steps")
    ap.add_argument("--files-lanes", type=int, default=6, help="job lanes for the file-in / file-out (e2e_files) legs")
    ap.add_argument("--files-steps", type=int, default=-1,
                    help="e2e_files leg (JPEG files in host memory -> exported file in host memory): jobs per output kind; "
                         "-1 = auto (24 for cfg2 / cfg5, 0 = skip for the other workloads)")
    ap.add_argument("--pool", type=int, default=0,
                    help="N > 0: run the batch sweep in ONE process through stitch_pool_* over N GPUs (no torchrun)")
    args = ap.parse_args()
    if args.pool > 0:
        args.workload = "cfg5" if args.workload == "cfg2" else args.workload
        return pool_sweep(args)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dims, direction, mode, platform, output, desc = WORKLOADS[args.workload]
    batch = args.workload == "cfg5"
    if args.steps is None:
        args.steps = (10000 + world - 1) // world if batch and args.impl == "ours" else (3 if args.impl == "reference" else 300)
    if args.slots is None:
        args.slots = 32 if batch else 3
    W = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        if rank != 0:
            return 0
        steps = max(1, args.steps)
        warm = max(0, min(args.warmup, 10))       # a warm-up step is one ~50 ms CPU job: run what was asked for
        val, ms, cores, mp = cpu_reference_run(args.workload, steps, warm)
        line = {
            "impl": "reference", "metric": "output MP/s, 9x12MP vertical stitch", "value": val, "unit": "MP/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "name": args.workload},
            "cpu_baseline": {"value": val, "unit": "MP/s", "cores": cores, "kind": "port",
                             "sample": f"{steps} full job(s) of the workload, C restatement (oracle/raster_ref.c fp32, OpenMP)"},
            "e2e": {"value": val, "unit": "MP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "the reference (WeChat mini-program, closed Canvas2D runtime) cannot run here; this times the CPU restatement of its path",
        }
        print(json.dumps(line))
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist

    from imagestitching_b200 import DeviceBitmap, StitchContext
    from imagestitching_b200 import _lib as L
    from imagestitching_b200.stitch import _profile_struct

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    affinity0 = os.sched_getaffinity(0)
    placement = bind_near_gpu(local_rank) if not args.no_numa else {"node": None}
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.Stream()
    ctx = StitchContext(local_rank, stream.cuda_stream)
    lib = ctx.lib
    profile = _profile_struct(platform)

    # ---- resident job slots (bitmaps generated on device by K0) ----
    slots = []
    n = len(dims)
    with torch.cuda.stream(stream):
        for s in range(max(1, args.slots)):
            imgs, keep = [], []
            for i, (w, h) in enumerate(dims):
                pitch = (w * 4 + 15) & ~15
                t = torch.empty(h * pitch, dtype=torch.uint8, device="cuda")
                st = lib.stitch_generate_image(ctx.handle, t.data_ptr(), w, h, pitch, SEED + 16 * s + i, 0)   # the same on every rank
                assert st == 0
                keep.append(t)
                imgs.append({"naturalWidth": w, "naturalHeight": h, "width": w, "height": h, "orientation": 1,
                             "fileSize": 0, "bitmap": DeviceBitmap(t.data_ptr(), w, h, pitch)})
            slots.append((imgs, keep))
    # layout (for sizes)
    from imagestitching_b200 import compute_layout
    lay = compute_layout(slots[0][0], direction, mode, 0, platform)
    cw, ch = int(lay["canvasW"]), int(lay["canvasH"])
    mp = cw * ch / 1e6
    rgba_bytes = cw * ch * 4
    yuv_bytes = cw * ch + 2 * ((cw + 1) // 2) * ((ch + 1) // 2)
    src_bytes = sum(w * h * 4 for (w, h) in dims)
    has_jpeg = output == "jpeg"
    has_png = output == "png"
    jpeg_cap = 64 << 20 if has_jpeg else 0        # quality 90: the 22 MP canvas compresses to a few MB
    png_cap = rgba_bytes + (rgba_bytes >> 2) if has_png else 0   # noisy synthetic content barely compresses
    out_bytes = (rgba_bytes if output in ("rgba", "both") else 0) + (yuv_bytes if output in ("yuv420", "both") else 0)
    algo_bytes = src_bytes + out_bytes

    jobs = []
    for (imgs, keep) in slots:
        o_rgba = torch.empty(rgba_bytes, dtype=torch.uint8, device="cuda") if output in ("rgba", "both") else None
        o_yuv = torch.empty(yuv_bytes, dtype=torch.uint8, device="cuda") if output in ("yuv420", "both") else None
        o_jpeg = torch.empty(jpeg_cap, dtype=torch.uint8, device="cuda") if has_jpeg else None
        job, keep2 = ctx.build_job(imgs, direction, mode, 0, profile, output, True,
                                   o_rgba.data_ptr() if o_rgba is not None else 0,
                                   o_yuv.data_ptr() if o_yuv is not None else 0, rgba_bytes, yuv_bytes,
                                   o_jpeg.data_ptr() if has_jpeg else 0, jpeg_cap, 90 if has_jpeg else 0, 0)
        o_png = torch.empty(png_cap, dtype=torch.uint8, device="cuda") if has_png else None
        if has_png:
            job.out_png, job.out_capacity_png = o_png.data_ptr(), png_cap
        jobs.append((job, keep2, o_rgba, o_yuv, o_jpeg, o_png))
    res = L.StitchResult()

    host_s = [0.0]

    # BASELINE configs[4]: job ids 0..9999 are sharded round-robin over the ranks (shard.py: static gpu = job_id % G,
    # SURVEY 8e); job j works on resident slot (j // G) % K, so every rank cycles through the same K slots and a job's
    # canvas does not depend on which GPU ran it -- checked below with on-device digests.
    from imagestitching_b200.shard import shard_jobs
    my_jobs = shard_jobs(args.steps * world, rank, world) if batch else None

    def step(i):
        job = jobs[(my_jobs[i % len(my_jobs)] // world if batch else i) % len(jobs)][0]
        t = C.c_uint64()
        h0 = time.perf_counter()
        st = lib.stitch_submit(ctx.handle, C.byref(job), C.byref(res), C.byref(t))
        host_s[0] += time.perf_counter() - h0
        if st != 0:
            raise RuntimeError(f"stitch_submit failed: {st} {lib.stitch_last_error(ctx.handle)}")
        return t.value

    def barrier():
        if world > 1:
            dist.barrier()

    # ---- device-resident timing ----
    # Steps are independent jobs; they rotate over `--lanes` lanes of the ctx so that one job's ramp-up
    # overlaps the previous job's drain.  ev0/ev1 sit on the ctx stream (lane 0); stitch_ctx_join makes
    # that stream wait for the other lanes before ev1 is recorded.
    ctx.set_lanes(max(1, args.lanes))
    last = 0
    # settle clocks / power state and every lane's lazily created buffers before the W warm-up steps: on a box that
    # has just come up the first milliseconds of GPU work run slower (one N = 2 run measured 120 us/job over 50 steps)
    for i in range(PREWARM_JOBS):
        last = step(i)
    for i in range(W):
        last = step(i)
    lib.stitch_wait(ctx.handle, last)
    torch.cuda.synchronize()
    barrier()
    sampler = ClockSampler(local_rank)
    launches0 = ctx.kernel_launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    sampler.start()
    ev0.record(stream)
    host_s[0] = 0.0
    for i in range(args.steps):
        last = step(i)
    ctx.join()
    ev1.record(stream)
    sampler.sample()
    lib.stitch_wait(ctx.handle, last)
    torch.cuda.synchronize()
    sampler.stop()
    barrier()
    elapsed_ms = ev0.elapsed_time(ev1)
    launches = ctx.kernel_launches - launches0
    # digests of the canvases the sweep left in the K job slots, on the device; equal on every rank or the run is void
    digests = None
    if batch and output == "rgba":
        mine = []
        for (_, _, o_rgba, _, _, _) in jobs:
            d = C.c_uint64(0)
            assert lib.stitch_device_digest(ctx.handle, o_rgba.data_ptr(), rgba_bytes, C.byref(d)) == 0
            mine.append(int(d.value))
        allv = [mine]
        if world > 1:
            allv = [None] * world
            dist.all_gather_object(allv, mine)
        digests = {"slots": len(mine), "ranks": world, "equal_across_ranks": all(v == allv[0] for v in allv),
                   "distinct": len(set(allv[0])), "slot0": f"{allv[0][0]:016x}"}
        assert digests["equal_across_ranks"], "canvases differ between GPUs"
    if world > 1:
        t = torch.tensor([elapsed_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
    value = mp * args.steps * world / (elapsed_ms / 1e3)
    ms_per_step = elapsed_ms / args.steps

    # ---- end-to-end through the C ABI with pinned host buffers ----
    e2e_steps = args.e2e_steps or max(3, min(args.steps, 40))   # long enough to amortise the last job's drain (1.7 ms)
    host_imgs, host_keep = [], []
    for i, (w, h) in enumerate(dims):
        p = C.c_void_p()
        assert lib.stitch_host_alloc(w * h * 4, C.byref(p)) == 0
        arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(h, w, 4))
        # fill from the device-generated bitmap of slot 0 (row pitch == w*4 for these widths or copy row-wise)
        t = slots[0][1][i]
        pitch = (w * 4 + 15) & ~15
        arr[:] = t.view(h, pitch)[:, : w * 4].reshape(h, w, 4).cpu().numpy()
        host_keep.append(p)
        host_imgs.append({"naturalWidth": w, "naturalHeight": h, "width": w, "height": h, "orientation": 1,
                          "fileSize": 0, "bitmap": arr})
    # one host output buffer per lane: jobs in flight must not share an export target
    e2e_lanes = max(1, args.e2e_lanes)
    lib.stitch_wait(ctx.handle, last)
    ctx.set_lanes(e2e_lanes)
    host_outs, hjobs = [], []
    for _ in range(e2e_lanes):
        ho_rgba, ho_yuv = C.c_void_p(), C.c_void_p()
        if output in ("rgba", "both"):
            assert lib.stitch_host_alloc(rgba_bytes, C.byref(ho_rgba)) == 0
        if output in ("yuv420", "both"):
            assert lib.stitch_host_alloc(yuv_bytes, C.byref(ho_yuv)) == 0
        host_outs += [ho_rgba, ho_yuv]
        ho_jpeg = C.c_void_p()
        if has_jpeg:
            assert lib.stitch_host_alloc(jpeg_cap, C.byref(ho_jpeg)) == 0
            host_outs.append(ho_jpeg)
        hjobs.append(ctx.build_job(host_imgs, direction, mode, 0, profile, output, False, ho_rgba.value or 0,
                                   ho_yuv.value or 0, rgba_bytes, yuv_bytes, ho_jpeg.value or 0, jpeg_cap,
                                   90 if has_jpeg else 0, 0))
        if has_png:
            ho_png = C.c_void_p()
            assert lib.stitch_host_alloc(png_cap, C.byref(ho_png)) == 0
            host_outs.append(ho_png)
            hjobs[-1][0].out_png, hjobs[-1][0].out_capacity_png = ho_png.value, png_cap
    hres = L.StitchResult()

    def e2e_loop(nsteps):
        tickets = []
        for i in range(nsteps):
            t = C.c_uint64()
            st = lib.stitch_submit(ctx.handle, C.byref(hjobs[i % e2e_lanes][0]), C.byref(hres), C.byref(t))
            assert st == 0, lib.stitch_last_error(ctx.handle)
            tickets.append(t.value)
            if i >= e2e_lanes - 1:          # the host buffers of job i+1's lane must have been drained
                lib.stitch_wait(ctx.handle, tickets[i - (e2e_lanes - 1)])
        lib.stitch_wait(ctx.handle, tickets[-1])

    e2e_loop(2 * e2e_lanes)   # warm-up: staging allocation, first-touch
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    e2e_loop(e2e_steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    # latency of ONE job through stitch_run (upload, draw, export strictly in sequence)
    t1 = time.perf_counter()
    for _ in range(3):
        assert lib.stitch_run(ctx.handle, C.byref(hjobs[0][0]), C.byref(hres)) == 0
    e2e_single_ms = (time.perf_counter() - t1) / 3 * 1e3
    # the same job with PAGEABLE buffers (plain numpy arrays: what stitch.py / the N-API shim really hold): the library
    # stages them through pinned memory with its copy pool (csrc/host_stage.h)
    pageable_ms = None
    if output == "rgba":
        pg_imgs = [dict(im, bitmap=np.array(im["bitmap"])) for im in host_imgs]
        pg_out = np.empty(rgba_bytes, dtype=np.uint8)
        pg_job = ctx.build_job(pg_imgs, direction, mode, 0, profile, output, False, pg_out.ctypes.data, 0, rgba_bytes, 0)
        assert lib.stitch_run(ctx.handle, C.byref(pg_job[0]), C.byref(hres)) == 0
        t1 = time.perf_counter()
        for _ in range(3):
            assert lib.stitch_run(ctx.handle, C.byref(pg_job[0]), C.byref(hres)) == 0
        pageable_ms = (time.perf_counter() - t1) / 3 * 1e3
        pinned_view = np.ctypeslib.as_array(C.cast(host_outs[0], C.POINTER(C.c_uint8)), shape=(rgba_bytes,))
        assert np.array_equal(pg_out, pinned_view), "pageable and pinned runs differ"
        del pg_job, pg_imgs
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = mp * e2e_steps * world / e2e_s

    # ---- end to end from FILES: 9 JPEG files in (pinned) host memory -> decode + stitch + encode on the GPU -> the
    # exported file in host memory (what the page does between P:1441 and P:1581).  Uploads 14x fewer bytes than e2e.
    files_steps = args.files_steps if args.files_steps >= 0 else (36 if args.workload in ("cfg2", "cfg5") else 0)
    f_lanes = max(1, min(8, args.files_lanes))   # file jobs are GPU-bound, not PCIe-bound: more of them in flight fill the gaps
    e2e_files = None
    shot_files = None
    prog_files = None
    file_bytes_list = []
    if files_steps > 0:
        lib.stitch_wait(ctx.handle, last)
        ctx.set_lanes(1)
        fbufs = []
        for i, (w, h) in enumerate(dims):   # the input files: each bitmap of slot 0 exported at its own size, quality 90
            o = ctx.run([slots[0][0][i]], "vertical", platform="other", deviceMaxCanvasSize=max(w, h),
                        deviceMaxCanvasPixels=w * h, output="jpeg", jpeg_quality=90)
            assert (o.width, o.height) == (w, h), (o.width, o.height, w, h)
            file_bytes_list.append(o.jpeg)
            pf = C.c_void_p()
            assert lib.stitch_host_alloc(len(o.jpeg), C.byref(pf)) == 0
            C.memmove(pf, o.jpeg, len(o.jpeg))
            fbufs.append((pf, len(o.jpeg)))
            host_outs.append(pf)
        in_bytes = sum(b for _, b in fbufs)
        ctx.set_lanes(f_lanes)
        e2e_files = {"unit": "MP/s", "steps": files_steps, "h2d_bytes_per_step": in_bytes, "lanes": f_lanes,
                     "inputs": f"{len(dims)} baseline JPEG files, quality 90, 4:2:0 ({in_bytes / 1e6:.1f} MB), pinned host memory"}
        for kind, flag in (("png", L.OUT_PNG), ("jpeg", L.OUT_JPEG)):
            cap = rgba_bytes + (rgba_bytes >> 2) if kind == "png" else (64 << 20)
            fjobs = []
            for _ in range(f_lanes):
                arr = (L.StitchImage * len(dims))()
                for i, (w, h) in enumerate(dims):
                    arr[i].desc.natural_width, arr[i].desc.natural_height = w, h
                    arr[i].desc.width, arr[i].desc.height = w, h
                    arr[i].desc.orientation = 1
                    arr[i].desc.file_size = fbufs[i][1]
                    arr[i].data = fbufs[i][0]
                    arr[i].pitch_bytes = fbufs[i][1]
                    arr[i].mem_kind = L.MEM_HOST_JPEG
                po = C.c_void_p()
                assert lib.stitch_host_alloc(cap, C.byref(po)) == 0
                host_outs.append(po)
                fj = L.StitchJob()
                fj.images, fj.n = arr, len(dims)
                fj.params = hjobs[0][0].params
                fj.output_flags = flag
                fj.out_mem_kind = L.MEM_HOST
                if kind == "png":
                    fj.out_png, fj.out_capacity_png = po.value, cap
                else:
                    fj.out_jpeg, fj.out_capacity_jpeg, fj.jpeg_quality = po.value, cap, 90
                fjobs.append((fj, arr, L.StitchResult()))

            def files_loop(nsteps):
                tickets = []
                for i in range(nsteps):
                    fj, _, fr = fjobs[i % f_lanes]
                    t = C.c_uint64()
                    st = lib.stitch_submit(ctx.handle, C.byref(fj), C.byref(fr), C.byref(t))
                    assert st == 0, lib.stitch_last_error(ctx.handle)
                    tickets.append(t.value)
                    if i >= f_lanes - 1:
                        assert lib.stitch_wait(ctx.handle, tickets[i - (f_lanes - 1)]) == 0, lib.stitch_last_error(ctx.handle)
                assert lib.stitch_wait(ctx.handle, tickets[-1]) == 0, lib.stitch_last_error(ctx.handle)

            files_loop(2 * f_lanes)
            torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            files_loop(files_steps)
            torch.cuda.synchronize()
            fs = time.perf_counter() - t0
            t1 = time.perf_counter()
            for _ in range(3):
                assert lib.stitch_run(ctx.handle, C.byref(fjobs[0][0]), C.byref(fjobs[0][2])) == 0
            single_ms = (time.perf_counter() - t1) / 3 * 1e3
            if world > 1:
                t = torch.tensor([fs], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                fs = float(t.item())
            fr = fjobs[0][2]
            fbytes = int(fr.png_bytes if kind == "png" else fr.jpeg_bytes)
            e2e_files[kind] = {"value": mp * files_steps * world / fs, "ms_per_step": fs / files_steps * 1e3,
                               "d2h_bytes_per_step": fbytes, "single_job_latency_ms": single_ms}
        e2e_files["value"] = e2e_files["png"]["value"]   # the page exports PNG (P:1579)

        # PNG inputs: nine phone screenshots (1290 x 2796 RGBA PNG files, what a long-screenshot stitch is fed) -> PNG file
        try:
            shots = [synth_screenshot_png(1290, 2796, seed=i) for i in range(3)]
            shots = [shots[i % 3] for i in range(9)]
            sbufs = []
            for f in shots:
                pf = C.c_void_p()
                assert lib.stitch_host_alloc(len(f), C.byref(pf)) == 0
                C.memmove(pf, f, len(f))
                sbufs.append((pf, len(f)))
                host_outs.append(pf)
            sjobs = []
            s_cap = 0
            for _ in range(f_lanes):
                arr = (L.StitchImage * 9)()
                for i in range(9):
                    arr[i].desc.natural_width, arr[i].desc.natural_height = 1290, 2796
                    arr[i].desc.width, arr[i].desc.height = 1290, 2796
                    arr[i].desc.orientation = 1
                    arr[i].desc.file_size = sbufs[i][1]
                    arr[i].data = sbufs[i][0]
                    arr[i].pitch_bytes = sbufs[i][1]
                    arr[i].mem_kind = L.MEM_HOST_PNG
                sj = L.StitchJob()
                sj.images, sj.n = arr, 9
                sj.params = hjobs[0][0].params
                sj.output_flags = L.OUT_PNG
                sj.out_mem_kind = L.MEM_HOST
                if not s_cap:
                    lay = L.StitchLayout()
                    descs = (L.StitchImageDesc * 9)(*[arr[i].desc for i in range(9)])
                    assert lib.stitch_layout(descs, 9, C.byref(sj.params), C.byref(lay)) == 0
                    s_cap = int(lib.stitch_png_capacity(C.byref(lay)))
                    s_mp = lay.canvas_w * lay.canvas_h / 1e6
                po = C.c_void_p()
                assert lib.stitch_host_alloc(s_cap, C.byref(po)) == 0
                host_outs.append(po)
                sj.out_png, sj.out_capacity_png = po.value, s_cap
                sjobs.append((sj, arr, L.StitchResult()))

            def shots_loop(nsteps):
                tickets = []
                for i in range(nsteps):
                    sj, _, sr = sjobs[i % f_lanes]
                    t = C.c_uint64()
                    st = lib.stitch_submit(ctx.handle, C.byref(sj), C.byref(sr), C.byref(t))
                    assert st == 0, lib.stitch_last_error(ctx.handle)
                    tickets.append(t.value)
                    if i >= f_lanes - 1:
                        assert lib.stitch_wait(ctx.handle, tickets[i - (f_lanes - 1)]) == 0, lib.stitch_last_error(ctx.handle)
                assert lib.stitch_wait(ctx.handle, tickets[-1]) == 0, lib.stitch_last_error(ctx.handle)

            shots_loop(2 * f_lanes)
            torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            shots_loop(files_steps)
            torch.cuda.synchronize()
            ss = time.perf_counter() - t0
            t1 = time.perf_counter()
            for _ in range(3):
                assert lib.stitch_run(ctx.handle, C.byref(sjobs[0][0]), C.byref(sjobs[0][2])) == 0
            s_single = (time.perf_counter() - t1) / 3 * 1e3
            if world > 1:
                t = torch.tensor([ss], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ss = float(t.item())
            e2e_files["png_in"] = {"value": s_mp * files_steps * world / ss, "unit": "MP/s", "ms_per_step": ss / files_steps * 1e3,
                                   "single_job_latency_ms": s_single, "output_mp_per_step": s_mp,
                                   "h2d_bytes_per_step": sum(b for _, b in sbufs), "d2h_bytes_per_step": int(sjobs[0][2].png_bytes),
                                   "inputs": "9 screenshot-like PNG files, 1290x2796 RGBA, written by PIL (zlib level 6), pinned host memory; "
                                             "inflated, unfiltered, stitched and re-encoded as one PNG on the GPU"}
            shot_files = shots
        except Exception as e:  # PIL missing on the box, or the leg failed: report, do not hide the other figures
            e2e_files["png_in"] = {"value": None, "note": f"failed: {e}"}
            shot_files = None

        # progressive JPEG inputs: nine 1600 x 1200 files (web-sized: cameras write baseline files) -> stitched JPEG file.
        # The refinement scans of a file are decoded by ONE warp each (jpeg_prog.cuh), so a job's latency is that of its
        # longest scan chain; few steps, the leg is slow by construction.
        try:
            pw_, ph_ = 1600, 1200
            progs = [synth_progressive_jpeg(pw_, ph_, seed=i) for i in range(3)]
            progs = [progs[i % 3] for i in range(9)]
            pbufs = []
            for f in progs:
                pf = C.c_void_p()
                assert lib.stitch_host_alloc(len(f), C.byref(pf)) == 0
                C.memmove(pf, f, len(f))
                pbufs.append((pf, len(f)))
                host_outs.append(pf)
            # a job keeps 45 warps busy (one per refinement scan) for tens of milliseconds: more jobs in flight than the
            # other legs use cost nothing, so this leg takes every lane a context has
            p_lanes = 8
            ctx.set_lanes(p_lanes)
            pjobs = []
            p_mp = 0.0
            for _ in range(p_lanes):
                arr = (L.StitchImage * 9)()
                for i in range(9):
                    arr[i].desc.natural_width, arr[i].desc.natural_height = pw_, ph_
                    arr[i].desc.width, arr[i].desc.height = pw_, ph_
                    arr[i].desc.orientation = 1
                    arr[i].desc.file_size = pbufs[i][1]
                    arr[i].data = pbufs[i][0]
                    arr[i].pitch_bytes = pbufs[i][1]
                    arr[i].mem_kind = L.MEM_HOST_JPEG
                pj_ = L.StitchJob()
                pj_.images, pj_.n = arr, 9
                pj_.params = hjobs[0][0].params
                pj_.output_flags = L.OUT_JPEG
                pj_.out_mem_kind = L.MEM_HOST
                if not p_mp:
                    lay = L.StitchLayout()
                    descs = (L.StitchImageDesc * 9)(*[arr[i].desc for i in range(9)])
                    assert lib.stitch_layout(descs, 9, C.byref(pj_.params), C.byref(lay)) == 0
                    p_mp = lay.canvas_w * lay.canvas_h / 1e6
                po = C.c_void_p()
                assert lib.stitch_host_alloc(32 << 20, C.byref(po)) == 0
                host_outs.append(po)
                pj_.out_jpeg, pj_.out_capacity_jpeg, pj_.jpeg_quality = po.value, 32 << 20, 90
                pjobs.append((pj_, arr, L.StitchResult()))

            def prog_loop(nsteps):
                tickets = []
                for i in range(nsteps):
                    pj_, _, pr = pjobs[i % p_lanes]
                    t = C.c_uint64()
                    st = lib.stitch_submit(ctx.handle, C.byref(pj_), C.byref(pr), C.byref(t))
                    assert st == 0, lib.stitch_last_error(ctx.handle)
                    tickets.append(t.value)
                    if i >= p_lanes - 1:
                        assert lib.stitch_wait(ctx.handle, tickets[i - (p_lanes - 1)]) == 0, lib.stitch_last_error(ctx.handle)
                for t_ in tickets[-(p_lanes - 1):]:
                    assert lib.stitch_wait(ctx.handle, t_) == 0, lib.stitch_last_error(ctx.handle)

            p_steps = min(files_steps, 3 * p_lanes)
            prog_loop(p_lanes)
            torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            prog_loop(p_steps)
            torch.cuda.synchronize()
            ps_ = time.perf_counter() - t0
            t1 = time.perf_counter()
            assert lib.stitch_run(ctx.handle, C.byref(pjobs[0][0]), C.byref(pjobs[0][2])) == 0
            p_single = (time.perf_counter() - t1) * 1e3
            if world > 1:
                t = torch.tensor([ps_], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ps_ = float(t.item())
            ctx.set_lanes(f_lanes)
            e2e_files["prog_in"] = {"value": p_mp * p_steps * world / ps_, "unit": "MP/s", "steps": p_steps, "lanes": p_lanes,
                                    "ms_per_step": ps_ / p_steps * 1e3, "single_job_latency_ms": p_single, "output_mp_per_step": p_mp,
                                    "h2d_bytes_per_step": sum(b for _, b in pbufs), "d2h_bytes_per_step": int(pjobs[0][2].jpeg_bytes),
                                    "inputs": "9 progressive JPEG files (SOF2, libjpeg's 10-scan script), 1600x1200, quality 90, 4:2:0, "
                                              "written by PIL, pinned host memory; first scans decoded in parallel, refinement "
                                              "scans one warp each (latency-bound by construction), stitched, JPEG file out"}
            prog_files = progs
        except Exception as e:
            e2e_files["prog_in"] = {"value": None, "note": f"failed: {e}"}
            prog_files = None
            try:
                ctx.set_lanes(f_lanes)
            except Exception:
                pass

    if has_jpeg or has_png:
        out_bytes = int(hres.jpeg_bytes if has_jpeg else hres.png_bytes)   # d2h of the e2e step = the file
        algo_bytes = src_bytes + out_bytes
    peak, peak_src = _peaks()
    achieved = algo_bytes / (ms_per_step * 1e-3) / 1e9   # one compose launch per step
    line = {
        "metric": "output MP/s, 9x12MP vertical stitch" if args.workload.startswith("cfg2") or args.workload == "cfg5" else "output MP/s",
        "value": value, "unit": "MP/s", "n_gpus": world, "steps": args.steps, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 in/out, f32 accumulate", "data": "synthetic",
        "config": {"workload": desc, "name": args.workload, "canvas": [cw, ch], "output_mp_per_step": mp,
                   "inputs": "device-resident RGBA8, generated on device (seed 0x5717C4)",
                   "l2": f"{len(jobs)} job slots rotated; each job reads {src_bytes/1e6:.0f} MB (> 126 MB L2)",
                   "parallelism": f"{world} rank(s), jobs sharded, no collective",
                   "lanes": f"{max(1, args.lanes)} job lane(s) per ctx for the device-resident steps, {e2e_lanes} for e2e, {max(1, min(8, args.files_lanes))} for e2e_files",
                   "host_placement": placement, "prewarm_jobs": PREWARM_JOBS},
        "gpu_launches": int(launches),
        "host_us_per_submit": round(host_s[0] / max(1, args.steps) * 1e6, 2),   # layout + plan + launch inside stitch_submit
        "clocks": sampler.summary(),
        "e2e": {"value": e2e_value, "unit": "MP/s", "h2d_bytes_per_step": src_bytes, "d2h_bytes_per_step": out_bytes,
                "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3, "single_job_latency_ms": e2e_single_ms,
                "single_job_latency_pageable_ms": pageable_ms},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": algo_bytes,
                     "frac_of_8TBps_nominal": achieved / 8000.0,
                     "launches_per_step": launches / max(1, args.steps)},
    }
    if e2e_files is not None:
        line["e2e_files"] = e2e_files
    if digests is not None:
        line["digests"] = digests
    prof_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof_path):
        try:
            line["roofline"]["traffic"] = json.load(open(prof_path)).get("cfg2" if args.workload == "cfg5" else args.workload)
        except Exception:
            pass
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            os.sched_setaffinity(0, affinity0)   # the CPU leg uses every host core again
            val, ms, cores, _ = cpu_reference_run(args.workload, max(1, args.cpu_steps), 1)
            line["cpu_baseline"] = {"value": val, "unit": "MP/s", "cores": cores, "kind": "port",
                                    "sample": f"{max(1, args.cpu_steps)} full job(s), C restatement fp32 + OpenMP ({ms:.0f} ms/job)",
                                    "note": "a scalar, un-vectorised port with a float canvas round trip: a soft baseline, stated "
                                            "and never the target -- the roofline fraction is what measures the kernel"}
            line["cpu_baseline"].update(cpu_sanity_baselines(args.workload))
        except Exception as e:  # the baseline is a report, never a blocker
            line["cpu_baseline"] = {"value": None, "unit": "MP/s", "cores": 0, "kind": "port", "sample": f"failed: {e}"}
        if e2e_files is not None:
            try:
                v, ms, cores = cpu_files_run(file_bytes_list, direction, mode, platform, 1, "png")
                e2e_files["cpu_baseline"] = {"value": v, "unit": "MP/s", "cores": cores, "kind": "port", "ms_per_step": ms,
                                             "sample": "1 job: oracle JPEG decoder (a thread per file) + C raster (OpenMP) + "
                                                       "oracle PNG writer (sequential)"}
            except Exception as e:
                e2e_files["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
            if isinstance(e2e_files.get("png_in"), dict) and e2e_files["png_in"].get("value") and shot_files:
                try:
                    v, ms, cores = cpu_png_files_run(shot_files, direction, mode, platform, 1)
                    e2e_files["png_in"]["cpu_baseline"] = {"value": v, "unit": "MP/s", "cores": cores, "kind": "port", "ms_per_step": ms,
                                                           "sample": "1 job: oracle PNG decoder (a thread per file) + C raster (OpenMP) + "
                                                                     "oracle PNG writer (sequential)"}
                except Exception as e:
                    e2e_files["png_in"]["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
            if isinstance(e2e_files.get("prog_in"), dict) and e2e_files["prog_in"].get("value") and prog_files:
                try:
                    v, ms, cores = cpu_files_run(prog_files, direction, mode, platform, 1, "jpeg")
                    e2e_files["prog_in"]["cpu_baseline"] = {"value": v, "unit": "MP/s", "cores": cores, "kind": "port", "ms_per_step": ms,
                                                            "sample": "1 job: oracle progressive JPEG decoder (a thread per file) + C raster "
                                                                      "(OpenMP) + oracle JPEG writer (sequential)"}
                except Exception as e:
                    e2e_files["prog_in"]["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
    if rank == 0:
        print(json.dumps(line))
    for p in host_keep + host_outs:
        if p:
            lib.stitch_host_free(p)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
