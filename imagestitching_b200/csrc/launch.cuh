// Host side of the draw loop: partition the canvas into regions (stitch_plan.h) and launch.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <vector>

#include "compose_kernel.cuh"
#include "gather_kernel.cuh"
#include "stitch_plan.h"
#include "tma_maps.h"

namespace stitch {

struct StreamScratch {
  void* p = nullptr;
  size_t cap = 0;
  TmaCache tma;   // tensor maps of the bitmaps this lane has streamed
};
inline void free_scratch(StreamScratch& s) {
  if (s.p) cudaFree(s.p);
  s.p = nullptr;
  s.cap = 0;
}

inline bool is_int(double v) { return std::floor(v) == v && std::fabs(v) < 1e9; }

inline bool stream_eligible(const ImgParam& im) {
  if (im.transpose) return false;
  const int64_t dw = (int64_t)im.W, dh = (int64_t)im.H;
  if (!(im.sx_len > dw && im.sy_len > dh)) return false;       // area filter on both axes
  if ((int64_t)im.sx_len > 16 * dw) return false;              // tails are collected over at most 4 groups: ratio_x <= 16
  if (((uintptr_t)im.data & 15) || (im.pitch & 15)) return false;  // 128-bit row loads
  // the kernel's geometry is unsigned 32-bit: every scaled coordinate it forms must stay below 2^31
  if (((int64_t)im.sx_len + 512) * (dw + 1) >= (1LL << 31) || ((int64_t)im.sy_len + 8) * (dh + 1) >= (1LL << 31)) return false;
  if (im.pitch * (int64_t)STREAM_TAB_MAX >= (1LL << 31)) return false;  // 32-bit row offsets inside a unit
  // at least one output row (plus unroll padding) must fit the per-warp row table
  if (im.sy_len / dh + 4 + STREAM_UNR > STREAM_TAB_MAX) return false;
  return true;
}

// Bilinear on both axes (device rect at least as large as the bitmap), EXIF 1-4: upscale_unit.cuh.
inline bool upscale_eligible(const ImgParam& im) {
  if (im.transpose) return false;
  if (!(im.W >= (double)im.sx_len && im.H >= (double)im.sy_len)) return false;
  if (im.sx_len < 5) return false;   // a lane loads a window of up to 5 source pixels per row
  if (((uintptr_t)im.data & 3) || (im.pitch & 3)) return false;
  if (!(std::fabs(im.X0) < 1e8 && std::fabs(im.Y0) < 1e8 && im.W < 1e8 && im.H < 1e8)) return false;
  return true;
}

// Largest tile_h (output rows per unit) whose source rows fit the row table.
inline int stream_max_tile_h(const ImgParam& im) {
  const int64_t dh = (int64_t)im.H;
  int64_t th = ((int64_t)(STREAM_TAB_MAX - 4 - STREAM_UNR) * dh) / im.sy_len;
  return (int)std::max<int64_t>(1, std::min<int64_t>(th, 1 << 20));
}

// Returns true and fills `plan` when the fused compose launch applies to this job: every device
// rect integer-aligned (superSample == 1 and gap*scaleDown integral, true for BASELINE configs 2-5)
// and the clipped rects pairwise disjoint, so painter's-order source-over degenerates to each
// pixel belonging to at most one image.
//
// mode = STREAM_MODE_*.  With chroma output the partition is made on the grid of 2x2 chroma blocks
// (cell size 2): a streamed image keeps the largest block-aligned box inside its rect and the odd
// rows / columns along its edges -- blocks shared with a neighbour or the background -- go to
// per-block GATHER regions, so no chroma block is ever split between two units.
inline bool build_plan(const JobParam& P, int mode, int hops, int resident_warps, Plan* plan) {
  const int cs = (mode == STREAM_MODE_RGBA) ? 1 : 2;                  // cell size in pixels
  const int stride = STREAM_STRIDE_OF(mode, hops);
  const int ccw = (P.cw + cs - 1) / cs, cch = (P.ch + cs - 1) / cs;   // canvas in cells
  // Per image: the OUTER box (every pixel the device rect touches) and the FULL box (pixels it covers
  // completely); both also on the cell grid.  A fast region (STREAM / UPSCALE) is the full box in whole
  // cells; it needs the image to be alone there, i.e. no other image's outer box may reach into it --
  // otherwise painter's-order compositing matters and the image is left to the gather pieces.
  struct Box { int x0, y0, x1, y1, img; int kind; int ix0, iy0, ix1, iy1, ox0, oy0, ox1, oy1; };
  std::vector<Box> boxes;
  bool any_fast = false;
  for (int k = 0; k < P.n; ++k) {
    const ImgParam& im = P.img[k];
    if (im.x_hi <= im.x_lo || im.y_hi <= im.y_lo) continue;  // nothing visible
    Box b{};
    b.x0 = im.x_lo; b.y0 = im.y_lo; b.x1 = im.x_hi; b.y1 = im.y_hi; b.img = k;
    // fully covered pixels, clipped to the canvas
    const int fx0 = (int)std::fmin(std::fmax(std::ceil(im.X0), 0.0), (double)P.cw);
    const int fy0 = (int)std::fmin(std::fmax(std::ceil(im.Y0), 0.0), (double)P.ch);
    const int fx1 = (int)std::fmin(std::fmax(std::floor(im.X0 + im.W), 0.0), (double)P.cw);
    const int fy1 = (int)std::fmin(std::fmax(std::floor(im.Y0 + im.H), 0.0), (double)P.ch);
    b.ix0 = (fx0 + cs - 1) / cs; b.iy0 = (fy0 + cs - 1) / cs; b.ix1 = fx1 / cs; b.iy1 = fy1 / cs;         // full, in cells
    b.ox0 = b.x0 / cs; b.oy0 = b.y0 / cs; b.ox1 = (b.x1 + cs - 1) / cs; b.oy1 = (b.y1 + cs - 1) / cs;   // outer, in cells
    const bool integral = is_int(im.X0) && is_int(im.Y0) && is_int(im.W) && is_int(im.H);
    b.kind = REGION_GATHER;
    if (b.ix0 < b.ix1 && b.iy0 < b.iy1) {
      if (integral && stream_eligible(im) && (cs == 1 || stream_max_tile_h(im) >= 2)) b.kind = REGION_STREAM;
      else if (upscale_eligible(im)) b.kind = REGION_UPSCALE;
    }
    boxes.push_back(b);
  }
  for (size_t a = 0; a < boxes.size(); ++a) {
    if (boxes[a].kind == REGION_GATHER) continue;
    for (size_t b = 0; b < boxes.size(); ++b) {
      if (a == b) continue;
      const Box &A = boxes[a], &B = boxes[b];
      if (A.ix0 < B.ox1 && B.ox0 < A.ix1 && A.iy0 < B.oy1 && B.oy0 < A.iy1) { boxes[a].kind = REGION_GATHER; break; }
    }
  }
  for (const Box& b : boxes) any_fast = any_fast || b.kind != REGION_GATHER;
  if (!any_fast) return false;   // nothing to fuse: the whole-canvas gather kernel does the job

  std::vector<Region> regs;
  // streaming regions: the inner box of every eligible image
  for (const Box& bx : boxes) {
    if (bx.kind != REGION_STREAM) continue;
    const ImgParam& im = P.img[bx.img];
    Region r{};
    r.img = bx.img;
    const int dw = (int)im.W, dh = (int)im.H, X0 = (int)im.X0, Y0 = (int)im.Y0;
    int rx0 = bx.ix0 * cs - X0, rx1 = bx.ix1 * cs - X0, ry0 = bx.iy0 * cs - Y0, ry1 = bx.iy1 * cs - Y0;  // device-relative
    if (im.flip_x) { int t = dw - rx1; rx1 = dw - rx0; rx0 = t; }
    if (im.flip_y) { int t = dh - ry1; ry1 = dh - ry0; ry0 = t; }
    r.kind = REGION_STREAM;
    r.x0 = rx0; r.x1 = rx1; r.y0 = ry0; r.y1 = ry1;
    // bands in group space: unit tx stores the outputs that start in groups g_base + stride*tx ...
    const int64_t sw = im.sx_len;
    r.g_base = (int)((((int64_t)rx0 * sw) / dw) >> 2);
    const int g_end = (int)std::min<int64_t>(((int64_t)rx1 * sw + 4 * dw - 1) / (4 * dw), (sw + 3) >> 2);
    r.tile_w = 0;
    r.tiles_x = std::max(1, (g_end - r.g_base + stride - 1) / stride);
    r.tile_h = stream_max_tile_h(im);   // refined below for load balance
    regs.push_back(r);
  }
  // enlarged images: the full box in device pixels, 32 aligned quads x UPSCALE_TILE_H rows per unit
  for (const Box& bx : boxes) {
    if (bx.kind != REGION_UPSCALE) continue;
    Region r{};
    r.kind = REGION_UPSCALE;
    r.img = bx.img;
    r.x0 = bx.ix0 * cs; r.x1 = bx.ix1 * cs; r.y0 = bx.iy0 * cs; r.y1 = bx.iy1 * cs;
    r.tile_w = UPSCALE_TILE_W;
    r.tile_h = UPSCALE_TILE_H;
    r.g_base = (P.img[r.img].W / (double)P.img[r.img].sx_len >= 1.5) ? 4 : 5;   // source window per lane
    r.tiles_x = (r.x1 - (r.x0 & ~3) + UPSCALE_TILE_W - 1) / UPSCALE_TILE_W;
    regs.push_back(r);
  }
  // the rest of the canvas, in cells: sweep the bands between all y edges; inside a band cut at all x
  // edges; a piece is GATHER when some image's rect reaches into it, FILL (white, P:1423-1424) otherwise
  std::vector<int> ys{0, cch};
  for (const Box& b : boxes) { ys.push_back(b.oy0); ys.push_back(b.oy1); ys.push_back(b.iy0); ys.push_back(b.iy1); }
  std::sort(ys.begin(), ys.end());
  ys.erase(std::unique(ys.begin(), ys.end()), ys.end());
  std::vector<Region> rest;
  for (size_t i = 0; i + 1 < ys.size(); ++i) {
    const int ya = ys[i], yb = ys[i + 1];
    if (ya < 0 || yb > cch || ya >= yb) continue;
    std::vector<int> xs{0, ccw};
    for (const Box& b : boxes)
      if (b.oy0 < yb && b.oy1 > ya) { xs.push_back(b.ox0); xs.push_back(b.ox1); xs.push_back(b.ix0); xs.push_back(b.ix1); }
    std::sort(xs.begin(), xs.end());
    xs.erase(std::unique(xs.begin(), xs.end()), xs.end());
    int run_kind = -1, run_x0 = 0, run_x1 = 0;
    auto flush = [&]() {
      if (run_kind < 0) return;
      // merge with a region of the same kind and x-range that ends where this band starts
      for (Region& f : rest)
        if (f.kind == run_kind && f.x0 == run_x0 && f.x1 == run_x1 && f.y1 == ya) { f.y1 = yb; run_kind = -1; return; }
      Region r{};
      r.kind = run_kind; r.img = -1;
      r.x0 = run_x0; r.x1 = run_x1; r.y0 = ya; r.y1 = yb;
      if (run_kind == REGION_FILL) { r.tile_w = 256; r.tile_h = 8; }
      else { r.tile_w = 32; r.tile_h = (cs == 1) ? 8 : 4; }
      rest.push_back(r);
      run_kind = -1;
    };
    for (size_t j = 0; j + 1 < xs.size(); ++j) {
      const int xa = xs[j], xb = xs[j + 1];
      if (xa < 0 || xb > ccw || xa >= xb) continue;
      int kind = REGION_FILL;
      for (const Box& b : boxes) {
        if (b.kind != REGION_GATHER && b.ix0 <= xa && b.ix1 >= xb && b.iy0 <= ya && b.iy1 >= yb) { kind = -1; break; }  // fast region
        if (b.ox0 < xb && b.ox1 > xa && b.oy0 < yb && b.oy1 > ya) kind = REGION_GATHER;
      }
      if (kind != run_kind || xa != run_x1) { flush(); run_kind = kind; run_x0 = xa; }
      run_x1 = xb;
    }
    flush();
  }
  for (Region& f : rest) regs.push_back(f);
  if (regs.size() > STITCH_MAX_REGIONS) return false;

  // Unit heights of the streaming regions -- a "guided" schedule.  Units are handed out in order to
  // `resident_warps` persistent warps.  Each image's rows are split into a BODY of large units
  // (about one per warp, so the per-unit setup -- row table, lane weights, the empty TMA ring -- is
  // amortised over the most rows) followed by a TAIL of smaller units that even out the different
  // speeds of the SMs (near / far L2 and HBM); all tails are queued after all bodies.
  // Tail units are 128 source rows: with the TMA ring and the integer vertical pass a unit's fixed cost
  // (~1.5 us until its first stage lands) weighs more than in round 1, when 32-row tails were best --
  // 9x12 MP job, two job lanes: 89.5 us with 32-row tails, 81.2 with 96, 80.4 with 128, 79.4 with 216
  // (profiles/r2_sched_sweep.md); one job at a time is 90 us either way.
  // tuning knobs (profiles/microbench sweeps); clamped so that no value can push a unit outside its region
  static const double body_frac = [] { const char* e = std::getenv("STITCH_BODY_FRAC"); return e ? std::min(1.0, std::max(0.0, atof(e))) : 0.75; }();
  static const int tail_rows = [] { const char* e = std::getenv("STITCH_TAIL_ROWS"); return e ? std::max(1, atoi(e)) : 128; }();
  static const double up_body_frac = [] { const char* e = std::getenv("STITCH_UP_BODY_FRAC"); return e ? std::min(1.0, std::max(0.0, atof(e))) : 0.8; }();
  static const int up_tail_h = [] { const char* e = std::getenv("STITCH_UP_TAIL_H"); return e ? std::min(32, std::max(2, atoi(e) & ~1)) : UPSCALE_TAIL_H; }();
  const int max_src_rows = STREAM_TAB_MAX - 4 - STREAM_UNR;
  auto tile_h_for = [&](const Region& r, int src_rows) {
    const ImgParam& im = P.img[r.img];
    const int64_t h = ((int64_t)src_rows * (int64_t)im.H) / im.sy_len;
    const int th = (int)std::max<int64_t>(1, std::min<int64_t>(h, stream_max_tile_h(im)));
    return cs == 1 ? th : std::max(2, th & ~1);   // chroma rows come in pairs
  };
  auto body_rows_of = [&](const Region& r, int th) {   // output rows of the region that go to whole body units
    const int64_t H = r.y1 - r.y0;
    return (int)((int64_t)(body_frac * (double)H) / th * th);
  };
  auto count_body = [&](int src_rows) {
    int64_t n = 0;
    for (const Region& r : regs)
      if (r.kind == REGION_STREAM) {
        const int th = tile_h_for(r, src_rows);
        n += (int64_t)r.tiles_x * (body_rows_of(r, th) / th);
      }
    return n;
  };
  int body_src_rows = max_src_rows;
  if (resident_warps > 0) {
    for (int k = 1; k <= 8; ++k) {
      // smallest src_rows with body units <= k * warps (non-increasing in src_rows up to rounding)
      if (count_body(max_src_rows) > (int64_t)k * resident_warps) continue;  // table caps the height: more units per warp
      int lo = 1, hi = max_src_rows;
      while (lo < hi) {
        const int mid = (lo + hi) / 2;
        if (count_body(mid) <= (int64_t)k * resident_warps) hi = mid; else lo = mid + 1;
      }
      body_src_rows = lo;
      break;
    }
  }
  {
    std::vector<Region> bodies, tails, others;
    for (const Region& r0 : regs) {
      if (r0.kind != REGION_STREAM) { others.push_back(r0); continue; }
      const int th_b = tile_h_for(r0, body_src_rows);
      const int th_t = tile_h_for(r0, std::min(tail_rows, body_src_rows));
      const int hb = (th_b > th_t) ? std::min(body_rows_of(r0, th_b), r0.y1 - r0.y0) : 0;   // no point in a body of tail-sized units
      if (hb > 0) {
        Region b = r0;
        b.y1 = r0.y0 + hb;
        b.tile_h = th_b;
        bodies.push_back(b);
      }
      if (r0.y0 + hb < r0.y1) {
        Region t = r0;
        t.y0 = r0.y0 + hb;
        t.tile_h = th_t;
        tails.push_back(t);
      }
    }
    regs.clear();
    // order = hand-out order: bodies, then the per-pixel gather regions, then the small units
    for (const Region& r : bodies) regs.push_back(r);
    // enlarged images get the same treatment: full-height units over the first rows, quarter-height
    // units over the rest, so the last round of units is short (configs[0]: 2.1 full units per warp)
    for (const Region& r : others) {
      if (r.kind != REGION_UPSCALE) continue;
      const int hb = std::min(r.y1 - r.y0, (int)((int64_t)(up_body_frac * (double)(r.y1 - r.y0)) / UPSCALE_TILE_H * UPSCALE_TILE_H));
      if (hb > 0) {
        Region b = r;
        b.y1 = r.y0 + hb;
        regs.push_back(b);
      }
      if (r.y0 + hb < r.y1) {
        Region t = r;
        t.y0 = r.y0 + hb;
        t.tile_h = up_tail_h;
        tails.push_back(t);
      }
    }
    for (const Region& r : others) if (r.kind == REGION_GATHER) regs.push_back(r);
    for (const Region& r : tails) regs.push_back(r);
    for (const Region& r : others) if (r.kind == REGION_FILL) regs.push_back(r);
    if (regs.size() > STITCH_MAX_REGIONS) return false;
  }
  int unit = 0;
  plan->n_regions = (int)regs.size();
  for (size_t i = 0; i < regs.size(); ++i) {
    Region& r = regs[i];
    if (r.kind == REGION_FILL || r.kind == REGION_GATHER) r.tiles_x = (r.x1 - r.x0 + r.tile_w - 1) / r.tile_w;
    r.tiles_y = (r.y1 - r.y0 + r.tile_h - 1) / r.tile_h;
    r.first_unit = unit;
    unit += r.tiles_x * r.tiles_y;
    plan->r[i] = r;
  }
  plan->n_units = unit;
  return unit > 0;
}

static constexpr int kMaxDevices = 64;

template <int MAXO, int HOPS, int MODE>
struct ComposeLaunch {
  // resident CTAs per SM of this instantiation, 0 on error.  Queried once PER DEVICE: the opt-in to more than 48 KB
  // of dynamic shared memory is a per-device function attribute, and one process may hold contexts on several GPUs.
  static int ctas_per_sm(cudaError_t* err) {
    static std::atomic<int> value[kMaxDevices];   // 0 = not queried yet
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) { *err = e != cudaSuccess ? e : cudaErrorInvalidDevice; return 0; }
    int v = value[dev].load(std::memory_order_acquire);
    if (v == 0) {
      e = cudaFuncSetAttribute(compose_kernel<MAXO, HOPS, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kComposeSmemBytes);
      int n = 0;
      if (e == cudaSuccess)
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, compose_kernel<MAXO, HOPS, MODE>, STREAM_WARPS * 32,
                                                          kComposeSmemBytes);
      if (e != cudaSuccess) { *err = e; return 0; }
      v = n < 1 ? 1 : n;
      value[dev].store(v, std::memory_order_release);
    }
    return v;
  }
  static cudaError_t launch(const JobParam& P, const Plan& plan, const TmaMaps& maps, int grid, unsigned int* sched,
                            cudaStream_t stream) {
    compose_kernel<MAXO, HOPS, MODE><<<grid, STREAM_WARPS * 32, kComposeSmemBytes, stream>>>(P, plan, maps, sched);
    return cudaGetLastError();
  }
};

template <int MODE>
struct LightLaunch {
  static int ctas_per_sm(cudaError_t* err) {
    static std::atomic<int> value{0};   // no per-device attribute involved: the occupancy is a property of the cubin
    int v = value.load(std::memory_order_acquire);
    if (v == 0) {
      int n = 0;
      const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, compose_light_kernel<MODE>, STREAM_WARPS * 32, 0);
      if (e != cudaSuccess) { *err = e; return 0; }
      v = n < 1 ? 1 : n;
      value.store(v, std::memory_order_release);
    }
    return v;
  }
  static cudaError_t launch(const JobParam& P, const Plan& plan, const TmaMaps&, int grid, unsigned int* sched, cudaStream_t stream) {
    compose_light_kernel<MODE><<<grid, STREAM_WARPS * 32, 0, stream>>>(P, plan, sched);
    return cudaGetLastError();
  }
};

#define STITCH_DISPATCH(M, H, Md, expr) \
  if (maxo == M && hops == H && mode == Md) { using K = ComposeLaunch<M, H, Md>; expr; }
#define STITCH_SHAPES(Md, expr)                                                               \
  STITCH_DISPATCH(2, 1, Md, expr) STITCH_DISPATCH(3, 1, Md, expr) STITCH_DISPATCH(4, 1, Md, expr) \
  STITCH_DISPATCH(2, 2, Md, expr) STITCH_DISPATCH(3, 2, Md, expr) STITCH_DISPATCH(4, 2, Md, expr) \
  STITCH_DISPATCH(4, 4, Md, expr)
#define STITCH_ALL(expr) STITCH_SHAPES(0, expr) STITCH_SHAPES(1, expr) STITCH_SHAPES(2, expr)
#define STITCH_LIGHT(expr)                                   \
  if (mode == 0) { using K = LightLaunch<0>; expr; }          \
  else if (mode == 1) { using K = LightLaunch<1>; expr; }     \
  else { using K = LightLaunch<2>; expr; }

// Everything about a job's draw that is decided on the host: the kernel shape, the fused plan, the tensor maps.
struct PreparedJob {
  bool fused = false;        // false: no fast region at all, the whole-canvas gather kernel draws the job
  int mode = 0, maxo = 2, hops = 1, per_sm = 0;
  bool any_stream = false;
  int path_info = 1 << 24;
  Plan plan;
  TmaMaps maps;
};

// path_info: streaming | gather << 8 | fill << 16 | upscale << 25 region counts of the fused plan, or 1 << 24 when the
// whole canvas takes the per-pixel gather kernel (every image transposed, overlapped or unaligned).
inline int prepare_job(const JobParam& P, bool want_rgba, bool want_yuv, int sm_count, StreamScratch& scratch, cudaStream_t stream,
                       PreparedJob* J) {
  const int mode = want_yuv ? (want_rgba ? STREAM_MODE_BOTH : STREAM_MODE_YUV) : STREAM_MODE_RGBA;
  // kernel variant: decided by the images alone (before the unit grid is chosen)
  int maxo = 2, hops = 1;
  bool any_stream = false;   // no streaming candidate at all -> the light kernel (no smem, more resident warps)
  for (int k = 0; k < P.n; ++k) {
    const ImgParam& im = P.img[k];
    if (!is_int(im.X0) || !is_int(im.Y0) || !is_int(im.W) || !is_int(im.H) || im.W < 1 || im.H < 1 || !stream_eligible(im)) continue;
    any_stream = true;
    const int64_t dw = (int64_t)im.W;
    maxo = std::max(maxo, (int)std::min<int64_t>(4, (4 * dw + im.sx_len - 1) / im.sx_len));
    hops = std::max(hops, (int)((im.sx_len + 4 * dw - 1) / (4 * dw)));
  }
  // instantiated (MAXO, HOPS) shapes: (2|3|4, 1) for ratios <= 4, (2, 2) for 2 < ratio <= 8, (3, 2) / (4, 2) for jobs that mix
  // ratios from 4/3 (or 1) up to 8 (mixed-aspect stitches to a common height: BASELINE configs[2] in 'min' mode), (4, 4) for the rest
  if (hops > 2) { maxo = 4; hops = 4; }
  cudaError_t err = cudaSuccess;
  int per_sm = 0;
  if (any_stream) { STITCH_ALL(per_sm = K::ctas_per_sm(&err)) }
  else { STITCH_LIGHT(per_sm = K::ctas_per_sm(&err)) }
  if (err != cudaSuccess) return (int)err;
  J->mode = mode; J->maxo = maxo; J->hops = hops; J->per_sm = per_sm; J->any_stream = any_stream;
  J->path_info = 1 << 24;
  J->fused = build_plan(P, mode, hops, sm_count * per_sm * STREAM_WARPS, &J->plan);
  if (!J->fused) return 0;
  if (!scratch.p) {  // unit scheduler counters {next unit, warps retired}; the kernel leaves them at zero
    cudaError_t e = cudaMalloc(&scratch.p, 256);
    if (e != cudaSuccess) return (int)e;
    scratch.cap = 256;
    e = cudaMemsetAsync(scratch.p, 0, 256, stream);
    if (e != cudaSuccess) return (int)e;
  }
  const Plan& plan = J->plan;
  int cnt[4] = {0, 0, 0, 0};
  uint32_t up_seen = 0;
  for (int i = 0; i < plan.n_regions; ++i) {
    const Region& r = plan.r[i];
    if (r.kind == REGION_UPSCALE) {  // body + tail of one enlarged image count once
      if (up_seen >> r.img & 1u) continue;
      up_seen |= 1u << r.img;
    }
    cnt[r.kind]++;
  }
  J->path_info = std::min(cnt[REGION_STREAM], 255) | (std::min(cnt[REGION_GATHER], 255) << 8) |
                 (std::min(cnt[REGION_FILL], 255) << 16) | (std::min(cnt[REGION_UPSCALE], 63) << 25);
  // tensor maps of the bitmaps that have a streaming region (the other slots are never dereferenced)
  uint32_t mapped = 0;
  for (int i = 0; i < plan.n_regions; ++i) {
    const Region& r = plan.r[i];
    if (r.kind != REGION_STREAM || (mapped >> r.img & 1u)) continue;
    mapped |= 1u << r.img;
    const ImgParam& im = P.img[r.img];
    const cudaError_t e = scratch.tma.get(im.data, im.pitch, im.bw, im.bh, &J->maps.box[r.img], &J->maps.row[r.img]);
    if (e != cudaSuccess) return (int)e;
  }
  return 0;
}

// One compose launch over `plan`: the job's whole plan, or a subset of its regions (sub_plan below).
inline int launch_plan(const JobParam& P, const PreparedJob& J, const Plan& plan, int sm_count, StreamScratch& scratch,
                       cudaStream_t stream) {
  const int mode = J.mode, maxo = J.maxo, hops = J.hops;
  // persistent grid: every resident warp slot of the chip, capped by the number of units
  int grid = sm_count * J.per_sm;
  const int need = (plan.n_units + STREAM_WARPS - 1) / STREAM_WARPS;
  if (grid > need) grid = need;
  unsigned int* sched = (unsigned int*)scratch.p;
  cudaError_t err = cudaSuccess;
  if (J.any_stream) { STITCH_ALL(err = K::launch(P, plan, J.maps, grid, sched, stream)) }
  else { STITCH_LIGHT(err = K::launch(P, plan, J.maps, grid, sched, stream)) }
  return (int)err;
}

// The regions of `plan` that `take` accepts, renumbered: same units, same unit geometry -- so drawing a job as
// several launches over disjoint subsets gives the same bytes as the one fused launch.
template <class Pred>
inline void sub_plan(const Plan& plan, Pred take, Plan* out) {
  int unit = 0, n = 0;
  for (int i = 0; i < plan.n_regions; ++i) {
    if (!take(plan.r[i])) continue;
    Region r = plan.r[i];
    r.first_unit = unit;
    unit += r.tiles_x * r.tiles_y;
    out->r[n++] = r;
  }
  out->n_regions = n;
  out->n_units = unit;
}

// The draw loop of one job as ONE launch.
inline int launch_job(const JobParam& P, bool want_rgba, bool want_yuv, int sm_count, StreamScratch& scratch,
                      cudaStream_t stream, int* launches, int* path_info) {
  *launches = 0;
  PreparedJob J;
  int st = prepare_job(P, want_rgba, want_yuv, sm_count, scratch, stream, &J);
  *path_info = J.path_info;
  if (st != 0) return st;
  *launches = 1;
  if (J.fused) return launch_plan(P, J, J.plan, sm_count, scratch, stream);
  if (want_yuv) {
    dim3 grid(((P.cw + 1) / 2 + 31) / 32, ((P.ch + 1) / 2 + 7) / 8);
    gather_yuv_kernel<<<grid, 256, 0, stream>>>(P);
  } else {
    dim3 grid((P.cw + 31) / 32, (P.ch + 7) / 8);
    gather_rgba_kernel<<<grid, 256, 0, stream>>>(P);
  }
  return (int)cudaGetLastError();
}

}  // namespace stitch
