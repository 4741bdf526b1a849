/* C restatement of the stated AB1 filter: fast checker (fp64) and timed CPU baseline (fp32).
 *
 * TEST INFRASTRUCTURE ONLY -- never linked into or called from libstitch.so.
 * PARITY UNPINNED for pixels: the reference's resampler is the closed WeChat
 * Canvas2D ctx.drawImage (utils/canvas.js:156-199); this restates the stated
 * filter of SURVEY.md App. C, identical to oracle/raster_ref.py (numpy), which
 * tests/test_oracle_raster.py checks it against.
 *
 * Reference lines restated (under /root/reference/miniprogram-stitch/miniprogram/):
 *   white background pages/index/index.js:1423-1424; painter's order :1439-1554;
 *   orientation placement utils/canvas.js:153-202 (resolved by the caller into
 *   X0,Y0,W,H,transpose,flip_u,flip_v -- see oracle/layout_ref.py placement_for);
 *   export of the whole canvas utils/canvas.js:206-209.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  const uint8_t* data;
  int64_t pitch;
  int32_t bw, bh;
  double X0, Y0, W, H;
  int32_t transpose, flip_u, flip_v, pad;
} OrImage;

#define REAL double
#define NAME(x) x##_f64
#include "raster_impl.inc"
#undef REAL
#undef NAME

#define REAL float
#define NAME(x) x##_f32
#include "raster_impl.inc"
#undef REAL
#undef NAME

int oracle_abi(void) { return 1; }

/* threads the OpenMP loops above use from now on (bench.py times the port on 1 thread and on all cores) */
#ifdef _OPENMP
#include <omp.h>
void oracle_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int oracle_max_threads(void) { return omp_get_max_threads(); }
#else
void oracle_set_threads(int n) { (void)n; }
int oracle_max_threads(void) { return 1; }
#endif
