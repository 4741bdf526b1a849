// libstitch.so: context, job planning and kernel launch behind the C ABI (include/stitch.h).
//
// Replaces, for the stitch path only (reference paths under
// /root/reference/miniprogram-stitch/miniprogram/, P = pages/index/index.js, U = utils/canvas.js):
//   P:1186-1633 onStitch()            -> stitch_run / stitch_submit + stitch_wait
//   U:131-150   safeCreateOffscreenCanvas -> caller/ctx-owned device canvas
//   U:153-202   drawWithOrientation   -> ImgParam placement + kernels
//   U:205-242   safeCanvasToTempFilePath  -> RGBA / I420 export into the caller's buffer
// There is no CPU pixel path in this library.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "gather_kernel.cuh"
#include "stitch.h"
#include "stitch_plan.h"
#include "launch.cuh"
#include "host_stage.h"
#include "jpeg_kernels.cuh"
#include "jpeg_decode.cuh"
#include "jpeg_prog.cuh"
#include "png_decode.h"
#include "bmp_decode.cuh"
#include "gif_decode.cuh"
#include "png_kernels.cuh"
#include "synth.cuh"

// A submitted job.  Jobs that decode files and / or export a file are DEFERRED: everything is enqueued without a host
// round trip at submit, and what needs the host (was the entropy decoder's fixed number of synchronisation rounds
// enough?  how long is the file?  copy it out) happens in finalize_job -- at stitch_wait, or when the job's lane is
// needed again.
struct PendingJob {
  uint64_t ticket = 0;
  cudaEvent_t done = nullptr;
  int32_t n = 0;
  int32_t lane = 0;
  bool deferred = false, finalized = false;
  int status = STITCH_OK;
  std::string error;
  StitchResult* res = nullptr;
  bool dec = false, pngdec = false, jpeg = false, png = false, out_host = true;
  const uint8_t* j_file = nullptr;   // device file staging (host outputs)
  const uint8_t* p_file = nullptr;
  StitchJob job;                     // for the synchronous redo when the decoder needed more rounds
  std::vector<StitchImage> images;
};

// words of a lane's pinned info block
enum { INFO_JPEG_BYTES = 0, INFO_JPEG_OVERFLOW = 1, INFO_PNG_BYTES = 2, INFO_PNG_OVERFLOW = 3, INFO_DEC_CHANGED = 4, INFO_PNGDEC_STATUS = 5, INFO_PDEC_CHANGED = 6, INFO_GIFDEC_STATUS = 7, INFO_WORDS = 8 };

// One job lane: a stream with its own unit-scheduler counters and its own staging, so that jobs
// submitted to different lanes overlap -- the next job's blocks fill the SM slots the draining one
// frees, and with host buffers lane A's device->host export runs against lane B's host->device upload.
struct Lane {
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  stitch::StreamScratch scratch;
  // grow-only device staging for host-resident inputs / outputs
  std::vector<void*> in_stage;
  std::vector<size_t> in_stage_cap;
  void* out_stage = nullptr;
  size_t out_stage_cap = 0;
  // JPEG export: one grow-only device arena (I420 planes when the caller did not ask for them, coefficients,
  // segment scratch, lengths / offsets / info, file staging for host outputs) and a pinned info word pair
  void* jpeg_arena = nullptr;
  size_t jpeg_arena_cap = 0;
  // stitch_run's pipelined form: canvas pieces go home on a second stream while later bitmaps still upload
  cudaStream_t dl_stream = nullptr, ul_stream = nullptr;
  cudaEvent_t ev_draw[STITCH_MAX_IMAGES + 1] = {};
  cudaEvent_t ev_up[STITCH_MAX_IMAGES] = {};
  cudaEvent_t ev_dl = nullptr, ev_start = nullptr;
  // pageable host buffers (host_stage.h): pinned upload ring, pinned copy of the canvas, events + counter of the
  // copy-out tasks the pool still owes this lane
  stitch::UploadRing* up_ring = nullptr;
  uint8_t* host_canvas = nullptr;
  size_t host_canvas_cap = 0;
  std::vector<cudaEvent_t> dl_events;
  size_t dl_events_used = 0;
  std::atomic<int>* host_copies = nullptr;
  uint64_t* info_host = nullptr;   // pinned, INFO_WORDS words: file sizes / overflow flags / decoder convergence
  uint64_t busy_ticket = 0;        // deferred job that still owns this lane's arenas and info block (0 = none)
  // JPEG decode: grow-only arena (scan bytes, tables, subsequence states, coefficients, component planes)
  void* dec_arena = nullptr;
  size_t dec_arena_cap = 0;
  void* prog_arena = nullptr;        // progressive files of the job (jpeg_prog.cuh): their own arena, same rules
  size_t prog_arena_cap = 0;
  void* bmp_arena = nullptr;         // BMP files of the job (bmp_decode.cuh): device copies of the files + palettes
  size_t bmp_arena_cap = 0;
  void* gif_arena = nullptr;         // GIF files of the job (gif_decode.cuh): files, LZW streams, index buffers
  size_t gif_arena_cap = 0;
  bool gifdec_active = false;        // the job on this lane decodes GIF files: check INFO_GIFDEC_STATUS when it is done
  stitch::pngdec::Arena pngdec_arena;   // PNG-input decoder (png_decode.cu)
  bool pngdec_active = false;        // the job on this lane decodes PNG files: check INFO_PNGDEC_STATUS when it is done
  // PNG export: grow-only arena (RGBA canvas when the caller did not ask for it, code tables, coded bands, file)
  void* png_arena = nullptr;
  size_t png_arena_cap = 0;
};

struct StitchCtx {
  int device = 0;
  std::vector<Lane> lanes;   // lanes[0] = the ctx stream (stitch_run); stitch_submit rotates over all
  StitchProgressFn progress = nullptr;
  void* progress_user = nullptr;
  std::string last_error;
  int64_t launches = 0;
  uint64_t next_ticket = 1;
  std::deque<PendingJob> pending;
  bool busy = false;
  int sm_count = 148;
};

namespace {

int fail(StitchCtx* ctx, int code, const std::string& msg) {
  if (ctx) ctx->last_error = msg;
  return code;
}

#define CUDA_TRY(ctx, expr)                                                                  \
  do {                                                                                       \
    cudaError_t e__ = (expr);                                                                \
    if (e__ != cudaSuccess) {                                                                \
      return fail(ctx, STITCH_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__)); \
    }                                                                                        \
  } while (0)

void emit(StitchCtx* ctx, int pct) {
  if (ctx->progress) ctx->progress(ctx->progress_user, pct);
}

// Resolve one drawWithOrientation call (U:153-202) + ctx.scale(ss,ss) (P:1426-1428).
void make_placement(const StitchRect& r, int orientation, double ss, int bw, int bh, int cw, int ch, ImgParam* p) {
  double dx = r.dx, dy = r.dy;
  const double dw = r.dw, dh = r.dh;
  bool transpose = false, flip_u = false, flip_v = false;  // bitmap-axis flips
  switch (orientation) {
    case 2: flip_u = true; break;                                   // U:161-165
    case 3: flip_u = true; flip_v = true; break;                    // U:166-170
    case 4: flip_v = true; break;                                   // U:171-175
    case 5: transpose = true; break;                                // U:176-181
    case 6: transpose = true; flip_v = true; break;                 // U:182-186
    case 7: transpose = true; flip_u = true; flip_v = true; dy = dy - dh; break;  // U:187-192 (lands above the rect)
    case 8: transpose = true; flip_u = true; break;                 // U:193-197
    default: break;                                                 // U:155-159,198-199
  }
  p->X0 = dx * ss;
  p->Y0 = dy * ss;
  p->W = dw * ss;
  p->H = dh * ss;
  p->bw = bw;
  p->bh = bh;
  p->transpose = transpose ? 1 : 0;
  if (transpose) {  // device x runs along bitmap y
    p->sx_len = bh;
    p->sy_len = bw;
    p->flip_x = flip_v;
    p->flip_y = flip_u;
  } else {
    p->sx_len = bw;
    p->sy_len = bh;
    p->flip_x = flip_u;
    p->flip_y = flip_v;
  }
  auto lo = [](double v, int hi) { double f = std::floor(v); return (int)std::fmin(std::fmax(f, 0.0), (double)hi); };
  auto hi = [](double v, int hi_) { double c = std::ceil(v); return (int)std::fmin(std::fmax(c, 0.0), (double)hi_); };
  p->x_lo = lo(p->X0, cw);
  p->x_hi = hi(p->X0 + p->W, cw);
  p->y_lo = lo(p->Y0, ch);
  p->y_hi = hi(p->Y0 + p->H, ch);
  if (!(p->W > 0) || !(p->H > 0)) p->x_hi = p->x_lo;  // empty
}

int ensure_info(StitchCtx* ctx, Lane& ln) {
  if (!ln.info_host) {
    CUDA_TRY(ctx, cudaMallocHost((void**)&ln.info_host, INFO_WORDS * sizeof(uint64_t)));
    std::memset(ln.info_host, 0, INFO_WORDS * sizeof(uint64_t));
  }
  return STITCH_OK;
}

int ensure_stage(StitchCtx* ctx, Lane& ln, size_t idx, size_t bytes) {
  if (ln.in_stage.size() <= idx) {
    ln.in_stage.resize(idx + 1, nullptr);
    ln.in_stage_cap.resize(idx + 1, 0);
  }
  if (ln.in_stage_cap[idx] < bytes) {
    if (ln.in_stage[idx]) CUDA_TRY(ctx, cudaFree(ln.in_stage[idx]));
    ln.in_stage[idx] = nullptr;
    ln.in_stage_cap[idx] = 0;
    CUDA_TRY(ctx, cudaMalloc(&ln.in_stage[idx], bytes));
    // row padding of staged bitmaps is read by the kernel's last 16-byte group: keep it opaque
    CUDA_TRY(ctx, cudaMemsetAsync(ln.in_stage[idx], 0xff, bytes, ln.stream));
    ln.in_stage_cap[idx] = bytes;
  }
  return STITCH_OK;
}

void free_lane(Lane& ln) {
  if (ln.stream) cudaStreamSynchronize(ln.stream);
  for (void* p : ln.in_stage)
    if (p) cudaFree(p);
  if (ln.out_stage) cudaFree(ln.out_stage);
  if (ln.jpeg_arena) cudaFree(ln.jpeg_arena);
  if (ln.dec_arena) cudaFree(ln.dec_arena);
  if (ln.prog_arena) cudaFree(ln.prog_arena);
  if (ln.bmp_arena) cudaFree(ln.bmp_arena);
  if (ln.gif_arena) cudaFree(ln.gif_arena);
  stitch::pngdec::free_arena(ln.pngdec_arena);
  if (ln.png_arena) cudaFree(ln.png_arena);
  if (ln.info_host) cudaFreeHost(ln.info_host);
  if (ln.dl_stream) { cudaStreamSynchronize(ln.dl_stream); cudaStreamDestroy(ln.dl_stream); }
  if (ln.ul_stream) { cudaStreamSynchronize(ln.ul_stream); cudaStreamDestroy(ln.ul_stream); }
  for (cudaEvent_t e : ln.ev_draw)
    if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : ln.ev_up)
    if (e) cudaEventDestroy(e);
  if (ln.ev_dl) cudaEventDestroy(ln.ev_dl);
  if (ln.ev_start) cudaEventDestroy(ln.ev_start);
  if (ln.host_copies) {
    while (ln.host_copies->load(std::memory_order_acquire) > 0) std::this_thread::yield();
    delete ln.host_copies;
  }
  if (ln.up_ring) { ln.up_ring->release(); delete ln.up_ring; }
  if (ln.host_canvas) cudaFreeHost(ln.host_canvas);
  for (cudaEvent_t e : ln.dl_events) cudaEventDestroy(e);
  stitch::free_scratch(ln.scratch);
  if (ln.own_stream && ln.stream) cudaStreamDestroy(ln.stream);
  ln = Lane();
}

// One JPEG file to decode into a device bitmap.
struct DecodeItem {
  const uint8_t* file;
  size_t bytes;
  stitch::jpegdec::Header hdr;
  uint8_t* rgba;     // device destination
  int64_t pitch;
};

// GPU decode of up to 9 files in one batch of launches on the lane's stream (jpeg_decode.cuh).
// deferred = false: synchronises the stream after every synchronisation round (their number is data dependent).
// deferred = true: enqueues kDecRounds rounds -- a round whose predecessor changed nothing returns at once -- and
// copies the last round's flag into the lane's info block; the caller checks it when the job is finalised and
// decodes again synchronously if the chains had not converged (no such file seen yet: photos take 2 rounds).
constexpr int kDecRounds = 6;
int decode_baseline_batch(StitchCtx* ctx, Lane& ln, DecodeItem* items, int n, bool deferred) {
  namespace jd = stitch::jpegdec;
  if (n <= 0) return STITCH_OK;
  if (n > JDEC_MAX_IMAGES) return fail(ctx, STITCH_E_TOO_MANY, "more than 9 images");
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  jd::DecBatch B;
  std::memset(&B, 0, sizeof B);
  B.n = n;
  size_t need = 256;  // the `changed` flag
  struct Carve { size_t scan, tab, state, prefix, coef, plane[3], dcsum, consumed, cp; } cv[JDEC_MAX_IMAGES];
  int max_mcu = 0;
  int max_sub = 0, max_blocks = 0, max_w = 0, max_h = 0;
  {
    size_t total_scan = 0;
    for (int i = 0; i < n; ++i) total_scan += items[i].hdr.scan_len;
    B.sub = jd::jdec_pick_sub(total_scan);
    static const bool cp_off = [] { const char* e = std::getenv("STITCH_JDEC_CP"); return e && e[0] == '0'; }();
    B.nseg = cp_off ? 1u : B.sub / JDEC_SEG;
  }
  for (int i = 0; i < n; ++i) {
    const jd::Header& H = items[i].hdr;
    jd::DecImage& I = B.img[i];
    if (H.scan_len >= (size_t)1 << 31) return fail(ctx, STITCH_E_BAD_IMAGE, "JPEG scan larger than 2 GiB");
    I.scan_len = (uint32_t)H.scan_len;
    I.w = H.w; I.h = H.h; I.ncomp = H.ncomp; I.hs = H.hs; I.vs = H.vs; I.mcux = H.mcux; I.mcuy = H.mcuy; I.ri = H.ri;
    I.bpm = H.ncomp == 1 ? 1 : H.hs * H.vs + 2;
    I.n_sub = (int)((H.scan_len + B.sub - 1) / B.sub);
    if (I.n_sub < 1) I.n_sub = 1;
    I.n_blocks = H.mcux * H.mcuy * I.bpm;
    for (int c = 0; c < 3; ++c) { I.td[c] = H.td[c]; I.ta[c] = H.ta[c]; I.tq[c] = H.tq[c]; }
    I.pw[0] = H.mcux * H.hs * 8; I.ph[0] = H.mcuy * H.vs * 8;
    I.pw[1] = I.pw[2] = H.ncomp == 3 ? H.mcux * 8 : 0;
    I.ph[1] = I.ph[2] = H.ncomp == 3 ? H.mcuy * 8 : 0;
    I.rgba = items[i].rgba;
    I.pitch = items[i].pitch;
    Carve& c = cv[i];
    c.scan = need; need += up(H.scan_len + 16);
    c.tab = need; need += up(sizeof(jd::Tables));
    c.state = need; need += up((size_t)I.n_sub * 16);
    c.prefix = need; need += up((size_t)I.n_sub * 4);
    c.coef = need; need += up((size_t)I.n_blocks * 128);
    for (int k = 0; k < 3; ++k) { c.plane[k] = need; need += up((size_t)I.pw[k] * I.ph[k] + 16); }
    c.dcsum = need; need += up((size_t)H.mcux * H.mcuy * 3 * 4);
    c.consumed = need; need += up((size_t)((I.n_sub + JDEC_THREADS - 1) / JDEC_THREADS) * 8);
    c.cp = need; need += up((size_t)I.n_sub * B.nseg * 16);
    max_mcu = std::max(max_mcu, H.mcux * H.mcuy);
    max_sub = std::max(max_sub, I.n_sub);
    max_blocks = std::max(max_blocks, I.n_blocks);
    max_w = std::max(max_w, H.w);
    max_h = std::max(max_h, H.h);
  }
  if (ln.dec_arena_cap < need) {
    if (ln.dec_arena) {
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
      CUDA_TRY(ctx, cudaFree(ln.dec_arena));
    }
    ln.dec_arena = nullptr;
    ln.dec_arena_cap = 0;
    if (cudaMalloc(&ln.dec_arena, need) != cudaSuccess)
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (JPEG decode buffers allocation failed)");
    ln.dec_arena_cap = need;
  }
  {
    const int st = ensure_info(ctx, ln);
    if (st != STITCH_OK) return st;
  }
  uint8_t* A = (uint8_t*)ln.dec_arena;
  B.changed = (uint32_t*)A;
  for (int i = 0; i < n; ++i) {
    jd::DecImage& I = B.img[i];
    const Carve& c = cv[i];
    const jd::Header& H = items[i].hdr;
    I.scan = A + c.scan;
    I.tab = (const jd::Tables*)(A + c.tab);
    I.state = (uint4*)(A + c.state);
    I.prefix = (uint32_t*)(A + c.prefix);
    I.coef = (int16_t*)(A + c.coef);
    for (int k = 0; k < 3; ++k) I.plane[k] = A + c.plane[k];
    I.dcsum = (int32_t*)(A + c.dcsum);
    I.consumed = (uint2*)(A + c.consumed);
    I.cp = B.nseg > 1 ? (uint4*)(A + c.cp) : nullptr;
    CUDA_TRY(ctx, cudaMemcpyAsync(A + c.scan, items[i].file + H.scan_off, H.scan_len, cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(A + c.tab, &H.tab, sizeof(jd::Tables), cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemsetAsync(A + c.coef, 0, (size_t)I.n_blocks * 128, ln.stream));
  }
  const dim3 gsub((max_sub + JDEC_THREADS - 1) / JDEC_THREADS, n);
  jd::jdec_first_kernel<<<gsub, JDEC_THREADS, 0, ln.stream>>>(B);
  int launches = 1, rounds = 0;
  if (deferred) {
    CUDA_TRY(ctx, cudaMemsetAsync(B.changed, 0, 4 * kDecRounds, ln.stream));
    for (; rounds < kDecRounds; ++rounds) {
      B.round = rounds;
      B.flag_cur = rounds;
      B.flag_prev = rounds - 1;
      jd::jdec_sync_kernel<jd::DecBatch><<<gsub, JDEC_THREADS, 0, ln.stream>>>(B);
      ++launches;
    }
    ln.info_host[INFO_DEC_CHANGED] = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_DEC_CHANGED], B.changed + (kDecRounds - 1), 4, cudaMemcpyDeviceToHost, ln.stream));
  } else {
    for (;; ++rounds) {
      if (rounds > 4096) return fail(ctx, STITCH_E_BAD_IMAGE, "JPEG entropy decoder did not synchronise");
      CUDA_TRY(ctx, cudaMemsetAsync(B.changed, 0, 4, ln.stream));
      B.round = rounds;
      B.flag_cur = 0;
      B.flag_prev = -1;
      jd::jdec_sync_kernel<jd::DecBatch><<<gsub, JDEC_THREADS, 0, ln.stream>>>(B);
      ++launches;
      ln.info_host[INFO_DEC_CHANGED] = 0;
      CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_DEC_CHANGED], B.changed, 4, cudaMemcpyDeviceToHost, ln.stream));
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
      if (*(volatile uint32_t*)&ln.info_host[INFO_DEC_CHANGED] == 0) break;
    }
  }
  jd::jdec_scan_kernel<jd::DecBatch><<<n, 1024, 0, ln.stream>>>(B);
  jd::jdec_write_kernel<<<gsub, JDEC_THREADS, 0, ln.stream>>>(B);
  jd::jdec_dc_sum_kernel<<<dim3((max_mcu + 255) / 256, n), 256, 0, ln.stream>>>(B);
  jd::jdec_dc_scan_kernel<<<dim3(n, 3), 1024, 0, ln.stream>>>(B);
  jd::jdec_dc_apply_kernel<<<dim3((max_mcu + 255) / 256, n), 256, 0, ln.stream>>>(B);
  jd::jdec_idct_kernel<<<dim3((max_blocks + 127) / 128, n), 128, 0, ln.stream>>>(B);
  jd::jdec_colour_kernel<<<dim3((max_w + 127) / 128, (max_h + 7) / 8, n), 256, 0, ln.stream>>>(B);
  launches += 7;
  CUDA_TRY(ctx, cudaGetLastError());
  ctx->launches += launches;
  return STITCH_OK;
}

// GPU decode of the PROGRESSIVE files of a job (jpeg_prog.cuh) on the lane's stream.  First scans (Ah = 0) of all
// files go through the self-synchronising decoder as one batch; the other scans follow in dependency levels, one
// thread per scan; then the shared dequantise / IDCT / colour kernels.  `deferred` as in decode_baseline_batch (the
// flag lands in INFO_PDEC_CHANGED).  STITCH_JPROG_SEQ=1 sends every scan down the one-thread path (tests).
int decode_prog_batch(StitchCtx* ctx, Lane& ln, DecodeItem* items, int n, bool deferred) {
  namespace jd = stitch::jpegdec;
  if (n <= 0) return STITCH_OK;
  if (n > JDEC_MAX_IMAGES) return fail(ctx, STITCH_E_TOO_MANY, "more than 9 images");
  const char* seq_env = std::getenv("STITCH_JPROG_SEQ");
  const bool all_seq = seq_env && seq_env[0] == '1';
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  struct ScanRef { int img, scan, level; bool par; size_t data, state, prefix, consumed, dcdiff; jd::PScan ps;
                   int deps[4]; int ndeps; bool deps_ok; };
  struct ImgCarve { size_t tabs, qtab, coef, plane[3], file; };
  std::vector<ScanRef> refs;
  ImgCarve cv[JDEC_MAX_IMAGES];
  jd::DecBatch B;   // back end (IDCT + colour) only
  std::memset(&B, 0, sizeof B);
  B.n = n;
  size_t need = 256;   // `changed` flags
  size_t par_bytes = 0;
  int max_blocks = 0, max_w = 0, max_h = 0, max_level = 0;
  for (int i = 0; i < n; ++i) {
    const jd::Header& H = items[i].hdr;
    jd::DecImage& I = B.img[i];
    I.w = H.w; I.h = H.h; I.ncomp = H.ncomp; I.hs = H.hs; I.vs = H.vs; I.mcux = H.mcux; I.mcuy = H.mcuy; I.ri = 0;
    I.bpm = H.ncomp == 1 ? 1 : H.hs * H.vs + 2;
    I.n_blocks = H.mcux * H.mcuy * I.bpm;
    for (int c = 0; c < 3; ++c) I.tq[c] = H.tq[c];
    I.pw[0] = H.mcux * H.hs * 8; I.ph[0] = H.mcuy * H.vs * 8;
    I.pw[1] = I.pw[2] = H.ncomp == 3 ? H.mcux * 8 : 0;
    I.ph[1] = I.ph[2] = H.ncomp == 3 ? H.mcuy * 8 : 0;
    I.rgba = items[i].rgba;
    I.pitch = items[i].pitch;
    ImgCarve& c = cv[i];
    c.file = need; need += up(items[i].bytes + 16);
    c.tabs = need; need += up(H.scan_tabs.size() * sizeof(jd::Tables));
    c.qtab = need; need += up(sizeof(jd::Tables));
    c.coef = need; need += up((size_t)I.n_blocks * 128);
    for (int k = 0; k < 3; ++k) { c.plane[k] = need; need += up((size_t)I.pw[k] * I.ph[k] + 16); }
    max_blocks = std::max(max_blocks, I.n_blocks);
    max_w = std::max(max_w, H.w);
    max_h = std::max(max_h, H.h);
    // dependency levels: a scan waits for every earlier scan that shares a component and overlaps its band
    const size_t first_ref = refs.size();
    for (size_t k = 0; k < H.scans.size(); ++k) {
      const jd::ProgScan& ps = H.scans[k];
      ScanRef r;
      std::memset(&r.ps, 0, sizeof r.ps);
      r.img = i;
      r.scan = (int)k;
      r.level = 0;
      r.ndeps = 0;
      r.deps_ok = true;
      for (size_t j = 0; j < k; ++j) {
        const jd::ProgScan& q = H.scans[j];
        const bool shares = ps.ns > 1 || q.ns > 1 || ps.comp == q.comp;
        if (shares && ps.ss <= q.se && q.ss <= ps.se) {
          r.level = std::max(r.level, refs[first_ref + j].level + 1);
          // a predecessor that is itself a one-warp scan (those of the parallel batch are finished before that launch):
          // chained through a progress word when both walk the same blocks in the same order
          if (!refs[first_ref + j].par) {
            const bool same_order = ps.ns == q.ns && (ps.ns > 1 || ps.comp == q.comp);
            if (!same_order || r.ndeps == 4) r.deps_ok = false;
            else r.deps[r.ndeps++] = (int)(first_ref + j);
          }
        }
      }
      r.par = !all_seq && ps.ah == 0 && r.level == 0;
      jd::pscan_describe(H, ps, &r.ps);
      r.data = need; need += up(ps.len + 16);
      r.state = r.prefix = r.consumed = r.dcdiff = 0;
      if (r.par) par_bytes += ps.len;
      refs.push_back(r);
    }
  }
  const uint32_t sub = jd::jdec_pick_sub(par_bytes);
  int n_par = 0, max_sub = 0;
  for (ScanRef& r : refs) {
    if (!r.par) { max_level = std::max(max_level, r.level); continue; }
    ++n_par;
    r.ps.n_sub = std::max(1, (int)((r.ps.len + sub - 1) / sub));
    max_sub = std::max(max_sub, r.ps.n_sub);
    r.state = need; need += up((size_t)r.ps.n_sub * 16);
    r.prefix = need; need += up((size_t)r.ps.n_sub * 4);
    r.consumed = need; need += up((size_t)((r.ps.n_sub + JDEC_THREADS - 1) / JDEC_THREADS) * 8);
    if (r.ps.mode == jd::PSCAN_DC_FIRST) { r.dcdiff = need; need += up((size_t)r.ps.n_blk * 4); }
  }
  // device copies of the scan records: the first scans in front, the others sorted by level behind them
  std::vector<int> order;
  for (size_t k = 0; k < refs.size(); ++k) if (refs[k].par) order.push_back((int)k);
  std::vector<int> level_off(max_level + 2, 0);
  for (int lv = 0; lv <= max_level; ++lv) {
    level_off[lv] = (int)order.size();
    for (size_t k = 0; k < refs.size(); ++k) if (!refs[k].par && refs[k].level == lv) order.push_back((int)k);
  }
  level_off[max_level + 1] = (int)order.size();
  const size_t recs = need;
  need += up(order.size() * sizeof(jd::PScan));
  const size_t prog_words = need;
  need += up(refs.size() * 4);
  const size_t piece_recs = need;
  need += up(refs.size() * sizeof(jd::ScanPiece));
  // One launch for all one-warp scans when every dependency among them is between scans of the same block order (they
  // are then chained through progress words); STITCH_JPROG_LEVELS=1 keeps a launch per dependency level.
  const char* lv_env = std::getenv("STITCH_JPROG_LEVELS");
  bool one_launch = !(lv_env && lv_env[0] == '1') && (int)order.size() - n_par <= 128;
  for (const ScanRef& r : refs)
    if (!r.par && !r.deps_ok) one_launch = false;
  if (ln.prog_arena_cap < need) {
    if (ln.prog_arena) {
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
      CUDA_TRY(ctx, cudaFree(ln.prog_arena));
    }
    ln.prog_arena = nullptr;
    ln.prog_arena_cap = 0;
    if (cudaMalloc(&ln.prog_arena, need) != cudaSuccess)
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (JPEG decode buffers allocation failed)");
    ln.prog_arena_cap = need;
  }
  {
    const int st = ensure_info(ctx, ln);
    if (st != STITCH_OK) return st;
  }
  uint8_t* A = (uint8_t*)ln.prog_arena;
  for (int i = 0; i < n; ++i) {
    const jd::Header& H = items[i].hdr;
    jd::DecImage& I = B.img[i];
    const ImgCarve& c = cv[i];
    I.tab = (const jd::Tables*)(A + c.qtab);
    I.coef = (int16_t*)(A + c.coef);
    for (int k = 0; k < 3; ++k) I.plane[k] = A + c.plane[k];
    CUDA_TRY(ctx, cudaMemcpyAsync(A + c.file, items[i].file, items[i].bytes, cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(A + c.tabs, H.scan_tabs.data(), H.scan_tabs.size() * sizeof(jd::Tables), cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(A + c.qtab, &H.tab, sizeof(jd::Tables), cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemsetAsync(A + c.coef, 0, (size_t)I.n_blocks * 128, ln.stream));
  }
  std::vector<jd::PScan> host_recs(order.size());
  std::vector<jd::ScanPiece> host_pieces;
  host_pieces.reserve(order.size());
  uint32_t max_piece = 0;
  for (size_t k = 0; k < order.size(); ++k) {
    ScanRef& r = refs[order[k]];
    const jd::Header& H = items[r.img].hdr;
    const jd::ProgScan& ps = H.scans[r.scan];
    r.ps.data = A + r.data;
    r.ps.tab = (const jd::Tables*)(A + cv[r.img].tabs) + ps.tab;
    r.ps.coef = (int16_t*)(A + cv[r.img].coef);
    if (r.par) {
      r.ps.state = (uint4*)(A + r.state);
      r.ps.prefix = (uint32_t*)(A + r.prefix);
      r.ps.consumed = (uint2*)(A + r.consumed);
      r.ps.dcdiff = r.dcdiff ? (int32_t*)(A + r.dcdiff) : nullptr;
    }
    if (ps.len) {
      host_pieces.push_back(jd::ScanPiece{A + cv[r.img].file + ps.off, A + r.data, (uint32_t)ps.len});
      max_piece = std::max(max_piece, (uint32_t)ps.len);
    }
    r.ps.progress = nullptr;
    r.ps.ndep = 0;
    if (!r.par && one_launch) {
      r.ps.progress = (uint32_t*)(A + prog_words) + order[k];
      for (int d = 0; d < r.ndeps; ++d) r.ps.dep[r.ps.ndep++] = (const uint32_t*)(A + prog_words) + r.deps[d];
    }
    host_recs[k] = r.ps;
  }
  if (one_launch) CUDA_TRY(ctx, cudaMemsetAsync(A + prog_words, 0, refs.size() * 4, ln.stream));
  int launches = 0;
  if (!host_pieces.empty()) {
    CUDA_TRY(ctx, cudaMemcpyAsync(A + piece_recs, host_pieces.data(), host_pieces.size() * sizeof(jd::ScanPiece), cudaMemcpyHostToDevice, ln.stream));
    const unsigned gy = std::min<unsigned>(64u, (max_piece + 4095u) / 4096u);
    jd::jprog_gather_kernel<<<dim3((unsigned)host_pieces.size(), std::max(1u, gy)), 256, 0, ln.stream>>>((const jd::ScanPiece*)(A + piece_recs));
    ++launches;
  }
  jd::PScan* d_recs = (jd::PScan*)(A + recs);
  if (!host_recs.empty())
    CUDA_TRY(ctx, cudaMemcpyAsync(d_recs, host_recs.data(), host_recs.size() * sizeof(jd::PScan), cudaMemcpyHostToDevice, ln.stream));
  ln.info_host[INFO_PDEC_CHANGED] = 0;
  if (n_par > 0) {
    jd::PBatch P;
    std::memset(&P, 0, sizeof P);
    P.scans = d_recs;
    P.changed = (uint32_t*)A;
    P.sub = sub;
    const dim3 gsub((max_sub + JDEC_THREADS - 1) / JDEC_THREADS, n_par);
    jd::jprog_first_kernel<<<gsub, JDEC_THREADS, 0, ln.stream>>>(P);
    ++launches;
    int rounds = 0;
    if (deferred) {
      CUDA_TRY(ctx, cudaMemsetAsync(P.changed, 0, 4 * kDecRounds, ln.stream));
      for (; rounds < kDecRounds; ++rounds) {
        P.round = rounds;
        P.flag_cur = rounds;
        P.flag_prev = rounds - 1;
        jd::jdec_sync_kernel<jd::PBatch><<<gsub, JDEC_THREADS, 0, ln.stream>>>(P);
        ++launches;
      }
      CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_PDEC_CHANGED], P.changed + (kDecRounds - 1), 4, cudaMemcpyDeviceToHost, ln.stream));
    } else {
      for (;; ++rounds) {
        if (rounds > 4096) return fail(ctx, STITCH_E_BAD_IMAGE, "JPEG entropy decoder did not synchronise");
        CUDA_TRY(ctx, cudaMemsetAsync(P.changed, 0, 4, ln.stream));
        P.round = rounds;
        P.flag_cur = 0;
        P.flag_prev = -1;
        jd::jdec_sync_kernel<jd::PBatch><<<gsub, JDEC_THREADS, 0, ln.stream>>>(P);
        ++launches;
        ln.info_host[INFO_PDEC_CHANGED] = 0;
        CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_PDEC_CHANGED], P.changed, 4, cudaMemcpyDeviceToHost, ln.stream));
        CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
        if (*(volatile uint32_t*)&ln.info_host[INFO_PDEC_CHANGED] == 0) break;
      }
    }
    jd::jdec_scan_kernel<jd::PBatch><<<n_par, 1024, 0, ln.stream>>>(P);
    jd::jprog_write_kernel<<<gsub, JDEC_THREADS, 0, ln.stream>>>(P);
    jd::jprog_dc_kernel<<<dim3(n_par, 3), 1024, 0, ln.stream>>>(d_recs);
    launches += 3;
  }
  if (one_launch) {
    const int cnt = (int)order.size() - n_par;       // sorted by level: producers get the lower CTA indices
    if (cnt > 0) {
      jd::jprog_seq_kernel<<<cnt, 32, 0, ln.stream>>>(d_recs + n_par);
      ++launches;
    }
  } else {
    for (int lv = 0; lv <= max_level; ++lv) {
      const int cnt = level_off[lv + 1] - level_off[lv];
      if (cnt <= 0) continue;
      jd::jprog_seq_kernel<<<cnt, 32, 0, ln.stream>>>(d_recs + level_off[lv]);
      ++launches;
    }
  }
  jd::jdec_idct_kernel<<<dim3((max_blocks + 127) / 128, n), 128, 0, ln.stream>>>(B);
  jd::jdec_colour_kernel<<<dim3((max_w + 127) / 128, (max_h + 7) / 8, n), 256, 0, ln.stream>>>(B);
  launches += 2;
  CUDA_TRY(ctx, cudaGetLastError());
  ctx->launches += launches;
  return STITCH_OK;
}

// The JPEG files of a job: baseline ones and progressive ones are two batches on the same stream.
int decode_batch(StitchCtx* ctx, Lane& ln, DecodeItem* items, int n, bool deferred) {
  if (n <= 0) return STITCH_OK;
  if (n > JDEC_MAX_IMAGES) return fail(ctx, STITCH_E_TOO_MANY, "more than 9 images");
  {
    const int st = ensure_info(ctx, ln);
    if (st != STITCH_OK) return st;
  }
  ln.info_host[INFO_DEC_CHANGED] = 0;
  ln.info_host[INFO_PDEC_CHANGED] = 0;
  std::vector<DecodeItem> base, prog;
  for (int i = 0; i < n; ++i) (items[i].hdr.progressive ? prog : base).push_back(items[i]);
  int st = decode_baseline_batch(ctx, ln, base.data(), (int)base.size(), deferred);
  if (st == STITCH_OK) st = decode_prog_batch(ctx, ln, prog.data(), (int)prog.size(), deferred);
  return st;
}

int parse_jpeg(StitchCtx* ctx, const void* file, size_t bytes, int idx, stitch::jpegdec::Header* H) {
  const stitch::jpegdec::ParseStatus ps = stitch::jpegdec::parse((const uint8_t*)file, bytes, H);
  if (ps == stitch::jpegdec::kOk) return STITCH_OK;
  char buf[96];
  std::snprintf(buf, sizeof buf, "图片%d解码异常 (%s)", idx,
                ps == stitch::jpegdec::kUnsupported ? "JPEG variant not supported: Huffman-coded baseline or progressive, 4:4:4 / 4:2:2 / 4:2:0 / gray only"
                : ps == stitch::jpegdec::kNotJpeg ? "not a JPEG file" : "corrupt JPEG headers");
  return fail(ctx, STITCH_E_BAD_IMAGE, buf);
}

// One BMP file to convert into a device bitmap.
struct BmpItem {
  const uint8_t* file;
  size_t bytes;
  stitch::bmpdec::Header hdr;
  uint8_t* rgba;
  int64_t pitch;
};

int parse_bmp(StitchCtx* ctx, const void* file, size_t bytes, int idx, stitch::bmpdec::Header* H) {
  const stitch::bmpdec::ParseStatus ps = stitch::bmpdec::parse((const uint8_t*)file, bytes, H);
  if (ps == stitch::bmpdec::kOk) return STITCH_OK;
  char buf[112];
  std::snprintf(buf, sizeof buf, "图片%d解码异常 (%s)", idx,
                ps == stitch::bmpdec::kUnsupported ? "BMP variant not supported: RLE, embedded JPEG / PNG, odd bit fields"
                : ps == stitch::bmpdec::kNotBmp ? "not a BMP file" : "corrupt BMP headers");
  return fail(ctx, STITCH_E_BAD_IMAGE, buf);
}

// The BMP files of a job: uploads + one kernel on the lane's stream, nothing waits.
int decode_bmp_batch(StitchCtx* ctx, Lane& ln, BmpItem* items, int n) {
  namespace bd = stitch::bmpdec;
  if (n <= 0) return STITCH_OK;
  if (n > bd::kMaxImages) return fail(ctx, STITCH_E_TOO_MANY, "more than 9 images");
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  size_t need = 0, off_file[bd::kMaxImages], off_pal[bd::kMaxImages];
  for (int i = 0; i < n; ++i) {
    off_file[i] = need; need += up(items[i].bytes + 16);
    off_pal[i] = need; need += 1024;
  }
  if (ln.bmp_arena_cap < need) {
    if (ln.bmp_arena) {
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
      CUDA_TRY(ctx, cudaFree(ln.bmp_arena));
    }
    ln.bmp_arena = nullptr;
    ln.bmp_arena_cap = 0;
    if (cudaMalloc(&ln.bmp_arena, need) != cudaSuccess)
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (BMP decode buffers allocation failed)");
    ln.bmp_arena_cap = need;
  }
  uint8_t* A = (uint8_t*)ln.bmp_arena;
  bd::Batch B;
  std::memset(&B, 0, sizeof B);
  B.n = n;
  int max_w = 0, max_h = 0;
  for (int i = 0; i < n; ++i) {
    const bd::Header& H = items[i].hdr;
    bd::Img& I = B.img[i];
    I.file = A + off_file[i];
    I.palette = (const uint32_t*)(A + off_pal[i]);
    I.rgba = items[i].rgba;
    I.pitch = items[i].pitch;
    I.w = H.w; I.h = H.h; I.bpp = H.bpp; I.top_down = H.top_down ? 1 : 0;
    I.data_off = H.data_off;
    I.row_stride = H.row_stride;
    for (int k = 0; k < 4; ++k) { I.shift[k] = H.shift[k]; I.bits[k] = H.bits[k]; }
    CUDA_TRY(ctx, cudaMemcpyAsync(A + off_file[i], items[i].file, items[i].bytes, cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(A + off_pal[i], H.palette, 1024, cudaMemcpyHostToDevice, ln.stream));
    max_w = std::max(max_w, H.w);
    max_h = std::max(max_h, H.h);
  }
  bd::bmp_rgba_kernel<<<dim3((max_w + 63) / 64, (max_h + 3) / 4, n), 256, 0, ln.stream>>>(B);
  CUDA_TRY(ctx, cudaGetLastError());
  ctx->launches += 1;
  return STITCH_OK;
}

// One GIF file (first frame) to decode into a device bitmap.
struct GifItem {
  const uint8_t* file;
  size_t bytes;
  stitch::gifdec::Header hdr;
  uint8_t* rgba;
  int64_t pitch;
};

int parse_gif(StitchCtx* ctx, const void* file, size_t bytes, int idx, stitch::gifdec::Header* H) {
  const stitch::gifdec::ParseStatus ps = stitch::gifdec::parse((const uint8_t*)file, bytes, H);
  if (ps == stitch::gifdec::kOk) return STITCH_OK;
  char buf[112];
  std::snprintf(buf, sizeof buf, "图片%d解码异常 (%s)", idx,
                ps == stitch::gifdec::kUnsupported ? "GIF variant not supported: first frame smaller than the screen"
                : ps == stitch::gifdec::kNotGif ? "not a GIF file" : "corrupt GIF blocks");
  return fail(ctx, STITCH_E_BAD_IMAGE, buf);
}

// The GIF files of a job: uploads + three kernels on the lane's stream; the status word follows them into the lane's
// info block (read when the job is waited for).
int decode_gif_batch(StitchCtx* ctx, Lane& ln, GifItem* items, int n) {
  namespace gd = stitch::gifdec;
  if (n <= 0) return STITCH_OK;
  if (n > gd::kMaxImages) return fail(ctx, STITCH_E_TOO_MANY, "more than 9 images");
  {
    const int st = ensure_info(ctx, ln);
    if (st != STITCH_OK) return st;
  }
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  size_t need = 256, off_file[gd::kMaxImages], off_stream[gd::kMaxImages], off_idx[gd::kMaxImages], off_pal[gd::kMaxImages];
  size_t n_pieces = 0;
  for (int i = 0; i < n; ++i) {
    const gd::Header& H = items[i].hdr;
    off_file[i] = need; need += up(items[i].bytes + 16);
    off_stream[i] = need; need += up((size_t)H.stream_len + 16);
    off_idx[i] = need; need += up((size_t)gd::kPrelude + (size_t)H.w * H.h + 16);
    off_pal[i] = need; need += 1024;
    n_pieces += H.pieces.size();
  }
  const size_t off_pieces = need; need += up(n_pieces * sizeof(gd::Piece));
  const size_t off_pimg = need; need += up(n_pieces);
  if (ln.gif_arena_cap < need) {
    if (ln.gif_arena) {
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
      CUDA_TRY(ctx, cudaFree(ln.gif_arena));
    }
    ln.gif_arena = nullptr;
    ln.gif_arena_cap = 0;
    if (cudaMalloc(&ln.gif_arena, need) != cudaSuccess)
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (GIF decode buffers allocation failed)");
    ln.gif_arena_cap = need;
  }
  uint8_t* A = (uint8_t*)ln.gif_arena;
  gd::Batch B;
  std::memset(&B, 0, sizeof B);
  B.n = n;
  B.status = (uint32_t*)A;
  B.pieces = (const gd::Piece*)(A + off_pieces);
  B.piece_img = A + off_pimg;
  CUDA_TRY(ctx, cudaMemsetAsync(A, 0, 256, ln.stream));
  std::vector<gd::Piece> pieces;
  std::vector<uint8_t> pimg;
  pieces.reserve(n_pieces);
  pimg.reserve(n_pieces);
  int max_w = 0, max_h = 0;
  for (int i = 0; i < n; ++i) {
    const gd::Header& H = items[i].hdr;
    gd::Img& I = B.img[i];
    I.file = A + off_file[i];
    I.stream = A + off_stream[i];
    I.idx = A + off_idx[i];
    I.palette = (const uint32_t*)(A + off_pal[i]);
    I.rgba = items[i].rgba;
    I.pitch = items[i].pitch;
    I.stream_len = H.stream_len;
    I.w = H.w; I.h = H.h; I.interlaced = H.interlaced ? 1 : 0; I.min_code = H.min_code;
    CUDA_TRY(ctx, cudaMemcpyAsync(A + off_file[i], items[i].file, items[i].bytes, cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(A + off_pal[i], H.palette, 1024, cudaMemcpyHostToDevice, ln.stream));
    CUDA_TRY(ctx, cudaMemsetAsync(A + off_stream[i] + (H.stream_len & ~(size_t)3), 0, 16, ln.stream));   // the bit reader fetches whole words
    for (const gd::Piece& pc : H.pieces) { pieces.push_back(pc); pimg.push_back((uint8_t)i); }
    max_w = std::max(max_w, H.w);
    max_h = std::max(max_h, H.h);
  }
  CUDA_TRY(ctx, cudaMemcpyAsync(A + off_pieces, pieces.data(), pieces.size() * sizeof(gd::Piece), cudaMemcpyHostToDevice, ln.stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(A + off_pimg, pimg.data(), pimg.size(), cudaMemcpyHostToDevice, ln.stream));
  gd::gif_gather_kernel<<<(unsigned)pieces.size(), 64, 0, ln.stream>>>(B);
  gd::gif_lzw_kernel<<<n, 32, 0, ln.stream>>>(B);
  gd::gif_rgba_kernel<<<dim3((max_w + 63) / 64, (max_h + 3) / 4, n), 256, 0, ln.stream>>>(B);
  CUDA_TRY(ctx, cudaGetLastError());
  ln.info_host[INFO_GIFDEC_STATUS] = 0;
  CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_GIFDEC_STATUS], B.status, 4, cudaMemcpyDeviceToHost, ln.stream));
  ln.gifdec_active = true;
  ctx->launches += 3;
  return STITCH_OK;
}

// What kind of file an input is: the caller's word, or -- STITCH_MEM_HOST_FILE -- the file's signature.
int file_kind_of(const StitchImage& im) {
  if (im.mem_kind != STITCH_MEM_HOST_FILE) return im.mem_kind;
  const uint8_t* f = (const uint8_t*)im.data;
  const int64_t n = im.pitch_bytes;
  if (!f || n < 8) return -1;
  if (f[0] == 0xFF && f[1] == 0xD8) return STITCH_MEM_HOST_JPEG;
  if (std::memcmp(f, "\x89PNG\r\n\x1a\n", 8) == 0) return STITCH_MEM_HOST_PNG;
  if (f[0] == 'B' && f[1] == 'M') return STITCH_MEM_HOST_BMP;
  if (std::memcmp(f, "GIF8", 4) == 0) return STITCH_MEM_HOST_GIF;
  return -1;   // WebP and the rest: probed, not decoded
}

int parse_png(StitchCtx* ctx, const void* file, size_t bytes, int idx, stitch::pngdec::Header* H) {
  const stitch::pngdec::ParseStatus ps = stitch::pngdec::parse((const uint8_t*)file, bytes, H);
  if (ps == stitch::pngdec::kOk) return STITCH_OK;
  char buf[128];
  std::snprintf(buf, sizeof buf, "图片%d解码异常 (%s)", idx,
                ps == stitch::pngdec::kUnsupported ? "PNG variant not supported: larger than 2 GiB inflated"
                : ps == stitch::pngdec::kNotPng ? "not a PNG file" : "corrupt PNG chunks");
  return fail(ctx, STITCH_E_BAD_IMAGE, buf);
}

// GPU decode of the PNG files of a job (png_decode.cu): all launches on the lane's stream, nothing waits; the
// decoder's status word arrives in the lane's info block behind them.
int decode_png_batch(StitchCtx* ctx, Lane& ln, stitch::pngdec::DecodeItem* items, int n) {
  if (n <= 0) return STITCH_OK;
  const int st = ensure_info(ctx, ln);
  if (st != STITCH_OK) return st;
  ln.info_host[INFO_PNGDEC_STATUS] = 0;
  int launches = 0;
  const cudaError_t e = stitch::pngdec::decode_batch(ln.pngdec_arena, items, n, ln.stream, (uint32_t*)&ln.info_host[INFO_PNGDEC_STATUS],
                                                     ctx->sm_count, &launches);
  if (e == cudaErrorMemoryAllocation) {
    cudaGetLastError();
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (PNG decode buffers allocation failed)");
  }
  if (e != cudaSuccess) return fail(ctx, STITCH_E_CUDA, std::string("PNG decode: ") + cudaGetErrorString(e));
  ctx->launches += launches;
  ln.pngdec_active = true;
  return STITCH_OK;
}

// After the lane's stream has drained: did the PNG decoder accept its streams?
int check_png_status(StitchCtx* ctx, Lane& ln) {
  if (ln.gifdec_active) {
    ln.gifdec_active = false;
    if ((uint32_t)ln.info_host[INFO_GIFDEC_STATUS]) {
      ln.pngdec_active = false;
      return fail(ctx, STITCH_E_BAD_IMAGE, "图片解码异常 (GIF data damaged: a code outside the dictionary)");
    }
  }
  if (!ln.pngdec_active) return STITCH_OK;
  ln.pngdec_active = false;
  const uint32_t bits = (uint32_t)ln.info_host[INFO_PNGDEC_STATUS];
  if (!bits) return STITCH_OK;
  char buf[96];
  std::snprintf(buf, sizeof buf, "图片解码异常 (PNG data damaged: status 0x%x)", bits);
  return fail(ctx, STITCH_E_BAD_IMAGE, buf);
}


// STITCH_OUT_JPEG (jpeg_kernels.cuh): carve() sizes and places the work buffers in the lane's arena before the
// compose launch, run() encodes the I420 planes of that launch into the file and fetches its size.
struct JpegExport {
  int jpeg_ri = 8, jpeg_q = 100, mcu_x = 0, mcu_y = 0, n_mcu = 0, n_seg = 0;
  int64_t seg_cap = 0, j_cap = 0;
  bool out_host = true;
  uint8_t* j_file = nullptr;
  int16_t* j_coef = nullptr;
  uint8_t* j_scratch = nullptr;
  uint32_t *j_len = nullptr, *j_bits = nullptr, *j_blen = nullptr;
  uint64_t *j_off = nullptr, *j_info = nullptr;

  int carve(StitchCtx* ctx, Lane& ln, const StitchJob* job, StitchResult* res, int cw, int ch, bool want_yuv, bool host,
            uint8_t** d_yuv) {
    out_host = host;
    // JPEG export: carve the lane's arena.  The encoder reads the I420 planes of this launch: the caller's
    // when STITCH_OUT_YUV420 is set too, planes inside the arena otherwise.
    jpeg_ri = job->jpeg_restart_interval > 0 ? job->jpeg_restart_interval : 8;
    jpeg_q = job->jpeg_quality > 0 ? (job->jpeg_quality > 100 ? 100 : job->jpeg_quality) : 100;
    mcu_x = (cw + 15) / 16; mcu_y = (ch + 15) / 16; n_mcu = mcu_x * mcu_y;
    n_seg = (n_mcu + jpeg_ri - 1) / jpeg_ri;
    seg_cap = (int64_t)jpeg_ri * 6 * JPEG_BLOCK_MAX_BYTES + 16;
    if (jpeg_ri > 65535) return fail(ctx, STITCH_E_BAD_ARG, "JPEG restart interval above 65535");
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const int64_t worst = 1024 + (int64_t)n_mcu * 6 * JPEG_BLOCK_MAX_BYTES + (int64_t)n_seg * 4 + 2;
    j_cap = job->out_capacity_jpeg < worst ? job->out_capacity_jpeg : worst;
    const size_t b_yuv = want_yuv ? 0 : up((size_t)res->yuv_bytes);
    const size_t b_coef = up((size_t)n_mcu * 6 * 64 * sizeof(int16_t));
    const size_t b_scr = up((size_t)n_seg * (size_t)seg_cap);
    const size_t b_len = up((size_t)n_seg * 4), b_off = up((size_t)n_seg * 8), b_info = 256;
    const size_t b_file = out_host ? up((size_t)j_cap) : 0;
    const size_t n_blk = (size_t)n_mcu * 6;
    const size_t b_bits = up((n_blk + 31) / 32 * JPEG_BLK_WORDS * 32 * sizeof(uint32_t)), b_blen = up(n_blk * 4);
    const size_t need = b_yuv + b_coef + b_scr + b_len + b_off + b_info + b_file + b_bits + b_blen;
    if (ln.jpeg_arena_cap < need) {
      if (ln.jpeg_arena) {
        CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
        CUDA_TRY(ctx, cudaFree(ln.jpeg_arena));
      }
      ln.jpeg_arena = nullptr;
      ln.jpeg_arena_cap = 0;
      if (cudaMalloc(&ln.jpeg_arena, need) != cudaSuccess)
        return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (JPEG work buffers allocation failed)");
      ln.jpeg_arena_cap = need;
    }
    {
      const int st = ensure_info(ctx, ln);
      if (st != STITCH_OK) return st;
    }
    uint8_t* a = (uint8_t*)ln.jpeg_arena;
    if (!want_yuv) *d_yuv = a;
    a += b_yuv;
    j_coef = (int16_t*)a; a += b_coef;
    j_scratch = a; a += b_scr;
    j_len = (uint32_t*)a; a += b_len;
    j_off = (uint64_t*)a; a += b_off;
    j_info = (uint64_t*)a; a += b_info;
    j_bits = (uint32_t*)a; a += b_bits;
    j_blen = (uint32_t*)a; a += b_blen;
    j_file = out_host ? a : (uint8_t*)job->out_jpeg;
    return STITCH_OK;
  }

  int run(StitchCtx* ctx, Lane& ln, const StitchJob* job, StitchResult* res, const JobParam& P, int cw, int ch, bool deferred) {
    stitch::jpeg::Tables tab;
    stitch::jpeg::build_tables(jpeg_q, &tab);
    stitch::jpeg::HeaderBytes hdr;
    const std::vector<uint8_t> hv = stitch::jpeg::header(cw, ch, jpeg_q, jpeg_ri);
    if (hv.size() > sizeof hdr.b) return fail(ctx, STITCH_E_BAD_ARG, "JPEG header too long");
    std::memcpy(hdr.b, hv.data(), hv.size());
    hdr.n = (int32_t)hv.size();
    const int n_blocks = n_mcu * 6;
    stitch::jpeg::jpeg_dct_kernel<<<(n_blocks + 127) / 128, 128, 0, ln.stream>>>(P.out_y, P.out_cb, P.out_cr, cw, ch, mcu_x,
                                                                              mcu_y, tab, j_coef);
    stitch::jpeg::jpeg_blockbits_kernel<<<(n_blocks + 127) / 128, 128, 0, ln.stream>>>(j_coef, n_blocks, jpeg_ri, tab, j_bits,
                                                                                    j_blen);
    stitch::jpeg::jpeg_pack_kernel<<<(n_seg + 3) / 4, 128, 0, ln.stream>>>(j_bits, j_blen, n_blocks, jpeg_ri, n_seg, j_scratch,
                                                                          seg_cap, j_len);
    stitch::jpeg::jpeg_offsets_kernel<<<1, 1024, 0, ln.stream>>>(j_len, n_seg, hdr, j_file, j_cap, j_off, j_info);
    stitch::jpeg::jpeg_gather_kernel<<<(n_seg + 7) / 8, 256, 0, ln.stream>>>(j_scratch, seg_cap, j_len, j_off, n_seg, j_info,
                                                                            j_file);
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches += 5;
    res->kernel_launches += 5;
    CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_JPEG_BYTES], j_info, 16, cudaMemcpyDeviceToHost, ln.stream));
    if (deferred) return STITCH_OK;   // finish() runs when the job is finalised
    CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
    return finish(ctx, ln, res, out_host, j_file, job->out_jpeg);
  }

  // The file size has reached the host: report it, and copy the file out of the lane's staging for host outputs.
  static int finish(StitchCtx* ctx, Lane& ln, StitchResult* res, bool out_host, const uint8_t* j_file, void* out_jpeg) {
    res->jpeg_bytes = (int64_t)ln.info_host[INFO_JPEG_BYTES];
    if (ln.info_host[INFO_JPEG_OVERFLOW])
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (JPEG output buffer too small; see jpeg_bytes)");
    if (out_host) {
      CUDA_TRY(ctx, cudaMemcpyAsync(out_jpeg, j_file, (size_t)res->jpeg_bytes, cudaMemcpyDeviceToHost, ln.stream));
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
    }
    return STITCH_OK;
  }
};

// STITCH_OUT_PNG (png_kernels.cuh): same two steps for the PNG file; the encoder reads the RGBA canvas.
struct PngExport {
  stitch::png::Params PP;
  bool out_host = true;

  int carve(StitchCtx* ctx, Lane& ln, const StitchJob* job, const StitchLayout& L, int cw, int ch, bool want_rgba, bool host,
            uint8_t** d_rgba, int64_t* d_rgba_pitch) {
    out_host = host;
    // PNG export: carve the lane's arena; the encoder reads the RGBA canvas of this launch (the caller's, or
    // one inside the arena when RGBA was not requested).
    std::memset(&PP, 0, sizeof PP);
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const int nb = (ch + PNG_BAND_ROWS - 1) / PNG_BAND_ROWS;
    const int64_t band_raw = ((int64_t)cw * 4 + 1) * PNG_BAND_ROWS;
    const int64_t band_cap = (int64_t)up((size_t)(band_raw * 2 + 1024));
    const int64_t worst = stitch_png_capacity(&L);
    const int64_t p_cap = job->out_capacity_png < worst ? job->out_capacity_png : worst;
    const int64_t c_pitch = ((int64_t)cw * 4 + 15) & ~15LL;
    const size_t b_canvas = want_rgba ? 0 : up((size_t)(c_pitch * ch));
    const size_t b_lens = up((size_t)nb * PNG_TAB_STRIDE), b_codes = up((size_t)nb * PNG_TAB_STRIDE * 2);
    const size_t b_adler = up((size_t)nb * 8), b_nm = up((size_t)nb * 4), b_scr = (size_t)nb * (size_t)band_cap;
    const int rounds = (int)(((int64_t)PNG_BAND_ROWS * cw + PNG_ROUND_PX - 1) / PNG_ROUND_PX);
    const size_t b_fpx = up((size_t)nb * PNG_BAND_ROWS * (size_t)cw * 4), b_desc = up((size_t)nb * rounds * PNG_THREADS * 8);
    const size_t b_blen = up((size_t)nb * 4), b_off = up((size_t)nb * 8), b_info = 256;
    const size_t b_file = out_host ? up((size_t)p_cap) : 0;
    const size_t need = b_canvas + b_lens + b_codes + b_adler + b_nm + b_fpx + b_desc + b_scr + b_blen + b_off + b_info + b_file;
    if (ln.png_arena_cap < need) {
      if (ln.png_arena) {
        CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
        CUDA_TRY(ctx, cudaFree(ln.png_arena));
      }
      ln.png_arena = nullptr;
      ln.png_arena_cap = 0;
      if (cudaMalloc(&ln.png_arena, need) != cudaSuccess)
        return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (PNG work buffers allocation failed)");
      ln.png_arena_cap = need;
    }
    {
      const int st = ensure_info(ctx, ln);
      if (st != STITCH_OK) return st;
    }
    uint8_t* a = (uint8_t*)ln.png_arena;
    if (!want_rgba) { *d_rgba = a; *d_rgba_pitch = c_pitch; }
    a += b_canvas;
    PP.lens = a; a += b_lens;
    PP.codes = (uint16_t*)a; a += b_codes;
    PP.adler = (uint32_t*)a; a += b_adler;
    PP.nmatch = (uint32_t*)a; a += b_nm;
    PP.fpx = (uint32_t*)a; a += b_fpx;
    PP.desc = (uint64_t*)a; a += b_desc;
    PP.rounds = rounds;
    PP.scratch = a; a += b_scr;
    PP.blen = (uint32_t*)a; a += b_blen;
    PP.off = (uint64_t*)a; a += b_off;
    PP.info = (uint64_t*)a; a += b_info;
    PP.file = out_host ? a : (uint8_t*)job->out_png;
    PP.cap = p_cap;
    PP.band_cap = band_cap;
    PP.w = cw; PP.h = ch; PP.nb = nb;
    PP.w_magic = (cw >= 2 && cw <= 16384) ? (uint32_t)(((1ull << 32) + (uint64_t)cw - 1) / (uint64_t)cw) : 0u;
    PP.rgba = *d_rgba;
    PP.pitch = *d_rgba_pitch;
    return STITCH_OK;
  }

  int run(StitchCtx* ctx, Lane& ln, const StitchJob* job, StitchResult* res, int cw, int ch, bool deferred) {
    stitch::png::HeaderBytes hdr;
    std::memset(&hdr, 0, sizeof hdr);
    const uint8_t sig[16] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A, 0, 0, 0, 13, 'I', 'H', 'D', 'R'};
    std::memcpy(hdr.b, sig, 16);
    auto be32 = [](uint8_t* q, uint32_t v) { q[0] = v >> 24; q[1] = v >> 16; q[2] = v >> 8; q[3] = (uint8_t)v; };
    be32(hdr.b + 16, (uint32_t)cw);
    be32(hdr.b + 20, (uint32_t)ch);
    hdr.b[24] = 8; hdr.b[25] = 6;   // 8 bit, colour type 6 (RGBA); compression 0, filter 0, no interlace
    {
      uint32_t c = 0xFFFFFFFFu;   // CRC-32 of 'IHDR' + 13 bytes
      for (int i = 12; i < 29; ++i) {
        c ^= hdr.b[i];
        for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      }
      be32(hdr.b + 29, c ^ 0xFFFFFFFFu);
    }
    hdr.n = 33;
    {
      // x^(2^k) mod P, reflected: x2n[0] = x, each next one the square of the previous
      auto mul = [](uint32_t a, uint32_t b) {
        uint32_t m = 1u << 31, p = 0;
        for (;;) {
          if (a & m) { p ^= b; if ((a & (m - 1)) == 0) break; }
          m >>= 1;
          b = (b & 1u) ? (b >> 1) ^ 0xEDB88320u : b >> 1;
        }
        return p;
      };
      uint32_t p = 1u << 30;
      PP.x2n[0] = p;
      for (int k = 1; k < 32; ++k) { p = mul(p, p); PP.x2n[k] = p; }
      // x^(-1) = x^(2^32 - 2) (x has order 2^32 - 1 modulo the CRC-32 polynomial), then its squares
      uint32_t inv = 1u << 31;
      for (int k = 1; k < 32; ++k) inv = mul(PP.x2n[k], inv);
      PP.xinv2n[0] = inv;
      for (int k = 1; k < 32; ++k) { inv = mul(inv, inv); PP.xinv2n[k] = inv; }
    }
    stitch::png::png_hist_kernel<<<PP.nb, PNG_THREADS, 0, ln.stream>>>(PP);
    stitch::png::png_encode_kernel<<<PP.nb, PNG_THREADS, 0, ln.stream>>>(PP);
    stitch::png::png_offsets_kernel<<<1, 1024, 0, ln.stream>>>(PP, hdr);
    stitch::png::png_gather_kernel<<<PP.nb, PNG_THREADS, 0, ln.stream>>>(PP);
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches += 4;
    res->kernel_launches += 4;
    CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_PNG_BYTES], PP.info, 16, cudaMemcpyDeviceToHost, ln.stream));
    if (deferred) return STITCH_OK;   // finish() runs when the job is finalised
    CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
    return finish(ctx, ln, res, out_host, PP.file, job->out_png);
  }

  static int finish(StitchCtx* ctx, Lane& ln, StitchResult* res, bool out_host, const uint8_t* p_file, void* out_png) {
    res->png_bytes = (int64_t)ln.info_host[INFO_PNG_BYTES];
    if (ln.info_host[INFO_PNG_OVERFLOW])
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (PNG output buffer too small; see png_bytes)");
    if (out_host) {
      CUDA_TRY(ctx, cudaMemcpyAsync(out_png, p_file, (size_t)res->png_bytes, cudaMemcpyDeviceToHost, ln.stream));
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
    }
    return STITCH_OK;
  }
};

// Uploads of PAGEABLE bitmaps through the lane's pinned ring (host_stage.h): the pool copies row chunks into ring slots,
// this thread sends each slot as soon as it is full and keeps the pool kSlots chunks ahead -- across image boundaries.
struct StagedUploads {
  struct Chunk { int img, r0, r1; };
  struct Src { uint8_t* dst; const uint8_t* src; int64_t dpitch, spitch; size_t row_bytes; };
  std::vector<Chunk> chunks;
  Src srcs[STITCH_MAX_IMAGES];
  size_t submitted = 0, issued = 0;
  stitch::UploadRing* ring = nullptr;
  uint64_t base = 0;

  void add(int img, void* dst, const void* src, int64_t dpitch, int64_t spitch, size_t row_bytes, int rows) {
    srcs[img] = Src{(uint8_t*)dst, (const uint8_t*)src, dpitch, spitch, row_bytes};
    const int per = (int)std::max<int64_t>(1, (int64_t)stitch::UploadRing::kChunk / dpitch);
    for (int r = 0; r < rows; r += per) chunks.push_back(Chunk{img, r, std::min(rows, r + per)});
  }
  void submit_ahead() {
    constexpr int K = stitch::UploadRing::kSlots;
    while (submitted < chunks.size() && submitted < issued + K) {
      const int slot = (int)((base + submitted) % K);
      if (ring->used[slot]) cudaEventSynchronize(ring->sent[slot]);   // the slot's previous contents have left
      ring->ready[slot].store(0, std::memory_order_relaxed);
      const Chunk c = chunks[submitted];
      const Src sr = srcs[c.img];
      uint8_t* pinned = ring->base + (size_t)slot * stitch::UploadRing::kChunk;
      std::atomic<int>* flag = &ring->ready[slot];
      stitch::CopyPool::get().submit([=] {
        stitch::copy_rows_in(pinned, sr.dpitch, sr.src, sr.spitch, sr.row_bytes, c.r0, c.r1);
        flag->store(1, std::memory_order_release);
      });
      ++submitted;
    }
  }
  // sends every chunk of `img` on stream `on`; returns when the last one has been handed to the DMA engine
  cudaError_t upload_image(int img, cudaStream_t on) {
    constexpr int K = stitch::UploadRing::kSlots;
    while (issued < chunks.size() && chunks[issued].img == img) {
      submit_ahead();
      const int slot = (int)((base + issued) % K);
      while (!ring->ready[slot].load(std::memory_order_acquire)) std::this_thread::yield();
      const Chunk c = chunks[issued];
      const Src& sr = srcs[c.img];
      cudaError_t e = cudaMemcpyAsync(sr.dst + (int64_t)c.r0 * sr.dpitch, ring->base + (size_t)slot * stitch::UploadRing::kChunk,
                                      (size_t)(c.r1 - c.r0) * (size_t)sr.dpitch, cudaMemcpyHostToDevice, on);
      if (e == cudaSuccess) e = cudaEventRecord(ring->sent[slot], on);
      if (e != cudaSuccess) return e;
      ring->used[slot] = true;
      ++issued;
    }
    submit_ahead();
    return cudaSuccess;
  }
  // on every way out: no pool task may still be writing into the ring, and the ring's counter moves on
  ~StagedUploads() {
    if (!ring) return;
    constexpr int K = stitch::UploadRing::kSlots;
    for (size_t g = issued; g < submitted; ++g) {
      const int slot = (int)((base + g) % K);
      while (!ring->ready[slot].load(std::memory_order_acquire)) std::this_thread::yield();
      ring->used[slot] = false;
    }
    ring->next = base + submitted;
  }
};

// defer == nullptr: everything that needs the host is done here (stitch_run).  Otherwise nothing waits for the GPU
// and *defer says what finalize_job still owes (stitch_submit).
int enqueue_job_impl(StitchCtx* ctx, Lane& ln, const StitchJob* job, StitchResult* res, PendingJob* defer);

// A job that fails half-way may already have copies in flight that read the caller's bitmaps or write the caller's
// canvas: the error is only returned once the lane's streams (and the host copy threads) are done with them.
int enqueue_job(StitchCtx* ctx, Lane& ln, const StitchJob* job, StitchResult* res, PendingJob* defer) {
  const int st = enqueue_job_impl(ctx, ln, job, res, defer);
  if (st != STITCH_OK && ctx && st != STITCH_E_BAD_ARG && st != STITCH_E_EMPTY && st != STITCH_E_TOO_MANY) {
    const std::string keep = ctx->last_error;
    if (ln.ul_stream) cudaStreamSynchronize(ln.ul_stream);
    if (ln.stream) cudaStreamSynchronize(ln.stream);
    if (ln.dl_stream) cudaStreamSynchronize(ln.dl_stream);
    if (ln.host_copies)
      while (ln.host_copies->load(std::memory_order_acquire) > 0) std::this_thread::yield();
    cudaGetLastError();
    ln.pngdec_active = false;
    ln.gifdec_active = false;
    ctx->last_error = keep;
  }
  return st;
}

int enqueue_job_impl(StitchCtx* ctx, Lane& ln, const StitchJob* job, StitchResult* res, PendingJob* defer) {
  if (!ctx || !job || !res) return fail(ctx, STITCH_E_BAD_ARG, "null argument");
  if (job->n <= 0) return fail(ctx, STITCH_E_EMPTY, "empty image list");  // P:1189
  if (!job->images) return fail(ctx, STITCH_E_BAD_ARG, "images == NULL");
  if (job->n > STITCH_MAX_IMAGES) return fail(ctx, STITCH_E_TOO_MANY, "more than 9 images");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int n = job->n;
  std::memset(res, 0, sizeof(*res));

  emit(ctx, 1);  // P:1193
  int32_t seq[2 * STITCH_MAX_IMAGES + 4];
  stitch_progress_sequence(n, seq, 2 * STITCH_MAX_IMAGES + 4);

  // [1] per-image prepare check (P:1225-1249).  A record without natural dims is re-probed by the page
  // (_prepareSingleImage -> getImageInfo, P:1227-1231, 724-734): for a FILE input the header is right here.
  StitchImageDesc descs[STITCH_MAX_IMAGES];
  int bw[STITCH_MAX_IMAGES], bh[STITCH_MAX_IMAGES];
  DecodeItem dec[STITCH_MAX_IMAGES];
  std::vector<stitch::pngdec::DecodeItem> pdec_store;
  std::vector<BmpItem> bdec_store;
  std::vector<GifItem> gdec_store;
  int n_dec = 0, n_pdec = 0, n_bdec = 0, n_gdec = 0;
  int kind[STITCH_MAX_IMAGES];
  for (int i = 0; i < n; ++i) {
    kind[i] = file_kind_of(job->images[i]);
    if (kind[i] < 0) {
      char buf[96];
      std::snprintf(buf, sizeof buf, "图片%d解码异常 (file type not decoded: JPEG, PNG, BMP and GIF are)", i);
      return fail(ctx, STITCH_E_BAD_IMAGE, buf);
    }
    if (kind[i] == STITCH_MEM_HOST_PNG && pdec_store.empty()) pdec_store.resize(STITCH_MAX_IMAGES);
    if (kind[i] == STITCH_MEM_HOST_BMP && bdec_store.empty()) bdec_store.resize(STITCH_MAX_IMAGES);
    if (kind[i] == STITCH_MEM_HOST_GIF && gdec_store.empty()) gdec_store.resize(STITCH_MAX_IMAGES);
  }
  stitch::pngdec::DecodeItem* pdec = pdec_store.data();
  BmpItem* bdec = bdec_store.data();
  GifItem* gdec = gdec_store.data();
  int st = STITCH_OK;
  for (int i = 0; i < n; ++i) {
    const StitchImage& im = job->images[i];
    descs[i] = im.desc;
    bw[i] = im.bmp_width;
    bh[i] = im.bmp_height;
    if (kind[i] == STITCH_MEM_HOST_GIF) {
      if (!im.data || im.pitch_bytes <= 0) {
        char buf[64];
        std::snprintf(buf, sizeof buf, "图片%d解码异常", i);
        return fail(ctx, STITCH_E_BAD_IMAGE, buf);
      }
      GifItem& d = gdec[n_gdec++];
      d.file = (const uint8_t*)im.data;
      d.bytes = (size_t)im.pitch_bytes;
      st = parse_gif(ctx, im.data, d.bytes, i, &d.hdr);
      if (st != STITCH_OK) return st;
      bw[i] = d.hdr.w;
      bh[i] = d.hdr.h;
      if (descs[i].natural_width <= 0 || descs[i].natural_height <= 0) {
        descs[i].natural_width = descs[i].width = d.hdr.w;      // P:729-732
        descs[i].natural_height = descs[i].height = d.hdr.h;
      }
      if (descs[i].file_size <= 0) descs[i].file_size = im.pitch_bytes;   // P:733
    } else if (kind[i] == STITCH_MEM_HOST_BMP) {
      if (!im.data || im.pitch_bytes <= 0) {
        char buf[64];
        std::snprintf(buf, sizeof buf, "图片%d解码异常", i);
        return fail(ctx, STITCH_E_BAD_IMAGE, buf);
      }
      BmpItem& d = bdec[n_bdec++];
      d.file = (const uint8_t*)im.data;
      d.bytes = (size_t)im.pitch_bytes;
      st = parse_bmp(ctx, im.data, d.bytes, i, &d.hdr);
      if (st != STITCH_OK) return st;
      bw[i] = d.hdr.w;
      bh[i] = d.hdr.h;
      if (descs[i].natural_width <= 0 || descs[i].natural_height <= 0) {
        descs[i].natural_width = descs[i].width = d.hdr.w;      // P:729-732
        descs[i].natural_height = descs[i].height = d.hdr.h;
      }
      if (descs[i].file_size <= 0) descs[i].file_size = im.pitch_bytes;   // P:733
    } else if (kind[i] == STITCH_MEM_HOST_JPEG) {
      if (!im.data || im.pitch_bytes <= 0) {
        char buf[64];
        std::snprintf(buf, sizeof buf, "图片%d解码异常", i);
        return fail(ctx, STITCH_E_BAD_IMAGE, buf);
      }
      DecodeItem& d = dec[n_dec++];
      d.file = (const uint8_t*)im.data;
      d.bytes = (size_t)im.pitch_bytes;
      st = parse_jpeg(ctx, im.data, d.bytes, i, &d.hdr);
      if (st != STITCH_OK) return st;
      bw[i] = d.hdr.w;
      bh[i] = d.hdr.h;
      if (descs[i].natural_width <= 0 || descs[i].natural_height <= 0) {
        descs[i].natural_width = descs[i].width = d.hdr.w;      // P:729-732
        descs[i].natural_height = descs[i].height = d.hdr.h;
      }
      if (descs[i].file_size <= 0) descs[i].file_size = im.pitch_bytes;   // P:733
    } else if (kind[i] == STITCH_MEM_HOST_PNG) {
      if (!im.data || im.pitch_bytes <= 0) {
        char buf[64];
        std::snprintf(buf, sizeof buf, "图片%d解码异常", i);
        return fail(ctx, STITCH_E_BAD_IMAGE, buf);
      }
      stitch::pngdec::DecodeItem& d = pdec[n_pdec++];
      d.file = (const uint8_t*)im.data;
      d.bytes = (size_t)im.pitch_bytes;
      st = parse_png(ctx, im.data, d.bytes, i, &d.hdr);
      if (st != STITCH_OK) return st;
      bw[i] = d.hdr.w;
      bh[i] = d.hdr.h;
      if (descs[i].natural_width <= 0 || descs[i].natural_height <= 0) {
        descs[i].natural_width = descs[i].width = d.hdr.w;      // P:729-732
        descs[i].natural_height = descs[i].height = d.hdr.h;
      }
      if (descs[i].file_size <= 0) descs[i].file_size = im.pitch_bytes;   // P:733
    }
    emit(ctx, seq[1 + i]);
  }
  // [2] layout (P:1251-1386)
  st = stitch_layout(descs, n, &job->params, &res->layout);
  if (st != STITCH_OK) return fail(ctx, st, stitch_strerror(st));
  emit(ctx, 30);  // P:1358
  const StitchLayout& L = res->layout;
  const int cw = (int)L.canvas_w, ch = (int)L.canvas_h;
  res->canvas_w = cw;
  res->canvas_h = ch;
  stitch_output_size(&L, &res->rgba_bytes, &res->yuv_bytes);

  int flags = job->output_flags ? job->output_flags : STITCH_OUT_RGBA;
  const bool want_rgba = (flags & STITCH_OUT_RGBA) != 0, want_yuv = (flags & STITCH_OUT_YUV420) != 0;
  const bool want_jpeg = (flags & STITCH_OUT_JPEG) != 0, want_png = (flags & STITCH_OUT_PNG) != 0;
  if (want_png && (!job->out_png || job->out_capacity_png < 128))
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (PNG output buffer missing or too small)");
  if (want_jpeg && (!job->out_jpeg || job->out_capacity_jpeg < 1024))
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (JPEG output buffer missing or too small)");
  if (want_jpeg && (cw > 65535 || ch > 65535)) return fail(ctx, STITCH_E_BAD_ARG, "JPEG export: canvas side above 65535");
  // [3] surface (U:131-150, P:1388-1413): the caller's buffer is the canvas.
  const int64_t rgba_pitch = job->out_rgba_pitch ? job->out_rgba_pitch : (int64_t)cw * 4;
  if (want_rgba) {
    if (!job->out_rgba || rgba_pitch < (int64_t)cw * 4 || (rgba_pitch & 3) ||
        job->out_capacity_rgba < rgba_pitch * (int64_t)(ch - 1) + (int64_t)cw * 4)
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (RGBA output buffer missing or too small)");
  }
  if (want_yuv) {
    if (!job->out_yuv || job->out_capacity_yuv < res->yuv_bytes)
      return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (YUV output buffer missing or too small)");
  }

  // bitmaps (P:1512-1514); a STITCH_MEM_HOST_JPEG image is a FILE the GPU decodes (loadImageFrom, U:27-121)
  for (int i = 0; i < n; ++i) {
    const StitchImage& im = job->images[i];
    if (kind[i] == STITCH_MEM_HOST_JPEG || kind[i] == STITCH_MEM_HOST_PNG || kind[i] == STITCH_MEM_HOST_BMP ||
        kind[i] == STITCH_MEM_HOST_GIF) continue;   // parsed above
    if (!im.data || im.bmp_width <= 0 || im.bmp_height <= 0) {
      char buf[64];
      std::snprintf(buf, sizeof buf, "图片%d解码异常", i);
      return fail(ctx, STITCH_E_BAD_IMAGE, buf);
    }
    const int64_t pitch = im.pitch_bytes ? im.pitch_bytes : (int64_t)im.bmp_width * 4;
    if (pitch < (int64_t)im.bmp_width * 4 || (pitch & 3)) return fail(ctx, STITCH_E_BAD_ARG, "bad bitmap pitch");
  }

  JobParam P;
  std::memset(&P, 0, sizeof P);
  P.n = n;
  P.cw = cw;
  P.ch = ch;
  const double ss = (L.super_sample != 1.0) ? L.super_sample : 1.0;  // P:1426

  // stage inputs.  Host bitmaps are only SCHEDULED here (uploads[]): the copies are issued below, either all in front
  // of the one fused launch or interleaved with per-image launches (stitch_run's pipelined form).
  struct Upload { void* dst; const void* src; int64_t dpitch, spitch; size_t row_bytes; int rows; bool linear; };
  Upload uploads[STITCH_MAX_IMAGES];
  bool has_upload[STITCH_MAX_IMAGES] = {};
  auto issue_upload = [&](int i, cudaStream_t on) -> cudaError_t {
    const Upload& u = uploads[i];
    if (u.linear) return cudaMemcpyAsync(u.dst, u.src, (size_t)u.dpitch * u.rows, cudaMemcpyHostToDevice, on);
    return cudaMemcpy2DAsync(u.dst, u.dpitch, u.src, u.spitch, u.row_bytes, u.rows, cudaMemcpyHostToDevice, on);
  };
  int k_dec = 0, k_pdec = 0, k_bdec = 0, k_gdec = 0, n_host = 0;
  int64_t host_bytes = 0;
  // pageable host memory is staged by the library itself, for the synchronous entry point (host_stage.h)
  static const bool host_stage_on = [] { const char* e = std::getenv("STITCH_HOST_STAGE"); return !(e && e[0] == '0'); }();
  const bool stage_ok = host_stage_on && defer == nullptr;
  StagedUploads staged;
  bool pend[STITCH_MAX_IMAGES] = {};
  if (stage_ok) ln.dl_events_used = 0;
  // stitch_run's pipelined form (see [4] below) is decided here so that bitmap 0 is on its way before the plan is made
  static const bool pipeline_on = [] { const char* e = std::getenv("STITCH_PIPELINE"); return !(e && e[0] == '0'); }();
  bool early = pipeline_on && defer == nullptr && n >= 2 && job->out_mem_kind != STITCH_MEM_DEVICE && want_rgba && !want_yuv &&
               !want_jpeg && !want_png;
  for (int i = 0; i < n && early; ++i) {
    const StitchImage& im = job->images[i];
    early = im.mem_kind == STITCH_MEM_HOST;
    host_bytes += (int64_t)im.bmp_width * 4 * im.bmp_height;
  }
  early = early && host_bytes >= (1 << 20);
  host_bytes = 0;
  if (early) {
    // three streams: uploads back to back on their own stream (the H2D engine never waits for a draw), draws on the
    // lane's stream behind their bitmap's event, downloads on a third (the D2H engine)
    if (!ln.dl_stream) CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ln.dl_stream, cudaStreamNonBlocking));
    if (!ln.ul_stream) CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ln.ul_stream, cudaStreamNonBlocking));
    if (!ln.ev_dl) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ln.ev_dl, cudaEventDisableTiming));
    if (!ln.ev_start) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ln.ev_start, cudaEventDisableTiming));
    for (int i = 0; i <= n; ++i)
      if (!ln.ev_draw[i]) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ln.ev_draw[i], cudaEventDisableTiming));
    for (int i = 0; i < n; ++i)
      if (!ln.ev_up[i]) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ln.ev_up[i], cudaEventDisableTiming));
  }
  for (int i = 0; i < n; ++i) {
    const StitchImage& im = job->images[i];
    int64_t pitch = im.pitch_bytes ? im.pitch_bytes : (int64_t)im.bmp_width * 4;
    const uint8_t* dptr = nullptr;
    if (kind[i] == STITCH_MEM_HOST_GIF) {
      const int64_t dpitch = ((int64_t)bw[i] * 4 + 15) & ~15LL;
      st = ensure_stage(ctx, ln, i, (size_t)(dpitch * bh[i]));
      if (st != STITCH_OK) return st;
      gdec[k_gdec].rgba = (uint8_t*)ln.in_stage[i];
      gdec[k_gdec].pitch = dpitch;
      ++k_gdec;
      dptr = (const uint8_t*)ln.in_stage[i];
      pitch = dpitch;
    } else if (kind[i] == STITCH_MEM_HOST_BMP) {
      const int64_t dpitch = ((int64_t)bw[i] * 4 + 15) & ~15LL;
      st = ensure_stage(ctx, ln, i, (size_t)(dpitch * bh[i]));
      if (st != STITCH_OK) return st;
      bdec[k_bdec].rgba = (uint8_t*)ln.in_stage[i];
      bdec[k_bdec].pitch = dpitch;
      ++k_bdec;
      dptr = (const uint8_t*)ln.in_stage[i];
      pitch = dpitch;
    } else if (kind[i] == STITCH_MEM_HOST_JPEG) {
      const int64_t dpitch = ((int64_t)bw[i] * 4 + 15) & ~15LL;
      st = ensure_stage(ctx, ln, i, (size_t)(dpitch * bh[i]));
      if (st != STITCH_OK) return st;
      dec[k_dec].rgba = (uint8_t*)ln.in_stage[i];
      dec[k_dec].pitch = dpitch;
      ++k_dec;
      dptr = (const uint8_t*)ln.in_stage[i];
      pitch = dpitch;
    } else if (kind[i] == STITCH_MEM_HOST_PNG) {
      const int64_t dpitch = ((int64_t)bw[i] * 4 + 15) & ~15LL;
      st = ensure_stage(ctx, ln, i, (size_t)(dpitch * bh[i]));
      if (st != STITCH_OK) return st;
      pdec[k_pdec].rgba = (uint8_t*)ln.in_stage[i];
      pdec[k_pdec].pitch = dpitch;
      ++k_pdec;
      dptr = (const uint8_t*)ln.in_stage[i];
      pitch = dpitch;
    } else if (im.mem_kind == STITCH_MEM_DEVICE) {
      dptr = (const uint8_t*)im.data;
    } else {
      // Host bitmap: stage with a 16B-aligned pitch so the streaming kernel's 128-bit loads apply.
      const int64_t dpitch = ((int64_t)im.bmp_width * 4 + 15) & ~15LL;
      st = ensure_stage(ctx, ln, i, (size_t)(dpitch * im.bmp_height));
      if (st != STITCH_OK) return st;
      // rows already contiguous: one linear DMA
      uploads[i] = Upload{ln.in_stage[i], im.data, dpitch, pitch, (size_t)im.bmp_width * 4, im.bmp_height,
                          pitch == dpitch && dpitch == (int64_t)im.bmp_width * 4};
      has_upload[i] = true;
      ++n_host;
      const int64_t img_bytes = (int64_t)im.bmp_width * 4 * im.bmp_height;
      host_bytes += img_bytes;
      if (early) {
        // the upload starts after whatever the lane's stream still holds for this buffer (a fresh staging's memset)
        CUDA_TRY(ctx, cudaEventRecord(ln.ev_start, ln.stream));
        CUDA_TRY(ctx, cudaStreamWaitEvent(ln.ul_stream, ln.ev_start, 0));
      }
      if (stage_ok && img_bytes >= (256 << 10) && stitch::is_pageable(im.data)) {
        // pageable memory: through the lane's pinned ring, filled by the copy pool (host_stage.h); sent below
        if (!ln.up_ring) ln.up_ring = new stitch::UploadRing();
        CUDA_TRY(ctx, ln.up_ring->ensure());
        if (!staged.ring) { staged.ring = ln.up_ring; staged.base = ln.up_ring->next; }
        staged.add(i, ln.in_stage[i], im.data, dpitch, pitch, (size_t)im.bmp_width * 4, im.bmp_height);
        pend[i] = true;
      } else if (early) {
        CUDA_TRY(ctx, issue_upload(i, ln.ul_stream));
        CUDA_TRY(ctx, cudaEventRecord(ln.ev_up[i], ln.ul_stream));
      }
      dptr = (const uint8_t*)ln.in_stage[i];
      pitch = dpitch;
    }
    P.img[i].data = dptr;
    P.img[i].pitch = pitch;
    make_placement(L.rects[i], im.desc.orientation, ss, bw[i], bh[i], cw, ch, &P.img[i]);
  }
  const bool deferred = defer != nullptr;
  if (n_dec > 0) {
    st = decode_batch(ctx, ln, dec, n_dec, deferred);
    if (st != STITCH_OK) return st;
  }
  if (n_pdec > 0) {
    st = decode_png_batch(ctx, ln, pdec, n_pdec);
    if (st != STITCH_OK) return st;
  }
  if (n_bdec > 0) {
    st = decode_bmp_batch(ctx, ln, bdec, n_bdec);
    if (st != STITCH_OK) return st;
  }
  if (n_gdec > 0) {
    st = decode_gif_batch(ctx, ln, gdec, n_gdec);
    if (st != STITCH_OK) return st;
  }

  // outputs
  uint8_t* d_rgba = nullptr;
  uint8_t* d_yuv = nullptr;
  int64_t d_rgba_pitch = rgba_pitch;
  const bool out_host = job->out_mem_kind != STITCH_MEM_DEVICE;
  if (out_host) {
    d_rgba_pitch = ((int64_t)cw * 4 + 15) & ~15LL;
    const size_t need_rgba = want_rgba ? (size_t)(d_rgba_pitch * ch) : 0;
    const size_t need = need_rgba + (want_yuv ? (size_t)res->yuv_bytes : 0) + 256;
    if (ln.out_stage_cap < need) {
      if (ln.out_stage) CUDA_TRY(ctx, cudaFree(ln.out_stage));
      ln.out_stage = nullptr;
      ln.out_stage_cap = 0;
      cudaError_t e = cudaMalloc(&ln.out_stage, need);
      if (e != cudaSuccess) return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (device canvas allocation failed)");
      ln.out_stage_cap = need;
    }
    if (want_rgba) d_rgba = (uint8_t*)ln.out_stage;
    if (want_yuv) d_yuv = (uint8_t*)ln.out_stage + ((need_rgba + 255) & ~(size_t)255);
  } else {
    if (want_rgba) d_rgba = (uint8_t*)job->out_rgba;
    if (want_yuv) d_yuv = (uint8_t*)job->out_yuv;
  }
  // JPEG / PNG exports carve their work buffers before the launch: the encoders read the I420 planes / the RGBA
  // canvas of this launch -- the caller's when requested too, buffers inside the lane's arena otherwise.
  const int64_t cwh = (cw + 1) / 2, chh = (ch + 1) / 2;
  JpegExport jx;
  PngExport px;
  if (want_jpeg) {
    st = jx.carve(ctx, ln, job, res, cw, ch, want_yuv, out_host, &d_yuv);
    if (st != STITCH_OK) return st;
  }
  if (want_png) {
    st = px.carve(ctx, ln, job, L, cw, ch, want_rgba, out_host, &d_rgba, &d_rgba_pitch);
    if (st != STITCH_OK) return st;
  }
  const bool planes = want_yuv || want_jpeg;
  const bool canvas = want_rgba || want_png;
  P.out_rgba = d_rgba;
  P.out_pitch = d_rgba_pitch;
  if (planes) {
    P.out_y = d_yuv;
    P.out_cb = d_yuv + (int64_t)cw * ch;
    P.out_cr = P.out_cb + cwh * chh;
  }

  // A pageable RGBA output (host_stage.h): rects go to a pinned copy of the canvas, and the pool copies them on into
  // the caller's memory behind a CUDA event -- in ~1 MB row slices, so a freshly allocated (never touched) output has its
  // page faults taken by several cores.  stitch_run waits for the slices after the stream.
  const bool out_pg = stage_ok && out_host && want_rgba && !want_png && (int64_t)cw * 4 * ch >= (256 << 10) &&
                      stitch::is_pageable(job->out_rgba);
  const int64_t hp = (int64_t)cw * 4;   // pitch of the pinned canvas
  if (out_pg) {
    const size_t need = (size_t)hp * (size_t)ch;
    if (ln.host_canvas_cap < need) {
      if (ln.host_canvas) CUDA_TRY(ctx, cudaFreeHost(ln.host_canvas));
      ln.host_canvas = nullptr;
      ln.host_canvas_cap = 0;
      CUDA_TRY(ctx, cudaMallocHost((void**)&ln.host_canvas, need));
      ln.host_canvas_cap = need;
    }
    if (!ln.host_copies) ln.host_copies = new std::atomic<int>(0);
  }
  auto download_staged = [&](int x0, int y0, int x1, int y1, cudaStream_t on) -> cudaError_t {
    if (x1 <= x0 || y1 <= y0) return cudaSuccess;
    const size_t wbytes = (size_t)(x1 - x0) * 4;
    uint8_t* pinned = ln.host_canvas + (int64_t)y0 * hp + (int64_t)x0 * 4;
    cudaError_t e = cudaMemcpy2DAsync(pinned, hp, d_rgba + (int64_t)y0 * d_rgba_pitch + (int64_t)x0 * 4, d_rgba_pitch, wbytes, y1 - y0,
                                      cudaMemcpyDeviceToHost, on);
    if (e != cudaSuccess) return e;
    if (ln.dl_events_used == ln.dl_events.size()) {
      cudaEvent_t ev;
      e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
      if (e != cudaSuccess) return e;
      ln.dl_events.push_back(ev);
    }
    const cudaEvent_t ev = ln.dl_events[ln.dl_events_used++];
    e = cudaEventRecord(ev, on);
    if (e != cudaSuccess) return e;
    const int per = (int)std::max<size_t>(1, ((size_t)1 << 20) / wbytes);
    uint8_t* user = (uint8_t*)job->out_rgba + (int64_t)y0 * rgba_pitch + (int64_t)x0 * 4;
    std::atomic<int>* counter = ln.host_copies;
    const int64_t up = rgba_pitch;
    for (int r = 0; r < y1 - y0; r += per) {
      const int r1 = std::min(y1 - y0, r + per);
      counter->fetch_add(1, std::memory_order_relaxed);
      stitch::CopyPool::get().submit([=] {
        cudaEventSynchronize(ev);
        if (up == hp && wbytes == (size_t)hp) stitch::copy_big(user + (int64_t)r * up, pinned + (int64_t)r * hp, (size_t)(r1 - r) * wbytes);
        else for (int y = r; y < r1; ++y) std::memcpy(user + (int64_t)y * up, pinned + (int64_t)y * hp, wbytes);
        counter->fetch_sub(1, std::memory_order_release);
      });
    }
    return cudaSuccess;
  };

  // [4] the draw loop (P:1439-1572)
  stitch::PreparedJob J;
  st = stitch::prepare_job(P, canvas, planes, ctx->sm_count, ln.scratch, ln.stream, &J);
  res->path_info = J.path_info;
  if (st != 0) return fail(ctx, STITCH_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString((cudaError_t)st));
  // stitch_run (the page's one job at a time, P:772,781) with host bitmaps in and a host RGBA canvas out is bound by the
  // PCIe copies: upload, draw and download in sequence take 9.6 ms for the 9 x 12 MP job against 8.0 ms of upload alone.
  // Pipelined form: image k's regions are drawn as soon as bitmap k has landed (a launch over a SUBSET of the fused
  // plan's regions: same units, same bytes) and its part of the canvas goes home on a second stream -- the other DMA
  // direction -- while bitmap k + 1 uploads; what is left (background, seams, images without a fast region) follows
  // the last upload.  The 30..90 progress values are then emitted when each image has really been drawn (P:1556-1557).
  const bool pipelined = early && J.fused;
  int launches = 0;
  if (!pipelined) {
    for (int i = 0; i < n; ++i) {
      if (!has_upload[i]) continue;
      if (pend[i]) {
        CUDA_TRY(ctx, staged.upload_image(i, early ? ln.ul_stream : ln.stream));
        if (early) CUDA_TRY(ctx, cudaEventRecord(ln.ev_up[i], ln.ul_stream));
      }
      if (early) CUDA_TRY(ctx, cudaStreamWaitEvent(ln.stream, ln.ev_up[i], 0));   // already on its way (no fast region after all)
      else if (!pend[i]) CUDA_TRY(ctx, issue_upload(i, ln.stream));
    }
    if (J.fused) {
      st = stitch::launch_plan(P, J, J.plan, ctx->sm_count, ln.scratch, ln.stream);
    } else {
      if (planes) {
        dim3 grid(((cw + 1) / 2 + 31) / 32, ((ch + 1) / 2 + 7) / 8);
        stitch::gather_yuv_kernel<<<grid, 256, 0, ln.stream>>>(P);
      } else {
        dim3 grid((cw + 31) / 32, (ch + 7) / 8);
        stitch::gather_rgba_kernel<<<grid, 256, 0, ln.stream>>>(P);
      }
      st = (int)cudaGetLastError();
    }
    if (st != 0) return fail(ctx, STITCH_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString((cudaError_t)st));
    launches = 1;
    ctx->launches += launches;
    res->kernel_launches = launches;
    for (int i = 0; i < n; ++i) emit(ctx, seq[2 + n + i]);  // P:1556-1557 (one fused launch covers all images)

    // [5] export (U:205-242): the canvas is copied out whole, unscaled (P:1577-1579)
    if (out_host) {
      if (want_rgba && out_pg) {
        CUDA_TRY(ctx, download_staged(0, 0, cw, ch, ln.stream));
      } else if (want_rgba) {
        if ((int64_t)rgba_pitch == (int64_t)d_rgba_pitch && (int64_t)rgba_pitch == (int64_t)cw * 4)
          CUDA_TRY(ctx, cudaMemcpyAsync(job->out_rgba, d_rgba, (size_t)cw * 4 * (size_t)ch, cudaMemcpyDeviceToHost, ln.stream));
        else
          CUDA_TRY(ctx, cudaMemcpy2DAsync(job->out_rgba, rgba_pitch, d_rgba, d_rgba_pitch, (size_t)cw * 4, ch,
                                          cudaMemcpyDeviceToHost, ln.stream));
      }
      if (want_yuv)
        CUDA_TRY(ctx, cudaMemcpyAsync(job->out_yuv, d_yuv, (size_t)res->yuv_bytes, cudaMemcpyDeviceToHost, ln.stream));
    }
  } else {
    // one canvas rect home (device pixel coordinates), on the download stream
    auto download = [&](int x0, int y0, int x1, int y1) -> cudaError_t {
      if (x1 <= x0 || y1 <= y0) return cudaSuccess;
      if (out_pg) return download_staged(x0, y0, x1, y1, ln.dl_stream);
      uint8_t* dst = (uint8_t*)job->out_rgba + (int64_t)y0 * rgba_pitch + (int64_t)x0 * 4;
      const uint8_t* src = d_rgba + (int64_t)y0 * d_rgba_pitch + (int64_t)x0 * 4;
      if (x0 == 0 && x1 == cw && rgba_pitch == d_rgba_pitch && rgba_pitch == (int64_t)cw * 4)   // whole rows: one linear DMA
        return cudaMemcpyAsync(dst, src, (size_t)cw * 4 * (size_t)(y1 - y0), cudaMemcpyDeviceToHost, ln.dl_stream);
      return cudaMemcpy2DAsync(dst, rgba_pitch, src, d_rgba_pitch, (size_t)(x1 - x0) * 4, y1 - y0, cudaMemcpyDeviceToHost, ln.dl_stream);
    };
    bool own_launch[STITCH_MAX_IMAGES] = {};
    Plan sub;
    for (int k = 0; k < n; ++k) {
      if (pend[k]) {
        CUDA_TRY(ctx, staged.upload_image(k, ln.ul_stream));
        CUDA_TRY(ctx, cudaEventRecord(ln.ev_up[k], ln.ul_stream));
      }
      CUDA_TRY(ctx, cudaStreamWaitEvent(ln.stream, ln.ev_up[k], 0));
      stitch::sub_plan(J.plan, [&](const Region& r) { return (r.kind == REGION_STREAM || r.kind == REGION_UPSCALE) && r.img == k; }, &sub);
      if (sub.n_units == 0) continue;   // no fast region: drawn with the rest
      st = stitch::launch_plan(P, J, sub, ctx->sm_count, ln.scratch, ln.stream);
      if (st != 0) return fail(ctx, STITCH_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString((cudaError_t)st));
      ++launches;
      own_launch[k] = true;
      CUDA_TRY(ctx, cudaEventRecord(ln.ev_draw[k], ln.stream));
      CUDA_TRY(ctx, cudaStreamWaitEvent(ln.dl_stream, ln.ev_draw[k], 0));
      // the canvas box of image k's fast regions (body + tail of one image make one box)
      int bx0 = cw, by0 = ch, bx1 = 0, by1 = 0;
      const ImgParam& im = P.img[k];
      for (int i = 0; i < sub.n_regions; ++i) {
        const Region& r = sub.r[i];
        int x0 = r.x0, x1 = r.x1, y0 = r.y0, y1 = r.y1;
        if (r.kind == REGION_STREAM) {   // source-order index ranges inside the image's device rect
          const int X0 = (int)im.X0, Y0 = (int)im.Y0, dw = (int)im.W, dh = (int)im.H;
          x0 = X0 + (im.flip_x ? dw - r.x1 : r.x0); x1 = X0 + (im.flip_x ? dw - r.x0 : r.x1);
          y0 = Y0 + (im.flip_y ? dh - r.y1 : r.y0); y1 = Y0 + (im.flip_y ? dh - r.y0 : r.y1);
        }
        bx0 = std::min(bx0, x0); by0 = std::min(by0, y0); bx1 = std::max(bx1, x1); by1 = std::max(by1, y1);
      }
      CUDA_TRY(ctx, download(bx0, by0, bx1, by1));
    }
    stitch::sub_plan(J.plan, [&](const Region& r) { return !((r.kind == REGION_STREAM || r.kind == REGION_UPSCALE) && own_launch[r.img]); }, &sub);
    if (sub.n_units > 0) {
      st = stitch::launch_plan(P, J, sub, ctx->sm_count, ln.scratch, ln.stream);
      if (st != 0) return fail(ctx, STITCH_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString((cudaError_t)st));
      ++launches;
    }
    CUDA_TRY(ctx, cudaEventRecord(ln.ev_draw[n], ln.stream));
    CUDA_TRY(ctx, cudaStreamWaitEvent(ln.dl_stream, ln.ev_draw[n], 0));
    for (int i = 0; i < sub.n_regions; ++i) {   // RGBA-only plans are made of pixel rects
      const Region& r = sub.r[i];
      CUDA_TRY(ctx, download(r.x0, r.y0, r.x1, r.y1));
    }
    CUDA_TRY(ctx, cudaEventRecord(ln.ev_dl, ln.dl_stream));
    CUDA_TRY(ctx, cudaStreamWaitEvent(ln.stream, ln.ev_dl, 0));   // the lane's stream ends when the last piece is home
    ctx->launches += launches;
    res->kernel_launches = launches;
    for (int k = 0; k < n; ++k) {   // P:1556-1557: image k's value when image k has been drawn
      if (ctx->progress) CUDA_TRY(ctx, cudaEventSynchronize(ln.ev_draw[own_launch[k] ? k : n]));
      emit(ctx, seq[2 + n + k]);
    }
  }
  // [5b] / [5c] export as a file (U:211-221): fileType 'jpg', or the 'png' the page asks for (P:1579).  The file
  // size has to reach the host: here for stitch_run, in finalize_job for a submitted job.
  if (want_jpeg) {
    st = jx.run(ctx, ln, job, res, P, cw, ch, deferred);
    if (st != STITCH_OK) return st;
  }
  if (want_png) {
    st = px.run(ctx, ln, job, res, cw, ch, deferred);
    if (st != STITCH_OK) return st;
  }
  if (deferred && (n_dec > 0 || want_jpeg || want_png)) {
    defer->deferred = true;
    defer->dec = n_dec > 0;
    defer->pngdec = n_pdec > 0;
    defer->jpeg = want_jpeg;
    defer->png = want_png;
    defer->out_host = out_host;
    defer->j_file = want_jpeg ? jx.j_file : nullptr;
    defer->p_file = want_png ? px.PP.file : nullptr;
    defer->job = *job;
    defer->images.assign(job->images, job->images + n);
  }
  return STITCH_OK;
}

// What a deferred job still needs from the host; leaves pj.status / pj.error set.  Frees the job's lane.
void finalize_job(StitchCtx* ctx, PendingJob& pj) {
  if (pj.finalized) return;
  pj.finalized = true;
  Lane& ln = ctx->lanes[pj.lane];
  cudaError_t e = cudaEventSynchronize(pj.done);
  cudaEventDestroy(pj.done);
  pj.done = nullptr;
  if (ln.busy_ticket == pj.ticket) ln.busy_ticket = 0;
  if (e != cudaSuccess) {
    pj.status = fail(ctx, STITCH_E_CUDA, std::string("event sync: ") + cudaGetErrorString(e));
  } else if (pj.deferred) {
    if (pj.dec && ((uint32_t)ln.info_host[INFO_DEC_CHANGED] | (uint32_t)ln.info_host[INFO_PDEC_CHANGED]) != 0u) {
      // the entropy decoder's chains had not converged after kDecRounds rounds: decode with as many as it takes
      pj.job.images = pj.images.data();
      StitchProgressFn keep = ctx->progress;
      ctx->progress = nullptr;   // the values were emitted at submit
      pj.status = enqueue_job(ctx, ln, &pj.job, pj.res, nullptr);
      ctx->progress = keep;
      if (pj.status == STITCH_OK && cudaStreamSynchronize(ln.stream) != cudaSuccess) pj.status = fail(ctx, STITCH_E_CUDA, "stream sync");
    } else {
      if (pj.jpeg) pj.status = JpegExport::finish(ctx, ln, pj.res, pj.out_host, pj.j_file, pj.job.out_jpeg);
      if (pj.png && pj.status == STITCH_OK) pj.status = PngExport::finish(ctx, ln, pj.res, pj.out_host, pj.p_file, pj.job.out_png);
    }
  }
  if (pj.status == STITCH_OK) pj.status = check_png_status(ctx, ln);
  ln.pngdec_active = false;
  ln.gifdec_active = false;
  if (pj.status != STITCH_OK) pj.error = ctx->last_error;
  if (pj.status == STITCH_OK) {
    emit(ctx, 96);   // P:1581
    emit(ctx, 100);  // P:1611
  }
}

// A lane's arenas and info block belong to its unfinalised deferred job: finish that one before the lane is reused.
void drain_lane(StitchCtx* ctx, Lane& ln) {
  if (!ln.busy_ticket) return;
  for (PendingJob& pj : ctx->pending)
    if (pj.ticket == ln.busy_ticket) { finalize_job(ctx, pj); break; }
  ln.busy_ticket = 0;
}

}  // namespace

extern "C" int stitch_ctx_create(int32_t device, void* stream, StitchCtx** out) {
  if (!out) return STITCH_E_BAD_ARG;
  *out = nullptr;
  // A context drives up to STITCH_MAX_LANES job streams (+ upload / download streams); with the default of 8 hardware
  // queues they alias, and a short kernel of one job waits behind another job's long one (progressive-JPEG jobs: 18
  // instead of 10 ms/job with eight in flight).  Takes effect when CUDA has not been initialised in the process yet;
  // a value the caller has set is left alone.
  setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count) return STITCH_E_CUDA;
  if (cudaSetDevice(device) != cudaSuccess) return STITCH_E_CUDA;
  StitchCtx* ctx = new StitchCtx();
  ctx->device = device;
  ctx->lanes.resize(1);
  Lane& l0 = ctx->lanes[0];
  if (stream) {
    l0.stream = (cudaStream_t)stream;
  } else {
    if (cudaStreamCreateWithFlags(&l0.stream, cudaStreamNonBlocking) != cudaSuccess) {
      delete ctx;
      return STITCH_E_CUDA;
    }
    l0.own_stream = true;
  }
  cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
  *out = ctx;
  return STITCH_OK;
}

extern "C" int stitch_ctx_destroy(StitchCtx* ctx) {
  if (!ctx) return STITCH_E_BAD_ARG;
  cudaSetDevice(ctx->device);
  for (Lane& ln : ctx->lanes)
    if (ln.stream) cudaStreamSynchronize(ln.stream);
  for (auto& p : ctx->pending)
    if (p.done) cudaEventDestroy(p.done);
  for (Lane& ln : ctx->lanes) free_lane(ln);
  delete ctx;
  return STITCH_OK;
}

extern "C" void* stitch_ctx_stream(StitchCtx* ctx) { return ctx ? (void*)ctx->lanes[0].stream : nullptr; }

extern "C" int stitch_ctx_set_lanes(StitchCtx* ctx, int32_t n) {
  if (!ctx || n < 1 || n > STITCH_MAX_LANES) return STITCH_E_BAD_ARG;
  if (ctx->busy || !ctx->pending.empty()) return fail(ctx, STITCH_E_BUSY, "jobs in flight");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  while ((int)ctx->lanes.size() > n) {
    free_lane(ctx->lanes.back());
    ctx->lanes.pop_back();
  }
  while ((int)ctx->lanes.size() < n) {
    Lane ln;
    CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ln.stream, cudaStreamNonBlocking));
    ln.own_stream = true;
    ctx->lanes.push_back(ln);
  }
  return STITCH_OK;
}

extern "C" int stitch_ctx_join(StitchCtx* ctx) {
  if (!ctx) return STITCH_E_BAD_ARG;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  for (size_t k = 1; k < ctx->lanes.size(); ++k) {
    Lane& ln = ctx->lanes[k];
    if (!ln.ev_start) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ln.ev_start, cudaEventDisableTiming));   // the lane's own event, reused
    CUDA_TRY(ctx, cudaEventRecord(ln.ev_start, ln.stream));
    CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->lanes[0].stream, ln.ev_start, 0));
  }
  return STITCH_OK;
}

extern "C" int stitch_set_progress_cb(StitchCtx* ctx, StitchProgressFn fn, void* user) {
  if (!ctx) return STITCH_E_BAD_ARG;
  ctx->progress = fn;
  ctx->progress_user = user;
  return STITCH_OK;
}

extern "C" int stitch_run(StitchCtx* ctx, const StitchJob* job, StitchResult* result) {
  if (!ctx) return STITCH_E_BAD_ARG;
  if (ctx->busy) return fail(ctx, STITCH_E_BUSY, "stitch already in progress");  // P:772,781
  ctx->busy = true;
  drain_lane(ctx, ctx->lanes[0]);
  int st = enqueue_job(ctx, ctx->lanes[0], job, result, nullptr);
  if (st == STITCH_OK) {
    cudaError_t e = cudaStreamSynchronize(ctx->lanes[0].stream);
    if (e != cudaSuccess) st = fail(ctx, STITCH_E_CUDA, std::string("stream sync: ") + cudaGetErrorString(e));
  }
  if (std::atomic<int>* hc = ctx->lanes[0].host_copies)   // slices of a pageable canvas still on their way (host_stage.h)
    while (hc->load(std::memory_order_acquire) > 0) std::this_thread::yield();
  if (st == STITCH_OK) st = check_png_status(ctx, ctx->lanes[0]);
  ctx->lanes[0].pngdec_active = false;
  ctx->lanes[0].gifdec_active = false;
  if (st == STITCH_OK) {
    emit(ctx, 96);   // P:1581
    emit(ctx, 100);  // P:1611
  } else if (st != STITCH_E_EMPTY) {
    emit(ctx, 0);  // P:1622
  }
  ctx->busy = false;
  return st;
}

// Preview draw (P:1593-1603): one image, cleared canvas, ctx.scale(dpr, dpr).  Reuses the fused compose
// launch with JobParam.transparent set: fully covered opaque pixels take the streaming / upscale units
// as in a stitch job, the fractional border and the empty margins become GATHER / FILL units.
extern "C" int stitch_preview(StitchCtx* ctx, const void* src, int32_t src_mem_kind, int32_t export_w, int32_t export_h,
                              int64_t src_pitch, double view_w, double view_h, double dpr, void* out_rgba,
                              int64_t out_capacity, int32_t out_mem_kind, StitchPreviewLayout* layout_out) {
  if (!ctx) return STITCH_E_BAD_ARG;
  if (ctx->busy) return fail(ctx, STITCH_E_BUSY, "context is busy");
  StitchPreviewLayout pl;
  int st = stitch_preview_layout(export_w, export_h, view_w, view_h, dpr, &pl);
  if (layout_out) *layout_out = pl;
  if (st != STITCH_OK) return fail(ctx, st, stitch_strerror(st));
  if (!src) return fail(ctx, STITCH_E_BAD_IMAGE, "图片0解码异常");
  int64_t pitch = src_pitch ? src_pitch : (int64_t)export_w * 4;
  if (pitch < (int64_t)export_w * 4 || (pitch & 3)) return fail(ctx, STITCH_E_BAD_ARG, "bad bitmap pitch");
  const int cw = pl.canvas_w, ch = pl.canvas_h;
  if (!out_rgba || out_capacity < (int64_t)cw * ch * 4)
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (preview output buffer missing or too small)");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Lane& ln = ctx->lanes[0];
  drain_lane(ctx, ln);

  JobParam P;
  std::memset(&P, 0, sizeof P);
  P.n = 1;
  P.cw = cw;
  P.ch = ch;
  P.transparent = 1;  // clearRect, P:1598
  const uint8_t* dptr = (const uint8_t*)src;
  if (src_mem_kind != STITCH_MEM_DEVICE) {
    const int64_t dpitch = ((int64_t)export_w * 4 + 15) & ~15LL;
    st = ensure_stage(ctx, ln, 0, (size_t)(dpitch * export_h));
    if (st != STITCH_OK) return st;
    CUDA_TRY(ctx, cudaMemcpy2DAsync(ln.in_stage[0], dpitch, src, pitch, (size_t)export_w * 4, export_h,
                                    cudaMemcpyHostToDevice, ln.stream));
    dptr = (const uint8_t*)ln.in_stage[0];
    pitch = dpitch;
  }
  P.img[0].data = dptr;
  P.img[0].pitch = pitch;
  StitchRect r;
  r.dx = pl.dx;
  r.dy = pl.dy;
  r.dw = (double)pl.pv_w;
  r.dh = (double)pl.pv_h;
  make_placement(r, 1, pl.dpr, export_w, export_h, cw, ch, &P.img[0]);

  const bool out_host = out_mem_kind != STITCH_MEM_DEVICE;
  int64_t d_pitch = (int64_t)cw * 4;
  uint8_t* d_out = (uint8_t*)out_rgba;
  if (out_host) {
    d_pitch = ((int64_t)cw * 4 + 15) & ~15LL;
    const size_t need = (size_t)(d_pitch * ch) + 256;
    if (ln.out_stage_cap < need) {
      CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
      if (ln.out_stage) CUDA_TRY(ctx, cudaFree(ln.out_stage));
      ln.out_stage = nullptr;
      ln.out_stage_cap = 0;
      if (cudaMalloc(&ln.out_stage, need) != cudaSuccess)
        return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (device canvas allocation failed)");
      ln.out_stage_cap = need;
    }
    d_out = (uint8_t*)ln.out_stage;
  }
  P.out_rgba = d_out;
  P.out_pitch = d_pitch;
  int launches = 0, path_info = 0;
  st = stitch::launch_job(P, true, false, ctx->sm_count, ln.scratch, ln.stream, &launches, &path_info);
  if (st != 0) return fail(ctx, STITCH_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString((cudaError_t)st));
  ctx->launches += launches;
  if (out_host)
    CUDA_TRY(ctx, cudaMemcpy2DAsync(out_rgba, (size_t)cw * 4, d_out, d_pitch, (size_t)cw * 4, ch, cudaMemcpyDeviceToHost,
                                    ln.stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
  return STITCH_OK;
}

// loadImageFrom (U:27-121) for one JPEG file: header parse on the host, everything else on the GPU.
extern "C" int stitch_decode_jpeg(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                                  int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info) {
  if (!ctx) return STITCH_E_BAD_ARG;
  if (ctx->busy) return fail(ctx, STITCH_E_BUSY, "context is busy");
  if (!file) return fail(ctx, STITCH_E_BAD_IMAGE, "图片0解码异常");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Lane& ln = ctx->lanes[0];
  drain_lane(ctx, ln);
  DecodeItem d;
  d.file = (const uint8_t*)file;
  d.bytes = bytes;
  int st = parse_jpeg(ctx, file, bytes, 0, &d.hdr);
  if (info) {
    std::memset(info, 0, sizeof(*info));
    StitchImageInfo probe;
    if (stitch_probe_image(file, bytes, &probe) == STITCH_OK) *info = probe;
  }
  if (st != STITCH_OK) return st;
  const int w = d.hdr.w, h = d.hdr.h;
  const int64_t pitch = out_pitch ? out_pitch : (int64_t)w * 4;
  if (!out_rgba || pitch < (int64_t)w * 4 || (pitch & 3) || out_capacity < pitch * (int64_t)(h - 1) + (int64_t)w * 4)
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (decode output buffer missing or too small)");
  const bool out_host = out_mem_kind != STITCH_MEM_DEVICE;
  if (out_host) {
    const int64_t dpitch = ((int64_t)w * 4 + 15) & ~15LL;
    st = ensure_stage(ctx, ln, 0, (size_t)(dpitch * h));
    if (st != STITCH_OK) return st;
    d.rgba = (uint8_t*)ln.in_stage[0];
    d.pitch = dpitch;
  } else {
    d.rgba = (uint8_t*)out_rgba;
    d.pitch = pitch;
  }
  st = decode_batch(ctx, ln, &d, 1, false);
  if (st != STITCH_OK) return st;
  if (out_host)
    CUDA_TRY(ctx, cudaMemcpy2DAsync(out_rgba, pitch, d.rgba, d.pitch, (size_t)w * 4, h, cudaMemcpyDeviceToHost, ln.stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
  return STITCH_OK;
}

extern "C" int stitch_decode_png(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                                 int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info) {
  if (!ctx) return STITCH_E_BAD_ARG;
  if (ctx->busy) return fail(ctx, STITCH_E_BUSY, "context is busy");
  if (!file) return fail(ctx, STITCH_E_BAD_IMAGE, "图片0解码异常");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Lane& ln = ctx->lanes[0];
  drain_lane(ctx, ln);
  ln.gifdec_active = ln.pngdec_active = false;   // nothing of an earlier, failed call is left to check
  stitch::pngdec::DecodeItem d;
  d.file = (const uint8_t*)file;
  d.bytes = bytes;
  int st = parse_png(ctx, file, bytes, 0, &d.hdr);
  if (info) {
    std::memset(info, 0, sizeof(*info));
    StitchImageInfo probe;
    if (stitch_probe_image(file, bytes, &probe) == STITCH_OK) *info = probe;
  }
  if (st != STITCH_OK) return st;
  const int w = d.hdr.w, h = d.hdr.h;
  const int64_t pitch = out_pitch ? out_pitch : (int64_t)w * 4;
  if (!out_rgba || pitch < (int64_t)w * 4 || (pitch & 3) || out_capacity < pitch * (int64_t)(h - 1) + (int64_t)w * 4)
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (decode output buffer missing or too small)");
  const bool out_host = out_mem_kind != STITCH_MEM_DEVICE;
  if (out_host) {
    const int64_t dpitch = ((int64_t)w * 4 + 15) & ~15LL;
    st = ensure_stage(ctx, ln, 0, (size_t)(dpitch * h));
    if (st != STITCH_OK) return st;
    d.rgba = (uint8_t*)ln.in_stage[0];
    d.pitch = dpitch;
  } else {
    d.rgba = (uint8_t*)out_rgba;
    d.pitch = pitch;
  }
  st = decode_png_batch(ctx, ln, &d, 1);
  if (st != STITCH_OK) return st;
  if (out_host)
    CUDA_TRY(ctx, cudaMemcpy2DAsync(out_rgba, pitch, d.rgba, d.pitch, (size_t)w * 4, h, cudaMemcpyDeviceToHost, ln.stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
  return check_png_status(ctx, ln);
}

extern "C" int stitch_decode_bmp(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                                 int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info) {
  if (!ctx) return STITCH_E_BAD_ARG;
  if (ctx->busy) return fail(ctx, STITCH_E_BUSY, "context is busy");
  if (!file) return fail(ctx, STITCH_E_BAD_IMAGE, "图片0解码异常");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Lane& ln = ctx->lanes[0];
  drain_lane(ctx, ln);
  BmpItem d;
  d.file = (const uint8_t*)file;
  d.bytes = bytes;
  int st = parse_bmp(ctx, file, bytes, 0, &d.hdr);
  if (info) {
    std::memset(info, 0, sizeof(*info));
    StitchImageInfo probe;
    if (stitch_probe_image(file, bytes, &probe) == STITCH_OK) *info = probe;
  }
  if (st != STITCH_OK) return st;
  const int w = d.hdr.w, h = d.hdr.h;
  const int64_t pitch = out_pitch ? out_pitch : (int64_t)w * 4;
  if (!out_rgba || pitch < (int64_t)w * 4 || (pitch & 3) || out_capacity < pitch * (int64_t)(h - 1) + (int64_t)w * 4)
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (decode output buffer missing or too small)");
  const bool out_host = out_mem_kind != STITCH_MEM_DEVICE;
  if (out_host) {
    const int64_t dpitch = ((int64_t)w * 4 + 15) & ~15LL;
    st = ensure_stage(ctx, ln, 0, (size_t)(dpitch * h));
    if (st != STITCH_OK) return st;
    d.rgba = (uint8_t*)ln.in_stage[0];
    d.pitch = dpitch;
  } else {
    d.rgba = (uint8_t*)out_rgba;
    d.pitch = pitch;
  }
  st = decode_bmp_batch(ctx, ln, &d, 1);
  if (st != STITCH_OK) return st;
  if (out_host)
    CUDA_TRY(ctx, cudaMemcpy2DAsync(out_rgba, pitch, d.rgba, d.pitch, (size_t)w * 4, h, cudaMemcpyDeviceToHost, ln.stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
  return STITCH_OK;
}

extern "C" int stitch_decode_gif(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                                 int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info) {
  if (!ctx) return STITCH_E_BAD_ARG;
  if (ctx->busy) return fail(ctx, STITCH_E_BUSY, "context is busy");
  if (!file) return fail(ctx, STITCH_E_BAD_IMAGE, "图片0解码异常");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Lane& ln = ctx->lanes[0];
  drain_lane(ctx, ln);
  ln.gifdec_active = ln.pngdec_active = false;   // nothing of an earlier, failed call is left to check
  GifItem d;
  d.file = (const uint8_t*)file;
  d.bytes = bytes;
  int st = parse_gif(ctx, file, bytes, 0, &d.hdr);
  if (info) {
    std::memset(info, 0, sizeof(*info));
    StitchImageInfo probe;
    if (stitch_probe_image(file, bytes, &probe) == STITCH_OK) *info = probe;
  }
  if (st != STITCH_OK) return st;
  const int w = d.hdr.w, h = d.hdr.h;
  const int64_t pitch = out_pitch ? out_pitch : (int64_t)w * 4;
  if (!out_rgba || pitch < (int64_t)w * 4 || (pitch & 3) || out_capacity < pitch * (int64_t)(h - 1) + (int64_t)w * 4)
    return fail(ctx, STITCH_E_NO_SURFACE, "无法获取绘图上下文 (decode output buffer missing or too small)");
  const bool out_host = out_mem_kind != STITCH_MEM_DEVICE;
  if (out_host) {
    const int64_t dpitch = ((int64_t)w * 4 + 15) & ~15LL;
    st = ensure_stage(ctx, ln, 0, (size_t)(dpitch * h));
    if (st != STITCH_OK) return st;
    d.rgba = (uint8_t*)ln.in_stage[0];
    d.pitch = dpitch;
  } else {
    d.rgba = (uint8_t*)out_rgba;
    d.pitch = pitch;
  }
  st = decode_gif_batch(ctx, ln, &d, 1);
  if (st != STITCH_OK) return st;
  if (out_host)
    CUDA_TRY(ctx, cudaMemcpy2DAsync(out_rgba, pitch, d.rgba, d.pitch, (size_t)w * 4, h, cudaMemcpyDeviceToHost, ln.stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
  return check_png_status(ctx, ln);
}

extern "C" int stitch_decode_image(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                                   int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info) {
  if (!ctx) return STITCH_E_BAD_ARG;
  StitchImage im;
  std::memset(&im, 0, sizeof im);
  im.data = file;
  im.pitch_bytes = (int64_t)bytes;
  im.mem_kind = STITCH_MEM_HOST_FILE;
  switch (file_kind_of(im)) {
    case STITCH_MEM_HOST_JPEG: return stitch_decode_jpeg(ctx, file, bytes, out_rgba, out_capacity, out_pitch, out_mem_kind, info);
    case STITCH_MEM_HOST_PNG: return stitch_decode_png(ctx, file, bytes, out_rgba, out_capacity, out_pitch, out_mem_kind, info);
    case STITCH_MEM_HOST_BMP: return stitch_decode_bmp(ctx, file, bytes, out_rgba, out_capacity, out_pitch, out_mem_kind, info);
    case STITCH_MEM_HOST_GIF: return stitch_decode_gif(ctx, file, bytes, out_rgba, out_capacity, out_pitch, out_mem_kind, info);
    default: break;
  }
  if (info) {
    std::memset(info, 0, sizeof(*info));
    StitchImageInfo probe;
    if (file && stitch_probe_image(file, bytes, &probe) == STITCH_OK) *info = probe;
  }
  return fail(ctx, STITCH_E_BAD_IMAGE, "图片0解码异常 (file type not decoded: JPEG, PNG, BMP and GIF are)");
}

extern "C" int stitch_submit(StitchCtx* ctx, const StitchJob* job, StitchResult* result, uint64_t* ticket) {
  if (!ctx || !ticket) return STITCH_E_BAD_ARG;
  const size_t li = ctx->next_ticket % ctx->lanes.size();   // jobs rotate over the lanes
  Lane& ln = ctx->lanes[li];
  drain_lane(ctx, ln);
  PendingJob pj;
  int st = enqueue_job(ctx, ln, job, result, &pj);
  if (st != STITCH_OK) return st;
  pj.ticket = ctx->next_ticket++;
  pj.n = job->n;
  pj.lane = (int32_t)li;
  pj.res = result;
  CUDA_TRY(ctx, cudaEventCreateWithFlags(&pj.done, cudaEventDisableTiming));
  CUDA_TRY(ctx, cudaEventRecord(pj.done, ln.stream));
  if (pj.deferred) ln.busy_ticket = pj.ticket;
  ctx->pending.push_back(std::move(pj));
  *ticket = ctx->pending.back().ticket;
  return STITCH_OK;
}

extern "C" int stitch_wait(StitchCtx* ctx, uint64_t ticket) {
  if (!ctx) return STITCH_E_BAD_ARG;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  // waiting for ticket T retires every earlier one too (tickets of one lane complete in order); the first failure
  // among them is what the call reports
  int st = STITCH_OK;
  while (!ctx->pending.empty() && ctx->pending.front().ticket <= ticket) {
    finalize_job(ctx, ctx->pending.front());
    PendingJob& pj = ctx->pending.front();
    if (st == STITCH_OK && pj.status != STITCH_OK) {
      st = pj.status;
      ctx->last_error = pj.error;
    }
    ctx->pending.pop_front();
  }
  return st;
}

extern "C" int stitch_host_alloc(size_t bytes, void** out) {
  if (!out) return STITCH_E_BAD_ARG;
  return cudaMallocHost(out, bytes) == cudaSuccess ? STITCH_OK : STITCH_E_CUDA;
}
extern "C" int stitch_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? STITCH_OK : STITCH_E_CUDA; }

extern "C" int stitch_device_alloc(StitchCtx* ctx, size_t bytes, void** out) {
  if (!ctx || !out) return STITCH_E_BAD_ARG;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaMalloc(out, bytes));
  return STITCH_OK;
}
extern "C" int stitch_device_free(StitchCtx* ctx, void* p) {
  if (!ctx) return STITCH_E_BAD_ARG;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaFree(p));
  return STITCH_OK;
}
extern "C" int stitch_memcpy_h2d(StitchCtx* ctx, void* dst, const void* src, size_t bytes) {
  if (!ctx) return STITCH_E_BAD_ARG;
  CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->lanes[0].stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->lanes[0].stream));
  return STITCH_OK;
}
extern "C" int stitch_memcpy_d2h(StitchCtx* ctx, void* dst, const void* src, size_t bytes) {
  if (!ctx) return STITCH_E_BAD_ARG;
  CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->lanes[0].stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->lanes[0].stream));
  return STITCH_OK;
}

extern "C" int stitch_generate_image(StitchCtx* ctx, void* dev_dst, int32_t width, int32_t height, int64_t pitch_bytes,
                                     uint64_t seed, int32_t pattern) {
  if (!ctx || !dev_dst || width <= 0 || height <= 0) return STITCH_E_BAD_ARG;
  if (pitch_bytes == 0) pitch_bytes = (int64_t)width * 4;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  cudaError_t e = stitch::launch_generate((uint8_t*)dev_dst, width, height, pitch_bytes, seed, pattern, ctx->lanes[0].stream);
  if (e != cudaSuccess) return fail(ctx, STITCH_E_CUDA, cudaGetErrorString(e));
  ctx->launches += 1;
  return STITCH_OK;
}

extern "C" int stitch_device_digest(StitchCtx* ctx, const void* dev_ptr, size_t bytes, uint64_t* out) {
  if (!ctx || !dev_ptr || !out) return STITCH_E_BAD_ARG;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Lane& ln = ctx->lanes[0];
  drain_lane(ctx, ln);
  int st = ensure_info(ctx, ln);
  if (st != STITCH_OK) return st;
  if (!ln.scratch.p) {   // same 256-byte block as the unit scheduler counters: words 0-1 are the scheduler's, word 8.. ours
    CUDA_TRY(ctx, cudaMalloc(&ln.scratch.p, 256));
    ln.scratch.cap = 256;
    CUDA_TRY(ctx, cudaMemsetAsync(ln.scratch.p, 0, 256, ln.stream));
  }
  unsigned long long* acc = (unsigned long long*)((uint8_t*)ln.scratch.p + 64);
  CUDA_TRY(ctx, cudaMemsetAsync(acc, 0, 8, ln.stream));
  stitch::digest_kernel<<<ctx->sm_count * 4, 256, 0, ln.stream>>>((const uint8_t*)dev_ptr, bytes, acc);
  CUDA_TRY(ctx, cudaGetLastError());
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaMemcpyAsync(&ln.info_host[INFO_WORDS - 1], acc, 8, cudaMemcpyDeviceToHost, ln.stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ln.stream));
  *out = ln.info_host[INFO_WORDS - 1];
  return STITCH_OK;
}

extern "C" const char* stitch_last_error(StitchCtx* ctx) { return ctx ? ctx->last_error.c_str() : ""; }
extern "C" int64_t stitch_kernel_launch_count(StitchCtx* ctx) { return ctx ? ctx->launches : 0; }
