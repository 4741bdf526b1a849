/*
 * libstitch.so -- C ABI of the B200-native stitch path.
 *
 * This is the drop-in boundary for ONE path of Iamctb/ImageStitching: the
 * `onStitch()` pipeline of the WeChat page (layout -> per-image drawImage loop
 * -> export).  The reference has no FFI of its own (it is a single JS Page), so
 * every entry point cites the reference lines it replaces.  Reference paths are
 * relative to /root/reference/miniprogram-stitch/miniprogram/ :
 *     P = pages/index/index.js      U = utils/canvas.js
 *
 * Rules of the ABI: extern "C", plain pointers and sizes, caller owns every
 * buffer it passes in, the library never frees caller memory, nothing throws
 * across the boundary (all functions return a StitchStatus), a StitchCtx is
 * bound to one GPU and is NOT thread-safe (use one ctx per thread/GPU).
 * There is no CPU fallback: every pixel is produced by sm_100a kernels.
 */
#ifndef STITCH_H_
#define STITCH_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STITCH_ABI_VERSION 8
#define STITCH_MAX_IMAGES 9 /* P:311  MAX_IMAGES: 9 */

/* ---- status codes: one per reference throw site (P:1618-1624 is the catch) ---- */
typedef enum StitchStatus {
  STITCH_OK = 0,
  STITCH_E_EMPTY = 1,      /* P:1189  empty list -> silent return upstream          */
  STITCH_E_BAD_DIMS = 2,   /* P:1254  '图片尺寸不可用'                               */
  STITCH_E_BAD_OUT = 3,    /* P:1320  '输出尺寸计算失败'                             */
  STITCH_E_NO_SURFACE = 4, /* P:1412  '无法获取绘图上下文' == canvas alloc failure    */
  STITCH_E_BAD_IMAGE = 5,  /* P:1513  '图片${idx}解码异常' == null/zero-size bitmap   */
  STITCH_E_TOO_MANY = 6,   /* P:1081-1091 intake caps the list at MAX_IMAGES        */
  STITCH_E_BAD_ARG = 7,    /* null pointer / unknown enum value at the ABI itself   */
  STITCH_E_CUDA = 8,       /* CUDA runtime error; text via stitch_last_error()      */
  STITCH_E_BUSY = 9        /* P:772,781 isStitching re-entrancy guard               */
} StitchStatus;

typedef enum StitchDirection { STITCH_VERTICAL = 0, STITCH_HORIZONTAL = 1 } StitchDirection; /* P:16,777,786 */
/* P:19-20; any string other than 'min'/'max' behaves as STITCH_MODE_OTHER (P:1280,1307 vs P:1527,1541) */
typedef enum StitchMode { STITCH_MODE_MIN = 0, STITCH_MODE_MAX = 1, STITCH_MODE_ORIGINAL = 2, STITCH_MODE_OTHER = 3 } StitchMode;
typedef enum StitchPlatform { STITCH_IOS = 0, STITCH_ANDROID = 1, STITCH_OTHER = 2 } StitchPlatform; /* wx.getSystemInfoSync().platform, P:1324 */

/* Device canvas-limit profile: this.deviceMaxCanvasSize / deviceMaxCanvasPixels (P:126-156).
 * 0 means "unset" (JS falsy) and takes the reference's `||` fallbacks (P:1325-1336). */
typedef struct StitchProfile {
  int32_t platform;           /* StitchPlatform */
  int32_t max_canvas_size;    /* 0 = unset */
  int64_t max_canvas_pixels;  /* 0 = unset */
} StitchProfile;

/* Fills the reference's fallback profile for a platform (P:128-138). */
int stitch_profile_preset(int32_t platform, StitchProfile* out);

/* The fields of the prepared-image record the layout reads (P:724-739). 0 = missing (JS falsy). */
typedef struct StitchImageDesc {
  int32_t natural_width, natural_height; /* naturalWidth / naturalHeight */
  int32_t width, height;                 /* width / height (fallbacks, P:1243-1244) */
  int32_t orientation;                   /* numeric EXIF 1..8; anything else draws plain (U:155,198) */
  int32_t reserved;
  int64_t file_size;                     /* fileSize, feeds bigTask (P:1211-1212) */
} StitchImageDesc;

typedef struct StitchParams {
  int32_t direction;     /* StitchDirection */
  int32_t mode;          /* StitchMode: verticalStitchMode / horizontalStitchMode (P:1257) */
  double gap;            /* this.data.gap (P:1187,1256); UI clamps to 0..20 (P:825,852) */
  StitchProfile profile;
} StitchParams;

typedef struct StitchRect { double dx, dy, dw, dh; } StitchRect; /* user-space drawImage dest rect (P:1526-1554) */

/* Everything P:1251-1386 + P:1432-1554 computes, as IEEE doubles exactly as the JS holds them. */
typedef struct StitchLayout {
  int32_t n;
  int32_t big_task;                 /* P:1212 */
  double max_side, max_pixels;      /* MAX_CANVAS_SIDE, MAX_TOTAL_PIXELS (P:1325-1336) */
  double pre_w, pre_h;              /* outW/outH after Math.round, before clamp (P:1318-1319) */
  double scale_down;                /* P:1337-1349 */
  double target_w, target_h;        /* P:1352-1353,1360-1361 */
  double super_sample;              /* P:1364-1383 */
  double canvas_w, canvas_h;        /* canvasOutW/H (P:1373-1383) */
  double scaled_gap;                /* P:1432 */
  StitchRect rects[STITCH_MAX_IMAGES];
  int32_t natural_w[STITCH_MAX_IMAGES], natural_h[STITCH_MAX_IMAGES]; /* P:1522-1523 */
} StitchLayout;

/* Pure CPU, no CUDA: bit-exact twin of P:1211-1213,1243-1386,1432-1433,1522-1554. */
int stitch_layout(const StitchImageDesc* images, int32_t n, const StitchParams* params, StitchLayout* out);

/* stitchProgress values of a successful run (P:1193,1247,1358,1556-1557,1581,1611).
 * Writes up to cap values, returns how many the sequence has (2n+4). */
int stitch_progress_sequence(int32_t n, int32_t* out, int32_t cap);

/* ---- execution ---- */
typedef struct StitchCtx StitchCtx;

/* One ctx per GPU (replaces the page's canvas node + OffscreenCanvas, U:131-150).
 * `stream` is a cudaStream_t the ctx should launch on, or NULL to create its own. */
int stitch_ctx_create(int32_t device, void* stream, StitchCtx** out);
int stitch_ctx_destroy(StitchCtx* ctx);
void* stitch_ctx_stream(StitchCtx* ctx); /* the cudaStream_t stitch_run launches on (lane 0) */

/* Job lanes.  The reference draws one job at a time (isStitching, P:772,781); a batch of independent
 * jobs (BASELINE configs[4]) is instead rotated by stitch_submit over `n` lanes -- streams with their
 * own staging and unit-scheduler counters -- so the next job's blocks fill the SMs the finishing job
 * frees and, with host buffers, one lane's export (D2H) overlaps another lane's upload (H2D).
 * Lane 0 is the ctx stream; lanes 1..n-1 are streams the ctx creates.  Default 1.
 * stitch_ctx_join makes lane 0's stream wait (on the device, without blocking the host) for everything
 * submitted so far on the other lanes, so one event on the ctx stream brackets a whole batch. */
#define STITCH_MAX_LANES 8
int stitch_ctx_set_lanes(StitchCtx* ctx, int32_t n);
int stitch_ctx_join(StitchCtx* ctx);

/* Progress callback == this.setData({stitchProgress}) (values of stitch_progress_sequence);
 * invoked on the calling thread inside stitch_run/stitch_wait. */
typedef void (*StitchProgressFn)(void* user, int32_t percent);
int stitch_set_progress_cb(StitchCtx* ctx, StitchProgressFn fn, void* user);

/* STITCH_MEM_HOST_JPEG: `data` is a JPEG FILE (baseline or progressive) in host memory and pitch_bytes its length in bytes; the
 * library decodes it on the GPU (loadImageFrom, U:27-121) -- bmp_width / bmp_height are read from the file. */
/* STITCH_MEM_HOST_PNG: the same for a PNG FILE (plain or Adam7-interlaced; gray / RGB / palette / alpha, 1-16 bit, tRNS). */
/* STITCH_MEM_HOST_BMP: the same for a BMP FILE (1 / 4 / 8 bit palettised, 16 / 24 / 32 bit, bit fields; no RLE). */
/* STITCH_MEM_HOST_GIF: the same for a GIF FILE: its first frame (what createImage + drawImage paint). */
/* STITCH_MEM_HOST_FILE: an image FILE of any of these types: the library looks at its signature (what the page does
 * with a picked path: SUPPORTED_IMAGE_TYPES, P:4). */
typedef enum StitchMemKind {
  STITCH_MEM_HOST = 0, STITCH_MEM_DEVICE = 1, STITCH_MEM_HOST_JPEG = 2, STITCH_MEM_HOST_PNG = 3, STITCH_MEM_HOST_BMP = 4,
  STITCH_MEM_HOST_FILE = 5, STITCH_MEM_HOST_GIF = 6
} StitchMemKind;
typedef enum StitchOutputFlags {
  STITCH_OUT_RGBA = 1, STITCH_OUT_YUV420 = 2, STITCH_OUT_JPEG = 4, STITCH_OUT_PNG = 8
} StitchOutputFlags;

/* A decoded bitmap (what loadImageFrom hands to drawWithOrientation, P:1445-1463,1532):
 * RGBA8, straight (non-premultiplied) alpha, row pitch in bytes (>= 4*bmp_width). */
typedef struct StitchImage {
  StitchImageDesc desc;
  const void* data;       /* NULL -> STITCH_E_BAD_IMAGE (P:1512-1514) */
  int32_t bmp_width, bmp_height; /* bmp.width / bmp.height; 0 -> STITCH_E_BAD_IMAGE */
  int64_t pitch_bytes;    /* 0 = tightly packed */
  int32_t mem_kind;       /* StitchMemKind */
  int32_t reserved;
} StitchImage;

typedef struct StitchJob {
  const StitchImage* images;
  int32_t n;
  StitchParams params;
  int32_t output_flags;   /* StitchOutputFlags bitmask; 0 == STITCH_OUT_RGBA */
  int32_t out_mem_kind;   /* where out_rgba / out_yuv live */
  void* out_rgba;         /* canvas_w*canvas_h*4 bytes (pitch = out_rgba_pitch or 4*canvas_w); may be NULL if not requested */
  int64_t out_rgba_pitch;
  void* out_yuv;          /* planar I420: Y[h][w], Cb[ceil(h/2)][ceil(w/2)], Cr[...]; tightly packed */
  int64_t out_capacity_rgba, out_capacity_yuv; /* bytes available; too small -> STITCH_E_NO_SURFACE */
  /* STITCH_OUT_JPEG: the export file itself (U:205-242 with fileType 'jpg'): baseline JFIF, YCbCr 4:2:0,
   * encoded on the GPU from the I420 planes of the same launch.  Lives where out_mem_kind says. */
  void* out_jpeg;
  int64_t out_capacity_jpeg;      /* bytes available; stitch_jpeg_capacity() is always enough */
  int32_t jpeg_quality;           /* 1..100 (IJG scale); 0 = 100, the reference's `quality: 1` */
  int32_t jpeg_restart_interval;  /* MCUs between restart markers; 0 = default (8) */
  /* STITCH_OUT_PNG: the export file the page actually asks for (fileType 'png', P:1577-1579): RGBA 8-bit PNG of
   * the canvas, filtered and deflate-coded on the GPU; lossless, so decoding it returns the RGBA canvas exactly. */
  void* out_png;
  int64_t out_capacity_png;       /* bytes available; stitch_png_capacity() is always enough */
} StitchJob;

typedef struct StitchResult {
  StitchLayout layout;
  int32_t canvas_w, canvas_h;
  int64_t rgba_bytes, yuv_bytes;
  int32_t kernel_launches;    /* kernels launched for this job */
  int32_t path_info;          /* diagnostics: bits 0-7 streaming regions, 8-15 gather regions, 16-23 fill
                                 regions of the fused compose launch; bit 24 = whole-canvas gather launch;
                                 bits 25-30 upscale regions */
  int64_t jpeg_bytes;         /* size of the JPEG file written to out_jpeg (or needed, on STITCH_E_NO_SURFACE) */
  int64_t png_bytes;          /* size of the PNG file written to out_png (or needed, on STITCH_E_NO_SURFACE) */
} StitchResult;

/* Required output sizes for a layout (export region == whole canvas, U:206-209, P:1577-1579). */
int stitch_output_size(const StitchLayout* layout, int64_t* rgba_bytes, int64_t* yuv_bytes);
/* Worst-case size of the JPEG export of this layout (every coefficient coded at full length, every byte
 * stuffed): a capacity that can never be too small.  Real files are a fraction of it; a smaller capacity
 * is accepted and reported through STITCH_E_NO_SURFACE + result.jpeg_bytes (the size needed) if exceeded. */
int64_t stitch_jpeg_capacity(const StitchLayout* layout, int32_t restart_interval);

/* Worst-case size of the PNG export of this layout (every byte coded with a 15-bit code). */
int64_t stitch_png_capacity(const StitchLayout* layout);

/* Synchronous onStitch(): layout -> draw all images -> export into the caller's buffer. (P:1186-1633) */
int stitch_run(StitchCtx* ctx, const StitchJob* job, StitchResult* result);

/* Asynchronous form for job-level sharding: enqueue on the next lane, wait later.  Nothing in stitch_submit waits
 * for the GPU -- also for STITCH_MEM_HOST_JPEG inputs (the entropy decoder runs a fixed number of synchronisation
 * rounds gated on a device flag) and STITCH_OUT_JPEG / STITCH_OUT_PNG outputs (the file size is fetched, and a host
 * file copied out, when the job is waited for -- or when its lane is needed by a later submit).
 * Buffers AND *result must stay valid until stitch_wait(ticket) returns; result->jpeg_bytes / png_bytes and the
 * "output buffer too small" error of a file job are delivered by stitch_wait, which returns the first failure among
 * the jobs it retires (every job submitted before `ticket`). */
int stitch_submit(StitchCtx* ctx, const StitchJob* job, StitchResult* result, uint64_t* ticket);
int stitch_wait(StitchCtx* ctx, uint64_t ticket);

/* ---- intake: what wx.getImageInfo gives _prepareSingleImage (P:678-741) ----
 * Reads width, height, container type and EXIF orientation from the first bytes of an image FILE
 * (SUPPORTED_IMAGE_TYPES, P:4: jpg/jpeg, png, gif, bmp, webp).  Pure CPU. */
typedef enum StitchImageType {
  STITCH_TYPE_UNKNOWN = 0, STITCH_TYPE_JPEG = 1, STITCH_TYPE_PNG = 2, STITCH_TYPE_GIF = 3, STITCH_TYPE_BMP = 4,
  STITCH_TYPE_WEBP = 5
} StitchImageType;
typedef struct StitchImageInfo {
  int32_t width, height;     /* info.width / info.height (stored size, before any EXIF rotation) */
  int32_t type;              /* StitchImageType; stitch_type_name() gives info.type */
  int32_t orientation_exif;  /* EXIF tag 0x0112, 1..8; 1 when absent.  stitch_orientation_name() gives the
                                string WeChat reports ('up', 'right', ...) */
} StitchImageInfo;
/* STITCH_E_BAD_IMAGE when the bytes are not a recognised image header (the page falls back to a transcode,
 * P:684-694, then fails with the load error). */
int stitch_probe_image(const void* file, size_t bytes, StitchImageInfo* out);
const char* stitch_orientation_name(int32_t exif);
const char* stitch_type_name(int32_t type);
/* The prepared record of P:724-739 from a probe result and the file size.  orientation is 1: the page
 * stores WeChat's orientation STRING, which drawWithOrientation's numeric switch (U:160-199) never matches. */
int stitch_prepare_record(const StitchImageInfo* info, int64_t file_size, StitchImageDesc* out);

/* ---- decode (loadImageFrom, U:27-121; P:1441-1520) ----
 * Decodes one JPEG file on the GPU into RGBA8 (alpha 255, EXIF orientation NOT applied -- the page hands the
 * un-rotated bitmap to drawWithOrientation).  Huffman-coded sequential (SOF0 / SOF1) and progressive (SOF2: any
 * script of interleaved DC scans and single-component scans) DCT, 8 bit, gray or YCbCr with 4:4:4 / 4:2:2 / 4:2:0
 * sampling, restart markers; integer IDCT, triangle-filter chroma upsampling and fixed-point colour conversion as
 * in libjpeg's default pipeline.  A progressive file's refinement scans are decoded by one warp each (a 12 MP
 * file takes 0.3 - 0.7 s against 3 ms for a baseline one).  Other variants (arithmetic coding, lossless, CMYK)
 * return STITCH_E_BAD_IMAGE: the page's answer to an undecodable file (P:1508-1513); there is no CPU fallback.
 * out_pitch 0 = packed.  `info` (optional) receives the probe result.  Synchronous. */
int stitch_decode_jpeg(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                       int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info);

/* The same for a PNG file (the other format photo lists hold: screenshots).  The zlib stream of the IDAT chunks is
 * inflated on the GPU -- every bit offset is tested for a deflate block header, the blocks are decoded in parallel
 * and the matches that reach back across a block boundary are resolved by pointer jumping (png_decode.cu) -- then
 * unfiltered as a wavefront over rows and converted to RGBA8 with straight alpha: gray / RGB / palette / gray+alpha
 * / RGBA at 1-16 bit, tRNS (palette alpha or colour key) applied, 16-bit samples reduced to their high byte, gAMA /
 * iCCP ignored -- what PIL's Image.convert("RGBA") returns.  Adam7-interlaced files take the unfilter and
 * conversion kernels once per pass.  Files that inflate to
 * >= 2 GiB and damaged files return STITCH_E_BAD_IMAGE (P:1508-1513); critical-chunk CRCs are checked, IDAT CRCs and
 * the Adler-32 are not.  There is no CPU fallback.  Synchronous. */
int stitch_decode_png(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                      int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info);

/* The same for a BMP file (P:4 lists 'bmp'): header and palette are read on the host, one kernel converts the pixel
 * array -- bottom-up or top-down rows, 1 / 4 / 8 bit palettised, 16 bit (5-5-5, or the bit fields the file names),
 * 24 bit, 32 bit (alpha only where a V3+ header's mask says so; a plain 32-bit file is opaque, as in PIL and the
 * browsers).  RLE4 / RLE8 files return STITCH_E_BAD_IMAGE. */
int stitch_decode_bmp(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                      int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info);

/* The same for a GIF file (P:4 lists 'gif'): the FIRST frame, which is what createImage + drawImage paint.  The
 * container is walked on the host; the LZW stream is decoded by one warp (every dictionary string is a substring of the
 * output already written: a code is a copy), indices go through the colour table, the transparent index of a graphic
 * control extension gets alpha 0, interlaced rows are put in place.  A first frame smaller than the logical screen and a
 * damaged stream return STITCH_E_BAD_IMAGE. */
int stitch_decode_gif(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                      int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info);

/* loadImageFrom for a file of whatever type (U:27-121): looks at the signature and calls the decoder above that takes
 * it; WebP files (probed by stitch_probe_image, not decoded) and anything else return STITCH_E_BAD_IMAGE. */
int stitch_decode_image(StitchCtx* ctx, const void* file, size_t bytes, void* out_rgba, int64_t out_capacity,
                        int64_t out_pitch, int32_t out_mem_kind, StitchImageInfo* info);

/* ---- preview of the exported image on the page's <canvas> (P:1593-1603) ----
 * createHighResCanvas(node, size.width, size.height) (U:17-24) restores the canvas to floor(size * dpr)
 * device pixels under ctx.scale(dpr, dpr); clearRect leaves it transparent; the exported bitmap is drawn
 * into the centred fit rect: scaleFit = min(sw/ew, sh/eh), pv = round(e * scaleFit), at ((s - pv) / 2). */
typedef struct StitchPreviewLayout {
  int32_t canvas_w, canvas_h;  /* floor(view * dpr), U:19-20 */
  int32_t pv_w, pv_h;          /* CSS px, P:1601-1602 */
  double scale_fit;            /* P:1600 */
  double dx, dy;               /* CSS px, P:1603 */
  double dpr;                  /* pixelRatio || 1, U:2-14 */
} StitchPreviewLayout;
/* Pure CPU.  STITCH_E_BAD_DIMS for a non-positive export size, STITCH_E_BAD_OUT when the canvas has no pixels. */
int stitch_preview_layout(int32_t export_w, int32_t export_h, double view_w, double view_h, double dpr,
                          StitchPreviewLayout* out);
/* Draws the preview: `src` is the exported RGBA8 canvas (export_w x export_h, straight alpha, pitch 0 =
 * packed), `out_rgba` receives canvas_w*canvas_h*4 bytes of STRAIGHT-alpha RGBA8 (transparent where the
 * image does not reach, P:1598).  One fused compose launch on the ctx stream; synchronous. */
int stitch_preview(StitchCtx* ctx, const void* src, int32_t src_mem_kind, int32_t export_w, int32_t export_h,
                   int64_t src_pitch, double view_w, double view_h, double dpr, void* out_rgba, int64_t out_capacity,
                   int32_t out_mem_kind, StitchPreviewLayout* layout_out);

/* Pinned host / device buffers for callers without a CUDA runtime of their own (e.g. the N-API shim). */
int stitch_host_alloc(size_t bytes, void** out);
int stitch_host_free(void* p);
int stitch_device_alloc(StitchCtx* ctx, size_t bytes, void** out);
int stitch_device_free(StitchCtx* ctx, void* p);
int stitch_memcpy_h2d(StitchCtx* ctx, void* dst, const void* src, size_t bytes); /* synchronous */
int stitch_memcpy_d2h(StitchCtx* ctx, void* dst, const void* src, size_t bytes); /* synchronous */

/* On-device synthetic bitmap generator (test/bench helper K0). */
int stitch_generate_image(StitchCtx* ctx, void* dev_dst, int32_t width, int32_t height, int64_t pitch_bytes,
                          uint64_t seed, int32_t pattern);

/* 64-bit digest of `bytes` bytes of device memory (order-independent sum of mixed 32-bit words): lets a batch sweep or a
 * multi-GPU test compare canvases without copying them back.  Synchronous, lane 0. */
int stitch_device_digest(StitchCtx* ctx, const void* dev_ptr, size_t bytes, uint64_t* out);

/* The same digest for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Allocates nothing (no cudaMalloc / cudaMallocHost / cudaFree, no ctx): its scratch -- one partial sum per 256 KiB of
 * input -- is a file-scope __device__ array of 32 KB that every GPU which loads the library holds until the process
 * exits, which bounds the input at STITCH_NOALLOC_DIGEST_MAX_BYTES; a larger `bytes` is refused with STITCH_E_BAD_ARG
 * and a message in `err` (optional, NUL-terminated, at most err_capacity bytes).  Two launches on `stream` (a
 * cudaStream_t, NULL = the default stream) of GPU `device`; the 64-bit result is stored to `dev_out`, DEVICE memory of
 * that GPU passed by the caller, when the stream reaches it -- the call does not synchronise.  Same value as
 * stitch_device_digest.  All calls on one GPU share the scratch, which keeps the last call's contents (nothing in it
 * has to start at zero: a call overwrites every slot it reads): calls that OVERLAP on that GPU -- two streams, two
 * threads -- race; serialise them.  The first call per device looks the scratch's device address up
 * (cudaGetSymbolAddress): make it once outside any graph capture. */
#define STITCH_NOALLOC_DIGEST_MAX_BYTES (1ull << 30)
int stitch_device_digest_noalloc(int32_t device, void* stream, const void* dev_ptr, size_t bytes, uint64_t* dev_out,
                                 char* err, size_t err_capacity);

/* ---- every GPU of a box from one process (north star: "independent stitch jobs are sharded across the 8 GPUs of one
 * box with no NCCL") ----
 * One host thread + one StitchCtx (with `lanes_per_device` job lanes) per GPU and a shared job queue: a worker takes the
 * next job when one of its lanes is free, so uneven jobs balance themselves.  No collective, no peer
 * access.  devices == NULL or n_devices <= 0: all GPUs.  A device may be listed more than once (one worker each).
 * stitch_pool_submit copies the job descriptor and its image array; pixel buffers and *result must stay valid until the
 * job has been waited for.  `device` = index into the pool's device list of the GPU that MUST run the job (its bitmaps or
 * outputs are STITCH_MEM_DEVICE memory there), or -1: any GPU (host buffers only, else STITCH_E_BAD_ARG).
 * stitch_pool_wait returns the job's status (text: stitch_pool_last_error) and which worker ran it.  Thread-safe. */
typedef struct StitchPool StitchPool;
int stitch_pool_create(const int32_t* devices, int32_t n_devices, int32_t lanes_per_device, StitchPool** out);
int stitch_pool_destroy(StitchPool* pool); /* completes everything queued first */
int32_t stitch_pool_device_count(StitchPool* pool);
/* The worker's ctx, e.g. to allocate device buffers or generate bitmaps on that GPU BEFORE jobs are queued (a ctx is
 * not thread-safe: do not use it while its worker has jobs). */
StitchCtx* stitch_pool_ctx(StitchPool* pool, int32_t index);
int stitch_pool_submit(StitchPool* pool, const StitchJob* job, StitchResult* result, int32_t device, uint64_t* ticket);
/* stitch_device_digest on GPU `device`, ordered after every job that worker has taken so far. */
int stitch_pool_digest(StitchPool* pool, int32_t device, const void* dev_ptr, size_t bytes, uint64_t* out, uint64_t* ticket);
int stitch_pool_wait(StitchPool* pool, uint64_t ticket, int32_t* device_out /* may be NULL */);
int stitch_pool_wait_all(StitchPool* pool);
int64_t stitch_pool_jobs_run(StitchPool* pool, int32_t index);
const char* stitch_pool_last_error(StitchPool* pool);

const char* stitch_strerror(int status);
const char* stitch_last_error(StitchCtx* ctx);
int stitch_abi_version(void);
/* Number of CUDA kernels this ctx has launched since creation (for bench accounting). */
int64_t stitch_kernel_launch_count(StitchCtx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* STITCH_H_ */
