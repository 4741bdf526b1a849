// TypeScript facade of the stitch path: keeps the page's interface (ordered image list +
// 'vertical' | 'horizontal' in, exported RGBA / I420 buffer out; pages/index/index.js:14-21,1186-1633)
// and calls libstitch.so through the N-API addon in ./napi/stitch_napi.cc.
//
// NOT EXECUTED IN THIS REPO'S CI: the build image has no node/tsc (DESIGN.md section 1).  The Python
// mirror (imagestitching_b200/stitch.py) is the executed equivalent and shares the golden vectors.
import { computeLayout, fallbackProfile, Direction, ImageRecord, Layout, Platform } from './layout';

// eslint-disable-next-line @typescript-eslint/no-var-requires
const native = require('./build/Release/stitch_napi.node') as {
  createContext(device: number): unknown;
  destroyContext(ctx: unknown): void;
  run(ctx: unknown, job: NativeJob, onProgress?: (pct: number) => void): Promise<NativeResult>;
  probe(file: Uint8Array): { width: number; height: number; type: string; orientation: string; orientationExif: number };
  preview(ctx: unknown, rgba: Uint8Array, exportW: number, exportH: number, viewW: number, viewH: number, dpr: number): Uint8Array;
  decodeImage(ctx: unknown, file: Uint8Array): { width: number; height: number; data: Uint8Array };
};

export interface StitchImage extends ImageRecord {
  /** decoded RGBA8 bitmap, straight alpha (what loadImageFrom hands to drawWithOrientation) ... */
  data?: Uint8Array;
  bmpWidth?: number; bmpHeight?: number; pitchBytes?: number;
  /** ... or the FILE itself -- JPEG (baseline or progressive), PNG, BMP or GIF (first frame): decoded on the GPU (STITCH_MEM_HOST_FILE, loadImageFrom canvas.js:27-121) */
  file?: Uint8Array;
}
export interface StitchOptions {
  gap?: number; mode?: 'min' | 'max' | 'original' | string;
  platform?: Platform; deviceMaxCanvasSize?: number; deviceMaxCanvasPixels?: number;
  output?: 'rgba' | 'yuv420' | 'both' | 'jpeg' | 'rgba+jpeg' | 'png' | 'rgba+png';
  /** canvasToTempFilePath({fileType:'jpg', quality}) on the IJG 1..100 scale; the page passes quality 1 = 100 */
  jpegQuality?: number; jpegRestartInterval?: number; device?: number; onProgress?: (pct: number) => void;
}
export interface StitchResult { width: number; height: number; layout: Layout; rgba?: Uint8Array; yuv?: Uint8Array; jpeg?: Uint8Array;
  /** the PNG file canvasToTempFilePath({fileType:'png'}) would write (index.js:1577-1579) */
  png?: Uint8Array; }
interface NativeJob {
  images: StitchImage[]; direction: number; mode: number; gap: number;
  platform: number; maxCanvasSize: number; maxCanvasPixels: number; output: number;
  jpegQuality: number; jpegRestartInterval: number;
}
interface NativeResult { width: number; height: number; rgba?: Uint8Array; yuv?: Uint8Array; jpeg?: Uint8Array; png?: Uint8Array; status: number; message: string; }

const MODES: Record<string, number> = { min: 0, max: 1, original: 2 };
const PLATFORMS: Record<string, number> = { ios: 0, android: 1 };
const contexts = new Map<number, unknown>();

export class StitchError extends Error {
  constructor(public status: number, message: string) { super(message); }
}

export { computeLayout, fallbackProfile, previewLayout } from './layout';

/** wx.getImageInfo for an in-memory file (stitch_probe_image): width, height, type and WeChat's orientation string. */
export function getImageInfo(file: Uint8Array): { width: number; height: number; type: string; orientation: string; orientationExif: number } {
  return native.probe(file);
}
/** The preview draw of index.js:1593-1603 (stitch_preview): straight-alpha RGBA8 of the fit-scaled export. */
export function preview(rgba: Uint8Array, exportW: number, exportH: number, viewW: number, viewH: number, pixelRatio = 1, device = 0): Uint8Array {
  let ctx = contexts.get(device);
  if (!ctx) { ctx = native.createContext(device); contexts.set(device, ctx); }
  return native.preview(ctx, rgba, exportW, exportH, viewW, viewH, pixelRatio);
}

/** loadImageFrom (canvas.js:27-121) for an in-memory file -- JPEG (baseline or progressive), PNG, BMP, GIF (first frame):
 *  decoded on the GPU (stitch_decode_image) into straight-alpha RGBA8; rejects like the page's 图片解码异常. */
export function decodeImage(file: Uint8Array, device = 0): { width: number; height: number; data: Uint8Array } {
  let ctx = contexts.get(device);
  if (!ctx) { ctx = native.createContext(device); contexts.set(device, ctx); }
  return native.decodeImage(ctx, file);
}

/** onStitch(): resolves to undefined for an empty list (index.js:1189), rejects with the reference's messages. */
export async function stitch(images: StitchImage[], direction: Direction, opts: StitchOptions = {}): Promise<StitchResult | undefined> {
  if (!images.length) return undefined;
  const platform = opts.platform || 'ios';
  const profile = fallbackProfile(platform, opts.deviceMaxCanvasSize, opts.deviceMaxCanvasPixels);
  const mode = opts.mode || 'min';
  const layout = computeLayout(images, direction, mode, opts.gap, profile);      // twin; the native side recomputes it
  const device = opts.device || 0;
  let ctx = contexts.get(device);
  if (!ctx) { ctx = native.createContext(device); contexts.set(device, ctx); }
  const res = await native.run(ctx, {
    images, direction: direction === 'vertical' ? 0 : 1, mode: mode in MODES ? MODES[mode] : 3, gap: opts.gap || 0,
    platform: platform in PLATFORMS ? PLATFORMS[platform] : 2,
    maxCanvasSize: profile.deviceMaxCanvasSize || 0, maxCanvasPixels: profile.deviceMaxCanvasPixels || 0,
    output: { rgba: 1, yuv420: 2, both: 3, jpeg: 4, 'rgba+jpeg': 5, png: 8, 'rgba+png': 9 }[opts.output || 'rgba'],
    jpegQuality: opts.jpegQuality || 0, jpegRestartInterval: opts.jpegRestartInterval || 0,
  }, opts.onProgress);
  if (res.status !== 0) throw new StitchError(res.status, res.message);
  if (res.width !== layout.canvasW || res.height !== layout.canvasH) throw new StitchError(-1, 'layout twin mismatch');
  return { width: res.width, height: res.height, layout, rgba: res.rgba, yuv: res.yuv, jpeg: res.jpeg, png: res.png };
}
