// Exact TypeScript twin of the reference's layout arithmetic (pages/index/index.js:1211-1213,
// 1243-1386, 1432-1433, 1522-1554).  JS numbers ARE the reference's arithmetic, so this file keeps the
// same operand order and Math.* calls; it must agree bit-for-bit with stitch_layout() in libstitch.so
// and with tests/golden/layout_golden.json.
//
// NOT EXECUTED IN THIS REPO'S CI: the build image has no node/tsc (DESIGN.md section 1).

export type Direction = 'vertical' | 'horizontal';
export type Platform = 'ios' | 'android' | string;

export interface ImageRecord {
  naturalWidth?: number; naturalHeight?: number; width?: number; height?: number;
  orientation?: number | string; fileSize?: number;
}
export interface Profile { platform: Platform; deviceMaxCanvasSize?: number; deviceMaxCanvasPixels?: number; }
export interface Rect { dx: number; dy: number; dw: number; dh: number; }
export interface Layout {
  n: number; bigTask: boolean; maxSide: number; maxPixels: number; preW: number; preH: number;
  scaleDown: number; targetW: number; targetH: number; superSample: number; canvasW: number; canvasH: number;
  scaledGap: number; rects: Rect[]; natural: [number, number][]; progress: number[];
}

/** index.js:128-138 applyFallback() */
export function fallbackProfile(platform: Platform, size?: number, pixels?: number): Profile {
  const s = size || (platform === 'ios' ? 12288 : 4096);
  let p = pixels;
  if (!p) {
    const cap = platform === 'android' ? s * Math.min(s, 4096) : s * Math.min(s, 12288);
    p = Math.max(cap, 4096 * 2048);
  }
  return { platform, deviceMaxCanvasSize: s, deviceMaxCanvasPixels: p };
}

/** stitchProgress values of a successful run (index.js:1193,1247,1358,1556-1557,1581,1611) */
export function progressSequence(n: number): number[] {
  const seq = [1];
  for (let done = 1; done <= n; done++) seq.push(Math.min(25, Math.round(1 + (done / n) * 24)));
  seq.push(30);
  for (let idx = 0; idx < n; idx++) seq.push(Math.min(90, 30 + Math.round(((idx + 1) / n) * 60)));
  seq.push(96, 100);
  return seq;
}

export function computeLayout(images: ImageRecord[], direction: Direction, modeIn: string | undefined,
                              gap: number | undefined, profile: Profile): Layout {
  if (!images.length) throw new Error('empty');                                   // :1189 (callers no-op)
  const totalBytes = images.reduce((s, it) => s + (it && it.fileSize ? it.fileSize : 0), 0);   // :1211
  const bigTask = images.length >= 7 || totalBytes >= 25 * 1024 * 1024;           // :1212
  const nat = images.map(i => [Math.max(1, i.naturalWidth || i.width || 1),       // :1243-1244
                               Math.max(1, i.naturalHeight || i.height || 1)] as [number, number]);
  const allW = nat.map(x => x[0]).filter(n => n > 0);
  const allH = nat.map(x => x[1]).filter(n => n > 0);
  if (!allW.length || !allH.length) throw new Error('图片尺寸不可用');              // :1254
  const gapPx = gap || 0;                                                         // :1256
  const mode = modeIn || 'min';                                                   // :1257
  let outW = 0, outH = 0;
  if (direction === 'vertical') {                                                 // :1261-1287
    if (mode === 'min' || mode === 'max') {
      outW = mode === 'min' ? Math.min.apply(null, allW) : Math.max.apply(null, allW);
      outH = nat.reduce((sum, [iw, ih], idx) => sum + ih * (outW / iw) + (idx ? gapPx : 0), 0);
    } else {
      outW = Math.max.apply(null, allW);
      outH = nat.reduce((sum, [, ih], idx) => sum + ih + (idx ? gapPx : 0), 0);
    }
  } else {                                                                        // :1288-1315
    if (mode === 'min' || mode === 'max') {
      outH = mode === 'min' ? Math.min.apply(null, allH) : Math.max.apply(null, allH);
      outW = nat.reduce((sum, [iw, ih], idx) => sum + iw * (outH / ih) + (idx ? gapPx : 0), 0);
    } else {
      outH = Math.max.apply(null, allH);
      outW = nat.reduce((sum, [iw], idx) => sum + iw + (idx ? gapPx : 0), 0);
    }
  }
  outW = Math.max(1, Math.round(outW));                                           // :1318
  outH = Math.max(1, Math.round(outH));                                           // :1319
  if (!outW || !outH) throw new Error('输出尺寸计算失败');                          // :1320
  const preW = outW, preH = outH;

  const platform = profile.platform;                                              // :1324-1336
  const SIDE = profile.deviceMaxCanvasSize || (platform === 'android' ? 4096 : 12288);
  let PIX: number;
  if (platform === 'ios') {
    const iosLimit = 16384 * 1400;
    PIX = Math.min(profile.deviceMaxCanvasPixels || iosLimit, iosLimit);
  } else if (platform === 'android') {
    const lim = SIDE * Math.min(SIDE, 8192);
    PIX = Math.min(profile.deviceMaxCanvasPixels || lim, lim);
  } else {
    const lim = SIDE * SIDE;
    PIX = Math.min(profile.deviceMaxCanvasPixels || lim, lim);
  }
  let scaleDown = 1;                                                              // :1337-1357
  if (outW > SIDE || outH > SIDE) scaleDown = Math.min(SIDE / outW, SIDE / outH);
  const totalPixels = outW * outH;
  if (totalPixels > PIX) scaleDown = Math.min(scaleDown, Math.sqrt(PIX / totalPixels));
  if (scaleDown < 1) {
    outW = Math.max(1, Math.floor(outW * scaleDown));
    outH = Math.max(1, Math.floor(outH * scaleDown));
  }
  const targetW = outW, targetH = outH;                                           // :1360-1386
  const basePixels = targetW * targetH;
  const MAX_SS = bigTask ? 1 : (platform === 'ios' ? 2.2 : 2.6);
  let ss = 1;
  if (basePixels > 0 && basePixels < PIX) {
    const ratio = Math.sqrt(PIX / basePixels);
    if (ratio > 1.01) ss = Math.min(MAX_SS, ratio, Math.min(SIDE / targetW, SIDE / targetH));
  }
  if (!Number.isFinite(ss) || ss < 1) ss = 1;
  let cw = Math.max(1, Math.round(targetW * ss));
  let ch = Math.max(1, Math.round(targetH * ss));
  let guard = 0;
  while (cw * ch > PIX && guard < 20) {
    ss *= 0.96;
    if (ss < 1) { ss = 1; break; }          // keeps cw/ch of the previous iteration (reference quirk)
    cw = Math.max(1, Math.floor(targetW * ss));
    ch = Math.max(1, Math.floor(targetH * ss));
    guard++;
  }
  const scaledGap = gapPx * scaleDown;                                            // :1432
  let cursorX = 0, cursorY = 0;
  const rects: Rect[] = [];
  for (const [nw, nh] of nat) {                                                   // :1522-1554
    if (direction === 'vertical') {
      if (mode === 'original') {
        const dw = Math.round(nw * scaleDown), dh = Math.round(nh * scaleDown);
        rects.push({ dx: Math.floor((outW - dw) / 2), dy: cursorY, dw, dh });
        cursorY += dh + scaledGap;
      } else {
        const dh = Math.round(nh * (outW / nw));
        rects.push({ dx: 0, dy: cursorY, dw: outW, dh });
        cursorY += dh + scaledGap;
      }
    } else {
      if (mode === 'original') {
        const dw = Math.round(nw * scaleDown), dh = Math.round(nh * scaleDown);
        rects.push({ dx: cursorX, dy: Math.floor((outH - dh) / 2), dw, dh });
        cursorX += dw + scaledGap;
      } else {
        const dw = Math.round(nw * (outH / nh));
        rects.push({ dx: cursorX, dy: 0, dw, dh: outH });
        cursorX += dw + scaledGap;
      }
    }
  }
  return { n: images.length, bigTask, maxSide: SIDE, maxPixels: PIX, preW, preH, scaleDown, targetW, targetH,
           superSample: ss, canvasW: cw, canvasH: ch, scaledGap, rects, natural: nat,
           progress: progressSequence(images.length) };
}

/** Preview geometry: createHighResCanvas (utils/canvas.js:17-24) + the fit-scale of index.js:1599-1603.
 *  Twin of stitch_preview_layout in libstitch.so. */
export interface PreviewLayout {
  canvasW: number; canvasH: number; scaleFit: number; pvW: number; pvH: number; dx: number; dy: number; dpr: number;
}
export function previewLayout(exportW: number, exportH: number, viewW: number, viewH: number, pixelRatio?: number): PreviewLayout {
  const dpr = pixelRatio || 1;                                                    // canvas.js:6,11
  const scaleFit = Math.min(viewW / exportW, viewH / exportH);                    // :1600
  const pvW = Math.round(exportW * scaleFit), pvH = Math.round(exportH * scaleFit); // :1601-1602
  return { canvasW: Math.floor(viewW * dpr), canvasH: Math.floor(viewH * dpr), scaleFit, pvW, pvH,
           dx: (viewW - pvW) / 2, dy: (viewH - pvH) / 2, dpr };                     // canvas.js:19-20, :1603
}
