// Host-side JPEG header parser of the GPU decoder (jpeg_decode.cuh): marker segments of ITU-T T.81 Annex B.
//
// Replaces the first half of loadImageFrom (U:27-121: canvas.createImage() + img.src = path); the entropy
// decoding, IDCT, upsampling and colour conversion run on the GPU.  Supported: SOF0 / SOF1 (Huffman, 8 bit) with
// 1 or 3 components in one interleaved scan, and SOF2 (progressive Huffman: any script of interleaved DC scans and
// single-component scans, tables redefined between scans); luma sampling 1x1 / 2x1 / 2x2 with 1x1 chroma, DRI.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

namespace stitch {
namespace jpegdec {

// zig-zag position k -> natural index (T.81 Figure A.6)
static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// Everything the kernels need from the headers; uploaded once per image.
struct Tables {
  uint16_t q[4][64];      // quantisation tables in ZIG-ZAG order (as DQT stores them)
  uint8_t bits[4][16];    // Huffman BITS, tables [class * 2 + id]
  uint8_t vals[4][256];   // HUFFVAL
};

// Huffman tables as the kernels hold them in shared memory (built from Tables by build_tables / build_tables_seq).
struct SmemTables {
  uint16_t lut[4][256];       // first 8 bits -> (length << 8 | symbol), 0 = longer than 8 bits
  int32_t maxcode[4][17];     // per length 1..16, -1 = no codes
  int32_t mincode[4][17];
  int32_t valptr[4][17];
  uint8_t vals[4][256];
};

// One scan of a progressive file (T.81 G.1): components, spectral band, successive-approximation bits, and where
// its entropy-coded bytes lie.  `tab` indexes Header::scan_tabs (the Huffman tables in force when the scan starts).
struct ProgScan {
  int ns = 0;                  // components in the scan: 1, or all of the frame's in frame order
  int comp = 0;                // frame index of the (first) component
  int td[3] = {0, 0, 0}, ta = 0;
  int ss = 0, se = 0, ah = 0, al = 0;
  int ri = 0;                  // restart interval in force, in MCUs of this scan
  int tab = 0;
  size_t off = 0, len = 0;
};

struct Header {
  bool progressive = false;    // SOF2: scans[] lists the file's scans, scan_off / scan_len are unused
  std::vector<ProgScan> scans;
  std::vector<Tables> scan_tabs;
  int w = 0, h = 0, ncomp = 0;
  int hs = 1, vs = 1;          // luma sampling factors (chroma is 1x1)
  int tq[3] = {0, 0, 0}, td[3] = {0, 0, 0}, ta[3] = {0, 0, 0};
  int ri = 0;                  // restart interval in MCUs (0 = none)
  int mcux = 0, mcuy = 0;
  size_t scan_off = 0, scan_len = 0;
  Tables tab;
};

enum ParseStatus { kOk = 0, kNotJpeg = 1, kCorrupt = 2, kUnsupported = 3 };

// End of the entropy-coded data that starts at o: the first marker that is neither a stuffed zero nor RSTn.
inline size_t scan_end(const uint8_t* p, size_t n, size_t o) {
  while (o + 1 < n) {
    const uint8_t* ff = (const uint8_t*)std::memchr(p + o, 0xFF, n - 1 - o);   // entropy-coded data holds few 0xFF bytes
    if (!ff) return n;
    o = (size_t)(ff - p);
    if (p[o + 1] != 0 && p[o + 1] != 0xFF && !(p[o + 1] >= 0xD0 && p[o + 1] <= 0xD7)) return o;
    ++o;
  }
  return n;
}

inline ParseStatus parse(const uint8_t* p, size_t n, Header* H) {
  std::memset(&H->tab, 0, sizeof(H->tab));
  H->progressive = false;
  H->scans.clear();
  H->scan_tabs.clear();
  bool tabs_dirty = true;       // a DHT since the last scan: the next scan gets a fresh snapshot
  if (n < 4 || p[0] != 0xFF || p[1] != 0xD8) return kNotJpeg;
  bool have_sof = false, have_h[4] = {false, false, false, false}, have_q[4] = {false, false, false, false};
  int comp_id[3] = {0, 0, 0};   // SOF component identifiers: the scan must name them in the same order
  int adobe_transform = -1;     // Adobe APP14 colour transform: 0 = the three components are RGB, not YCbCr
  bool saw_jfif = false;        // a JFIF APP0 marker settles it: YCbCr
  size_t o = 2;
  while (o + 4 <= n) {
    if (p[o] != 0xFF) return kCorrupt;
    const int m = p[o + 1];
    if (m == 0xFF) { ++o; continue; }
    if (m == 0xD9 && H->progressive && !H->scans.empty()) return kOk;   // EOI
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { o += 2; continue; }
    const size_t len = ((size_t)p[o + 2] << 8) | p[o + 3];
    if (len < 2 || o + 2 + len > n) return (H->progressive && !H->scans.empty()) ? kOk : kCorrupt;
    const uint8_t* s = p + o + 4;
    const size_t sl = len - 2;
    if (m == 0xDB) {
      if (H->progressive && !H->scans.empty()) return kUnsupported;   // quantisation tables changed between scans
      for (size_t i = 0; i < sl;) {
        const int pq = s[i] >> 4, t = s[i] & 15;
        if (pq != 0) return kUnsupported;  // 16-bit tables belong to 12-bit JPEG
        if (t > 3 || i + 65 > sl) return kCorrupt;
        for (int k = 0; k < 64; ++k) H->tab.q[t][k] = s[i + 1 + k];
        have_q[t] = true;
        i += 65;
      }
    } else if (m == 0xE0) {
      if (sl >= 5 && std::memcmp(s, "JFIF", 5) == 0) saw_jfif = true;
    } else if (m == 0xEE) {
      if (sl >= 12 && std::memcmp(s, "Adobe", 5) == 0) adobe_transform = s[11];
    } else if (m == 0xC0 || m == 0xC1 || m == 0xC2) {
      if (have_sof) return kCorrupt;
      H->progressive = (m == 0xC2);
      if (sl < 6) return kCorrupt;
      if (s[0] != 8) return kUnsupported;
      H->h = (s[1] << 8) | s[2];
      H->w = (s[3] << 8) | s[4];
      H->ncomp = s[5];
      if (H->w == 0 || H->h == 0) return kCorrupt;
      if (H->ncomp != 1 && H->ncomp != 3) return kUnsupported;
      if (sl < (size_t)(6 + 3 * H->ncomp)) return kCorrupt;
      int hs[3] = {1, 1, 1}, vs[3] = {1, 1, 1};
      for (int c = 0; c < H->ncomp; ++c) {
        comp_id[c] = s[6 + 3 * c];
        hs[c] = s[7 + 3 * c] >> 4;
        vs[c] = s[7 + 3 * c] & 15;
        H->tq[c] = s[8 + 3 * c];
        if (H->tq[c] > 3) return kCorrupt;
      }
      if (H->ncomp == 1) {
        H->hs = H->vs = 1;
      } else {
        if (hs[1] != 1 || vs[1] != 1 || hs[2] != 1 || vs[2] != 1) return kUnsupported;
        if (!((hs[0] == 1 && vs[0] == 1) || (hs[0] == 2 && vs[0] == 1) || (hs[0] == 2 && vs[0] == 2))) return kUnsupported;
        H->hs = hs[0];
        H->vs = vs[0];
      }
      H->mcux = (H->w + 8 * H->hs - 1) / (8 * H->hs);
      H->mcuy = (H->h + 8 * H->vs - 1) / (8 * H->vs);
      have_sof = true;
    } else if (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
      return kUnsupported;  // lossless, hierarchical, arithmetic coding
    } else if (m == 0xC4) {
      tabs_dirty = true;
      for (size_t i = 0; i < sl;) {
        if (i + 17 > sl) return kCorrupt;
        const int tc = s[i] >> 4, th = s[i] & 15;
        if (tc > 1 || th > 1) return kUnsupported;  // baseline: two tables per class
        const int t = tc * 2 + th;
        int cnt = 0;
        for (int l = 0; l < 16; ++l) { H->tab.bits[t][l] = s[i + 1 + l]; cnt += s[i + 1 + l]; }
        if (cnt > 256 || i + 17 + cnt > sl) return kCorrupt;
        std::memset(H->tab.vals[t], 0, 256);
        std::memcpy(H->tab.vals[t], s + i + 17, cnt);
        have_h[t] = true;
        i += 17 + cnt;
      }
    } else if (m == 0xDD) {
      if (sl < 2) return kCorrupt;
      H->ri = (s[0] << 8) | s[1];
    } else if (m == 0xDA && have_sof && H->progressive) {
      // a scan of a progressive file (T.81 B.2.3, G.1.1): recorded, its data stepped over
      if (sl < 1) return kCorrupt;
      ProgScan S;
      S.ns = s[0];
      if (S.ns < 1 || S.ns > 4 || sl < (size_t)(1 + 2 * S.ns + 3)) return kCorrupt;
      if (S.ns != 1 && S.ns != H->ncomp) return kUnsupported;            // partial interleaves
      if (H->scans.empty() && H->ncomp == 3 && !saw_jfif &&
          (adobe_transform == 0 || (adobe_transform < 0 && comp_id[0] == 'R' && comp_id[1] == 'G' && comp_id[2] == 'B')))
        return kUnsupported;                                            // RGB-in-JPEG, as for baseline files
      for (int j = 0; j < S.ns; ++j) {
        int c = -1;
        for (int k = 0; k < H->ncomp; ++k)
          if (comp_id[k] == s[1 + 2 * j]) c = k;
        if (c < 0) return kCorrupt;
        if (S.ns > 1 && c != j) return kUnsupported;                     // scan components in another order than the frame's
        if (j == 0) S.comp = c;
        S.td[j] = s[2 + 2 * j] >> 4;
        if (j == 0) S.ta = s[2 + 2 * j] & 15;
        if (S.td[j] > 1 || (s[2 + 2 * j] & 15) > 1) return kUnsupported;
        if (!have_q[H->tq[c]]) return kCorrupt;
      }
      S.ss = s[1 + 2 * S.ns];
      S.se = s[2 + 2 * S.ns];
      S.ah = s[3 + 2 * S.ns] >> 4;
      S.al = s[3 + 2 * S.ns] & 15;
      if (S.al > 13 || S.se > 63 || S.ss > S.se) return kCorrupt;
      if (S.ss == 0 && S.se != 0) return kCorrupt;                       // DC and AC never share a scan (G.1.1.1.1)
      if (S.ss > 0 && S.ns != 1) return kCorrupt;                        // AC scans code one component
      if (S.ss == 0 && S.ah == 0)
        for (int j = 0; j < S.ns; ++j)
          if (!have_h[S.td[j]]) return kCorrupt;
      if (S.ss > 0 && !have_h[2 + S.ta]) return kCorrupt;
      S.ri = H->ri;
      if (tabs_dirty || H->scan_tabs.empty()) { H->scan_tabs.push_back(H->tab); tabs_dirty = false; }
      S.tab = (int)H->scan_tabs.size() - 1;
      S.off = o + 2 + len;
      const size_t end = scan_end(p, n, S.off);
      S.len = end - S.off;
      if (S.len >= ((size_t)1 << 31)) return kUnsupported;
      H->scans.push_back(S);
      if (H->scans.size() > 256) return kUnsupported;
      o = end;
      continue;
    } else if (m == 0xDA) {
      if (!have_sof) return kCorrupt;
      if (sl < 1 || s[0] != H->ncomp) return kUnsupported;  // non-interleaved scans
      if (sl < (size_t)(1 + 2 * H->ncomp + 3)) return kCorrupt;
      // The decoder converts three components as YCbCr.  libjpeg (jdapimin.c default_decompress_parms) takes them as
      // RGB when there is no JFIF marker and an Adobe marker says transform 0, or -- with neither marker -- when the
      // component ids are 'R','G','B': such files are refused here rather than decoded to wrong colours.
      if (H->ncomp == 3 && !saw_jfif &&
          (adobe_transform == 0 || (adobe_transform < 0 && comp_id[0] == 'R' && comp_id[1] == 'G' && comp_id[2] == 'B')))
        return kUnsupported;
      for (int c = 0; c < H->ncomp; ++c) {
        if (s[1 + 2 * c] != comp_id[c]) return kUnsupported;   // scan components in another order than the frame's
        H->td[c] = s[2 + 2 * c] >> 4;
        H->ta[c] = s[2 + 2 * c] & 15;
        if (H->td[c] > 1 || H->ta[c] > 1) return kUnsupported;
        if (!have_h[H->td[c]] || !have_h[2 + H->ta[c]] || !have_q[H->tq[c]]) return kCorrupt;
      }
      H->scan_off = o + 2 + len;
      H->scan_len = n - H->scan_off;
      return kOk;
    }
    o += 2 + len;
  }
  return (H->progressive && !H->scans.empty()) ? kOk : kCorrupt;   // a progressive file cut short keeps the scans it has
}

}  // namespace jpegdec
}  // namespace stitch
