// CPU drive of imagestitching_b200/csrc/jpeg_prog_core.h: the sequential scan decoder a GPU thread runs for the
// refinement scans of a progressive JPEG, and the scan-order -> buffer-order block map, run scan after scan on the host.
// TEST CODE ONLY: tests/test_jpeg_progressive.py holds its coefficients against the oracle decoder's; nothing in the
// library links it.
#include <cstdint>
#include <cstring>
#include <vector>

#include "jpeg_prog_core.h"

using namespace stitch::jpegdec;

// info: [0] w, [1] h, [2] ncomp, [3] hs, [4] vs, [5] scans, [6] blocks in the buffer
extern "C" int jprog_host_info(const uint8_t* file, size_t n, int32_t* info) {
  Header H;
  const ParseStatus ps = parse(file, n, &H);
  if (ps != kOk) return -(int)ps;
  if (!H.progressive) return -10;
  info[0] = H.w; info[1] = H.h; info[2] = H.ncomp; info[3] = H.hs; info[4] = H.vs;
  info[5] = (int32_t)H.scans.size();
  info[6] = H.mcux * H.mcuy * (H.ncomp == 1 ? 1 : H.hs * H.vs + 2);
  return 0;
}

// coef: [blocks][64] int16, buffer order (interleaved MCU order, zig-zag inside a block), zeroed here.
// scan_desc (optional): per scan {mode, ns, comp, ss, se, ah, al, n_blk}.
static int g_by_mask = 0;   // 0: coefficient by coefficient, 1: pj_refine_parse, 2: pj_refine_parse_z

// AC / DC refinement scans the way the GPU's warp kernel runs them (jprog_seq_kernel): block by block through
// pj_refine_parse (one thread) + pj_refine_apply (every coefficient) with the block's non-zero mask.
static void refine_by_mask(const PScan& S, const SmemTables& T) {
  SeqBits b;
  std::memset(&b, 0, sizeof b);
  b.p = S.data;
  b.n = S.len;
  int rst_left = S.ri, rst_next = 0;
  uint32_t eobrun = 0;
  const int p1 = 1 << S.al;
  int b_in_mcu = 0;
  for (int blk = 0; blk < S.n_blk; ++blk) {
    if (b_in_mcu == 0 && S.ri) {
      if (rst_left == 0) { br_restart(b, &rst_next); eobrun = 0; rst_left = S.ri; }
      rst_left--;
    }
    int16_t* c = S.coef + pscan_block(S, blk) * 64;
    if (S.mode == PSCAN_DC_REFINE) {
      if (br_bits(b, 1)) c[0] = (int16_t)(c[0] | p1);
    } else {
      uint64_t nz = 0;
      for (int k = 0; k < 64; ++k) nz |= (uint64_t)(c[k] != 0) << k;
      const uint64_t band = (S.se >= 63 ? ~0ull : ((1ull << (S.se + 1)) - 1)) & ~((1ull << S.ss) - 1);
      RefineOut o;
      if (g_by_mask == 2) {
        uint8_t zpos[64];
        int nz0 = 0;
        for (int k = S.ss; k <= S.se; ++k)
          if (!((nz >> k) & 1)) zpos[nz0++] = (uint8_t)k;
        pj_refine_parse_z(b, T, 2 + S.ta, nz, band, zpos, nz0, S.ss, S.se, &eobrun, &o);
      } else {
        pj_refine_parse(b, T, 2 + S.ta, nz, band, &eobrun, &o);
      }
      for (int k = 0; k < 64; ++k) c[k] = (int16_t)pj_refine_apply(c[k], k, p1, o);      // on the GPU: two per lane
    }
    b_in_mcu = (b_in_mcu + 1 == S.bpm) ? 0 : b_in_mcu + 1;
  }
}

extern "C" void jprog_host_set_mask_path(int on) { g_by_mask = on; }

extern "C" int jprog_host_decode(const uint8_t* file, size_t n, int16_t* coef, int32_t* scan_desc) {
  Header H;
  const ParseStatus st = parse(file, n, &H);
  if (st != kOk) return -(int)st;
  if (!H.progressive) return -10;
  const size_t blocks = (size_t)H.mcux * H.mcuy * (H.ncomp == 1 ? 1 : H.hs * H.vs + 2);
  std::memset(coef, 0, blocks * 128);
  std::vector<uint8_t> data;
  for (size_t i = 0; i < H.scans.size(); ++i) {
    const ProgScan& ps = H.scans[i];
    PScan S;
    std::memset(&S, 0, sizeof S);
    pscan_describe(H, ps, &S);
    data.assign(file + ps.off, file + ps.off + ps.len);
    data.resize(ps.len + 16, 0);
    S.data = data.data();
    S.tab = &H.scan_tabs[ps.tab];
    S.coef = coef;
    SmemTables T;
    build_tables_seq(S.tab, T);
    if (g_by_mask && (S.mode == PSCAN_AC_REFINE || S.mode == PSCAN_DC_REFINE)) refine_by_mask(S, T);
    else pscan_sequential(S, T);
    if (scan_desc) {
      int32_t* d = scan_desc + 8 * i;
      d[0] = S.mode; d[1] = S.ns; d[2] = S.comp; d[3] = S.ss; d[4] = S.se; d[5] = ps.ah; d[6] = S.al; d[7] = S.n_blk;
    }
  }
  return 0;
}

// walk_index (the division-free block map of the warp kernel) against pscan_block, every block of every scan.
extern "C" int jprog_host_walk_check(const uint8_t* file, size_t n) {
  Header H;
  if (parse(file, n, &H) != kOk || !H.progressive) return -1;
  for (const ProgScan& ps : H.scans) {
    PScan S;
    std::memset(&S, 0, sizeof S);
    pscan_describe(H, ps, &S);
    BlockWalk w;
    walk_start(S, 0, &w);
    for (int blk = 0; blk < S.n_blk; ++blk) {
      if (walk_index(S, w, blk) != pscan_block(S, blk)) return blk + 1;
      walk_next(S, &w);
    }
  }
  return 0;
}
