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
    pscan_sequential(S, T);
    if (scan_desc) {
      int32_t* d = scan_desc + 8 * i;
      d[0] = S.mode; d[1] = S.ns; d[2] = S.comp; d[3] = S.ss; d[4] = S.se; d[5] = ps.ah; d[6] = S.al; d[7] = S.n_blk;
    }
  }
  return 0;
}
