// CPU drive of imagestitching_b200/csrc/png_inflate_core.h: the same inline functions the GPU kernels of
// png_decode.cu call, run one "thread" after another in the kernels' order (find -> validate -> count -> walk ->
// write -> expand -> resolve -> bytes).  TEST CODE ONLY: tests/test_png_decode.py holds its output against zlib; nothing in
// the library links it.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>

#include "png_inflate_core.h"

using namespace stitch::pnginf;

namespace {
struct Thread {
  std::vector<uint16_t> lit, dist, slow;
  HuffTables T;
  uint8_t lens[320];
  Thread() : lit(kLitEntries), dist(kDistEntries), slow(kSlowWords) {
    T.lit = lit.data();
    T.dist = dist.data();
    T.stride = 1;
    bind_slow(T, slow.data());
  }
};
}  // namespace

// stats: [0] prefilter survivors, [1] candidates, [2] chain blocks, [3] blocks the walker had to decode itself,
//        [4] resolve rounds, [5] bytes that are copies
extern "C" int pinf_host_inflate(const uint8_t* z_in, uint32_t z_len, uint32_t raw_len, uint8_t* out, uint32_t* stats) {
  std::vector<uint8_t> zbuf((size_t)z_len + 32, 0);
  std::memcpy(zbuf.data(), z_in, z_len);
  const uint8_t* z = zbuf.data();
  const uint32_t z_bits = z_len * 8u;
  Thread th;
  std::memset(stats, 0, 6 * sizeof(uint32_t));
  // find + validate
  std::vector<uint32_t> surv;
  for (uint32_t p = 16; p + 32 <= z_bits; ++p)
    if (header_prefilter(bits_at(z, p), bits_at(z, p + 64))) surv.push_back(p);
  stats[0] = (uint32_t)surv.size();
  std::map<uint32_t, BlockRec> cand;
  for (uint32_t p : surv) {
    BitReader br;
    br.init(z, p);
    br.take(3);
    int nlen, ndist;
    if (!read_dynamic_lengths(br, th.lens, &nlen, &ndist, z_bits)) continue;
    // count
    BlockRec rec;
    if (decode_block<false>(z, z_len, p, th.T, th.lens, &rec, 0, raw_len, nullptr, nullptr) != kBlockOk) continue;
    cand[p] = rec;
  }
  stats[1] = (uint32_t)cand.size();
  // walk
  struct Real { uint32_t bitpos, out, tok; };
  std::vector<Real> chain;
  uint32_t pos = 16, outp = 0, tokp = 0;
  for (;;) {
    BlockRec rec;
    auto it = cand.find(pos);
    if (it != cand.end()) {
      rec = it->second;
    } else {
      if (decode_block<false>(z, z_len, pos, th.T, th.lens, &rec, 0, raw_len - outp, nullptr, nullptr) != kBlockOk) return 1;
      stats[3]++;
    }
    if (rec.n_out > raw_len - outp) return 2;
    chain.push_back(Real{pos, outp, tokp});
    outp += rec.n_out;
    tokp += rec.n_tok;
    pos = rec.end_bit;
    if (rec.flags & kRecFinal) break;
  }
  stats[2] = (uint32_t)chain.size();
  if (outp != raw_len) return 3;
  // write: literals + tokens
  std::vector<uint32_t> src(raw_len, kRoot);
  std::vector<Token> tok(tokp + 1);
  for (const Real& b : chain)
    if (decode_block<true>(z, z_len, b.bitpos, th.T, th.lens, nullptr, b.out, raw_len, out, tok.data() + b.tok) != kBlockOk) return 4;
  // expand
  for (uint32_t t = 0; t < tokp; ++t) {
    const uint32_t len = tok[t].len_dist >> 15, dist = (tok[t].len_dist & 0x7FFFu) + 1u;
    for (uint32_t k = 0; k < len; ++k) src[tok[t].pos + k] = tok[t].pos - dist + (dist >= len ? k : k % dist);   // as pinf_expand_kernel
  }
  for (uint32_t i = 0; i < raw_len; ++i) stats[5] += src[i] != kRoot;
  // resolve
  for (;;) {
    bool changed = false;
    for (uint32_t i = 0; i < raw_len; ++i) {
      const uint32_t s = src[i];
      if (s == kRoot) continue;
      const uint32_t s2 = src[s];
      if (s2 != kRoot) { src[i] = s2; changed = true; }
    }
    if (!changed) break;
    if (++stats[4] > 64) return 5;
  }
  for (uint32_t i = 0; i < raw_len; ++i)
    if (src[i] != kRoot) out[i] = out[src[i]];
  return 0;
}

// Self-synchronisation statistics of the first Huffman block of a stream: subsequences of `sub_bits` bits are decoded
// from their nominal (guessed) starts; out[0] = subsequences tried, out[1] = those whose end is a true symbol boundary.
extern "C" int pinf_host_sync_stats(const uint8_t* z_in, uint32_t z_len, uint32_t sub_bits, uint32_t* out) {
  std::vector<uint8_t> zbuf((size_t)z_len + 32, 0);
  std::memcpy(zbuf.data(), z_in, z_len);
  const uint8_t* z = zbuf.data();
  Thread th;
  DecState st;
  block_begin<false>(st, z, z_len, 16, th.T, th.lens, 0, 0x7FFFFFFFu, nullptr);
  out[0] = out[1] = 0;
  if (st.status != kRunning) return 1;
  const uint32_t p0 = st.br.pos();
  std::vector<uint8_t> boundary((size_t)z_len * 8 + 64, 0);
  while (st.status == kRunning) {
    boundary[st.br.pos()] = 1;
    symbol_step<false>(st, th.T, nullptr, nullptr);
  }
  if (st.status != kDone) return 2;
  const uint32_t end = st.end_bit;
  for (uint32_t s = p0 + sub_bits; s + sub_bits < end; s += sub_bits) {
    DecState sp;
    sp.z_bits = z_len * 8u;
    sp.pos = sp.out_start = 0;
    sp.out_limit = 0x7FFFFFFFu;
    sp.n_tok = 0;
    sp.status = kRunning;
    sp.rep_nbits = 0;
    sp.br.init(z, s);
    while (sp.status == kRunning && sp.br.pos() < s + sub_bits) symbol_step<false>(sp, th.T, nullptr, nullptr);
    out[0]++;
    if (sp.status == kRunning && boundary[sp.br.pos()]) out[1]++;
  }
  return 0;
}

// header_prefilter_tab (what the finder kernel runs) against header_prefilter on n pseudo-random 128-bit windows and on
// every bit offset of a stream; returns the number of disagreements, *hits = windows that pass.
extern "C" uint32_t pinf_host_prefilter_check(const uint8_t* z_in, uint32_t z_len, uint32_t n_random, uint32_t* hits) {
  uint8_t tab[512];
  for (int e = 0; e < 512; ++e) tab[e] = kraft3_entry(e);
  uint32_t bad = 0, h = 0;
  uint64_t x = 0x9E3779B97F4A7C15ull;
  auto next = [&]() { x ^= x << 13; x ^= x >> 7; x ^= x << 17; return x; };
  for (uint32_t i = 0; i < n_random; ++i) {
    uint64_t b = next(), b2 = next();
    if (i & 1) b = (b & ~6ull) | 4ull;   // half of them with block type 2, so the interesting path is taken often
    const bool a = header_prefilter(b, b2), t = header_prefilter_tab(b, b2, tab);
    bad += a != t;
    h += a;
  }
  std::vector<uint8_t> zbuf((size_t)z_len + 32, 0);
  std::memcpy(zbuf.data(), z_in, z_len);
  for (uint32_t p = 16; p + 32 <= z_len * 8u; ++p) {
    const uint64_t b = bits_at(zbuf.data(), p), b2 = bits_at(zbuf.data(), p + 64);
    const bool a = header_prefilter(b, b2), t = header_prefilter_tab(b, b2, tab);
    bad += a != t;
    h += a;
  }
  *hits = h;
  return bad;
}
