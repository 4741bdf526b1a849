// DEFLATE (RFC 1951) pieces shared by the PNG-input decoder's kernels (png_decode.cu): bit reader, Huffman table
// construction with zlib's acceptance rules, the dynamic-block header parser that doubles as the block FINDER's test,
// and the block decoder in its two modes (count / write).  Everything here is a __host__ __device__ inline function on
// plain pointers so that tests/native/pinf_host.cpp can drive exactly the same code sequentially on the CPU (a unit
// test of the pieces -- the library never runs them there).
//
// The page decodes PNG files inside loadImageFrom (U:27-121, called at P:1441-1463) with the platform's libpng + zlib;
// the stream format is RFC 1950 / 1951 and the acceptance rules below (over-subscribed / incomplete code sets, the
// missing end-of-block code, distances beyond the start) are zlib's inflate.c / inftrees.c rules restated.
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define PINF_HD __host__ __device__ __forceinline__
#else
#define PINF_HD inline
#endif

namespace stitch {
namespace pnginf {

constexpr int kLBits = 10;                 // direct-lookup bits of the literal/length table
constexpr int kDBits = 7;                  // ... of the distance table
constexpr int kLSub = 320, kDSub = 288;    // second-level entries (zlib's enough.c: 1332 - 1024 and 400 - 128 suffice)
constexpr int kLitEntries = (1 << kLBits) + kLSub, kDistEntries = (1 << kDBits) + kDSub;
constexpr uint32_t kRoot = 0xFFFFFFFFu;    // src[] value of a byte that is not copied from anywhere (a literal)
constexpr int kMaxLit = 288, kMaxDist = 32;
constexpr uint16_t kLink = 0x8000;         // table entry: second-level table at ((e >> 4) & 0x7FF), indexed by (e & 15) more bits

enum BlockStatus : int { kBlockOk = 0, kBlockBad = 1 };
enum BlockType : uint32_t { kStored = 0, kFixed = 1, kDynamic = 2 };

// LSB-first bit reader over 32-bit little-endian words.  The stream buffer is 4-byte aligned and padded with >= 16
// readable bytes past its end; running past the end is detected by the callers through pos().
struct BitReader {
  const uint32_t* w;
  uint64_t buf;
  uint32_t nbits;
  uint32_t next;   // index of the next word to load
  PINF_HD void init(const uint8_t* z, uint32_t bitpos) {
    w = reinterpret_cast<const uint32_t*>(z);
    next = bitpos >> 5;
    buf = (uint64_t)w[next++] >> (bitpos & 31);
    nbits = 32 - (bitpos & 31);
    refill();
  }
  PINF_HD void refill() {   // afterwards 33 <= nbits <= 64
    if (nbits <= 32) {
      buf |= (uint64_t)w[next++] << nbits;
      nbits += 32;
    }
  }
  PINF_HD uint32_t peek(int n) const { return (uint32_t)buf & ((1u << n) - 1u); }
  PINF_HD void skip(int n) { buf >>= n; nbits -= (uint32_t)n; }
  PINF_HD uint32_t take(int n) { const uint32_t v = peek(n); skip(n); return v; }
  PINF_HD uint32_t pos() const { return next * 32u - nbits; }
};

// One thread's decoding tables.  lit / dist are two-level lookup tables whose element e lives at [e * stride] (the
// kernels interleave the tables of a CTA's threads in shared memory).  A first-level element (indexed by the next
// kLBits / kDBits stream bits) is (symbol << 4) | code length, or kLink | (offset << 4) | n for codes longer than the
// first level: the symbol is then at [first-level size + offset + next n bits] as (symbol << 4) | remaining length.
// 0 = not a code, or a second level that did not fit (never seen; zlib's bound says it cannot happen): the canonical
// bit-by-bit path through count / sym (global scratch) decides.
struct HuffTables {
  uint16_t* lit;
  uint16_t* dist;
  uint32_t stride;
  uint16_t* lcount;   // [16]
  uint16_t* dcount;   // [16]
  uint16_t* lsym;     // [288] symbols ordered by (length, symbol)
  uint16_t* dsym;     // [32]
};
constexpr int kSlowWords = 16 + 16 + kMaxLit + kMaxDist;   // uint16 elements of the scratch block behind one HuffTables

PINF_HD void bind_slow(HuffTables& T, uint16_t* scratch) {
  T.lcount = scratch;
  T.dcount = scratch + 16;
  T.lsym = scratch + 32;
  T.dsym = scratch + 32 + kMaxLit;
}

PINF_HD uint32_t bitrev(uint32_t v, int n) {   // reverse the low n bits (n <= 16)
  v = ((v & 0x5555u) << 1) | ((v >> 1) & 0x5555u);
  v = ((v & 0x3333u) << 2) | ((v >> 2) & 0x3333u);
  v = ((v & 0x0F0Fu) << 4) | ((v >> 4) & 0x0F0Fu);
  v = ((v & 0x00FFu) << 8) | ((v >> 8) & 0x00FFu);
  return v >> (16 - n);
}

// zlib's verdict on a set of code lengths (inftrees.c): over-subscribed is always an error; incomplete is an error
// except for a literal/length or distance set whose only lengths are 1 (max == 1); an empty set is accepted.
// `allow_single` = false for the code-length alphabet.
PINF_HD bool lengths_ok(const uint16_t* count, bool allow_single) {
  int max = 15;
  while (max >= 1 && count[max] == 0) --max;
  if (max == 0) return true;
  int left = 1;
  for (int len = 1; len <= 15; ++len) {
    left <<= 1;
    left -= (int)count[len];
    if (left < 0) return false;
  }
  if (left > 0 && (!allow_single || max != 1)) return false;
  return true;
}

// Canonical Huffman tables from code lengths: count[], sym[] (ordered by length then symbol) and the two-level lookup
// table (indexed by stream bits, LSB first = code bits reversed).  The caller has checked lengths_ok.
PINF_HD void build_table(const uint8_t* lens, int n, uint16_t* count, uint16_t* sym, uint16_t* lut, uint32_t stride, int bits, int sub_cap) {
  for (int i = 0; i < 16; ++i) count[i] = 0;
  for (int i = 0; i < n; ++i) count[lens[i]]++;
  count[0] = 0;
  uint16_t offs[16];
  uint16_t first_code[16];
  uint16_t next_code[16];
  {
    uint32_t o = 0, code = 0;
    offs[0] = 0;
    first_code[0] = 0;
    for (int len = 1; len <= 15; ++len) {
      offs[len] = (uint16_t)o;
      o += count[len];
      code = (code + count[len - 1]) << 1;
      first_code[len] = (uint16_t)code;
    }
  }
  const uint32_t size = 1u << bits;
  for (uint32_t e = 0; e < size; ++e) lut[e * stride] = 0;
  bool any_long = false;
  for (int i = 0; i < 16; ++i) next_code[i] = first_code[i];
  for (int s = 0; s < n; ++s) {
    const int len = lens[s];
    if (!len) continue;
    sym[offs[len]++] = (uint16_t)s;
    const uint32_t rev = bitrev(next_code[len]++, len);
    if (len <= bits) {
      const uint16_t entry = (uint16_t)((s << 4) | len);
      for (uint32_t e = rev; e < size; e += 1u << len) lut[e * stride] = entry;
    } else {   // remember the longest code behind this first-level slot
      const uint32_t slot = (rev & (size - 1u)) * stride;
      const uint16_t extra = (uint16_t)(len - bits);
      if ((lut[slot] & 15) < extra) lut[slot] = (uint16_t)(kLink | extra);
      any_long = true;
    }
  }
  if (!any_long) return;
  uint32_t next = 0;
  for (uint32_t e = 0; e < size; ++e) {
    const uint16_t v = lut[e * stride];
    if (!(v & kLink)) continue;
    const uint32_t extra = v & 15u, sz = 1u << extra;
    if (next + sz > (uint32_t)sub_cap) {
      lut[e * stride] = 0;   // does not fit: the canonical path takes these codes
      continue;
    }
    lut[e * stride] = (uint16_t)(kLink | (next << 4) | extra);
    for (uint32_t i = 0; i < sz; ++i) lut[(size + next + i) * stride] = 0;
    next += sz;
  }
  for (int i = 0; i < 16; ++i) next_code[i] = first_code[i];
  for (int s = 0; s < n; ++s) {
    const int len = lens[s];
    if (!len) continue;
    const uint32_t rev = bitrev(next_code[len]++, len);
    if (len <= bits) continue;
    const uint16_t v = lut[(rev & (size - 1u)) * stride];
    if (!(v & kLink)) continue;
    const uint32_t off = (v >> 4) & 0x7FFu, sz = 1u << (v & 15u);
    const uint16_t entry = (uint16_t)((s << 4) | (len - bits));
    for (uint32_t i = rev >> bits; i < sz; i += 1u << (len - bits)) lut[(size + off + i) * stride] = entry;
  }
}

// Canonical decode one symbol bit by bit (codes longer than the lookup, and the holes of an incomplete set).
// Returns the symbol and its length through *used, or -1 when the bits are not a code.
PINF_HD int slow_symbol(uint64_t bits, const uint16_t* count, const uint16_t* sym, int* used) {
  int code = 0, first = 0, index = 0;
  for (int len = 1; len <= 15; ++len) {
    code |= (int)(bits & 1);
    bits >>= 1;
    const int c = count[len];
    if (code - c < first) {
      *used = len;
      return sym[index + (code - first)];
    }
    index += c;
    first += c;
    first <<= 1;
    code <<= 1;
  }
  return -1;
}

// Order in which the code-length-code lengths are stored (RFC 1951 3.2.7).
PINF_HD int clc_order(int i) {
  // 16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15 packed 5 bits each
  const uint64_t lo = 16ull | (17ull << 5) | (18ull << 10) | (0ull << 15) | (8ull << 20) | (7ull << 25) | (9ull << 30) | (6ull << 35) |
                      (10ull << 40) | (5ull << 45) | (11ull << 50) | (4ull << 55);
  const uint64_t hi = 12ull | (3ull << 5) | (13ull << 10) | (2ull << 15) | (14ull << 20) | (1ull << 25) | (15ull << 30);
  return i < 12 ? (int)((lo >> (5 * i)) & 31) : (int)((hi >> (5 * (i - 12))) & 31);
}

// The cheap part of the dynamic-header test, run for EVERY bit offset by the finder: block type 2, HLIT <= 29 (zlib:
// "too many length or distance symbols" for nlen > 286 || ndist > 30), and a complete code-length code.
// `b` holds the 64 stream bits at the offset, `b2` the 64 after them.
PINF_HD bool header_prefilter(uint64_t b, uint64_t b2) {
  if (((b >> 1) & 3) != 2) return false;
  const uint32_t hlit = (uint32_t)(b >> 3) & 31, hdist = (uint32_t)(b >> 8) & 31, hclen = (uint32_t)(b >> 13) & 15;
  if (hlit > 29 || hdist > 29) return false;
  // (hclen + 4) 3-bit lengths from bit 17: up to 57 bits, ending at bit 74
  uint64_t v = (b >> 17) | (b2 << 47);   // 111 valid bits
  int left = 128;                        // Kraft sum in units of 2^-7
  const int n = (int)hclen + 4;
  for (int i = 0; i < 19; ++i) {
    if (i < n) {
      const int len = (int)(v & 7);
      v >>= 3;
      if (len) left -= 128 >> len;
    }
  }
  return left == 0;
}

// The same test with the Kraft sum taken three 3-bit lengths at a time from a 512-entry table (kraft3_entry(e) for
// e = 0 .. 511; the finder keeps it in shared memory): 7 look-ups instead of the 19-step loop.
PINF_HD uint8_t kraft3_entry(int e) {
  int sum = 0;
  for (int k = 0; k < 3; ++k) {
    const int len = (e >> (3 * k)) & 7;
    if (len) sum += 128 >> len;
  }
  return (uint8_t)sum;
}
PINF_HD bool header_prefilter_tab(uint64_t b, uint64_t b2, const uint8_t* kraft3) {
  const uint32_t lo = (uint32_t)b;
  if (((lo >> 1) & 3u) != 2u || ((lo >> 3) & 31u) > 29u || ((lo >> 8) & 31u) > 29u) return false;
  const uint32_t n3 = (((lo >> 13) & 15u) + 4u) * 3u;   // bits of code-length lengths, 12 .. 57
  uint64_t v = (b >> 17) | (b2 << 47);
  v &= (1ull << n3) - 1ull;
  uint32_t sum = 0;
  for (int k = 0; k < 7; ++k) sum += kraft3[(uint32_t)(v >> (9 * k)) & 511u];
  return sum == 128u;
}

// Full parse of a dynamic block header at the reader's position (which is just past the 3 header bits): reads the
// code lengths into lens[0 .. nlen + ndist) and checks what zlib checks.  Returns false for an invalid header.
PINF_HD bool read_dynamic_lengths(BitReader& br, uint8_t* lens, int* nlen_out, int* ndist_out, uint32_t z_bits) {
  br.refill();
  const int nlen = (int)br.take(5) + 257, ndist = (int)br.take(5) + 1, ncode = (int)br.take(4) + 4;
  if (nlen > 286 || ndist > 30) return false;
  uint8_t cl[19];
  for (int i = 0; i < 19; ++i) cl[i] = 0;
  for (int i = 0; i < ncode; ++i) {
    br.refill();
    cl[clc_order(i)] = (uint8_t)br.take(3);
  }
  uint16_t ccount[16];
  for (int i = 0; i < 16; ++i) ccount[i] = 0;
  for (int i = 0; i < 19; ++i) ccount[cl[i]]++;
  ccount[0] = 0;
  if (!lengths_ok(ccount, false)) return false;
  // 7-bit direct table for the code-length code
  uint8_t clut[128];
  {
    uint32_t code = 0;
    uint16_t next_code[8];
    next_code[0] = 0;
    for (int len = 1; len <= 7; ++len) {
      code = (code + ccount[len - 1]) << 1;
      next_code[len] = (uint16_t)code;
    }
    for (int e = 0; e < 128; ++e) clut[e] = 0;
    for (int s = 0; s < 19; ++s) {
      const int len = cl[s];
      if (!len) continue;
      const uint32_t c = next_code[len]++;
      for (uint32_t e = bitrev(c, len); e < 128; e += 1u << len) clut[e] = (uint8_t)((s << 3) | len);
    }
  }
  const int total = nlen + ndist;
  int have = 0;
  while (have < total) {
    br.refill();
    const uint8_t e = clut[br.peek(7)];
    if (!e) return false;
    br.skip(e & 7);
    const int s = e >> 3;
    if (s < 16) {
      lens[have++] = (uint8_t)s;
    } else {
      int rep, val = 0;
      if (s == 16) {
        if (have == 0) return false;
        val = lens[have - 1];
        rep = 3 + (int)br.take(2);
      } else if (s == 17) {
        rep = 3 + (int)br.take(3);
      } else {
        rep = 11 + (int)br.take(7);
      }
      if (have + rep > total) return false;
      while (rep--) lens[have++] = (uint8_t)val;
    }
    if (br.pos() > z_bits) return false;
  }
  if (lens[256] == 0) return false;   // zlib: "invalid code -- missing end-of-block"
  uint16_t count[16];
  for (int i = 0; i < 16; ++i) count[i] = 0;
  for (int i = 0; i < nlen; ++i) count[lens[i]]++;
  count[0] = 0;
  if (!lengths_ok(count, true)) return false;
  for (int i = 0; i < 16; ++i) count[i] = 0;
  for (int i = 0; i < ndist; ++i) count[lens[nlen + i]]++;
  count[0] = 0;
  if (!lengths_ok(count, true)) return false;
  *nlen_out = nlen;
  *ndist_out = ndist;
  return true;
}

// Code lengths of the fixed Huffman block (RFC 1951 3.2.6).
PINF_HD void fixed_lengths(uint8_t* lens) {
  for (int i = 0; i < 144; ++i) lens[i] = 8;
  for (int i = 144; i < 256; ++i) lens[i] = 9;
  for (int i = 256; i < 280; ++i) lens[i] = 7;
  for (int i = 280; i < 288; ++i) lens[i] = 8;
  for (int i = 0; i < 30; ++i) lens[288 + i] = 5;
}

// One symbol through the two-level table; -1 = not a code.
PINF_HD int lookup_symbol(BitReader& br, const uint16_t* lut, uint32_t stride, int bits, const uint16_t* count, const uint16_t* sym) {
  uint16_t e = lut[br.peek(bits) * stride];
  if (e & kLink) {
    br.skip(bits);
    e = lut[((1u << bits) + ((e >> 4) & 0x7FFu) + br.peek(e & 15)) * stride];
    if (!e) return -1;
  } else if (!e) {
    int used;
    const int s = slow_symbol(br.buf, count, sym, &used);
    if (s >= 0) br.skip(used);
    return s;
  }
  br.skip(e & 15);
  return e >> 4;
}

// A match: bytes [pos, pos + len) are copies of the bytes `dist` before them.
struct Token {
  uint32_t pos;
  uint32_t len_dist;   // (len << 15) | (dist - 1)
};

// What a finished block leaves behind.
struct BlockRec {
  uint32_t bitpos;    // of its 3 header bits
  uint32_t end_bit;   // first bit after its end-of-block code (stored: after its last byte)
  uint32_t n_out;     // bytes it produces
  uint32_t flags;     // bit 0: BFINAL, bits 1-2: type, bit 3: valid
  uint32_t n_tok;     // matches in it
  uint32_t first_sub; // index of its first window of subsequence records (kNoSub: decoded by one thread)
  uint32_t next;      // index of the candidate that starts where this block ends (kNoSub: none / not looked up)
};
constexpr uint32_t kRecFinal = 1, kRecValid = 8, kNoSub = 0xFFFFFFFFu;

// A block being decoded by one thread.  The work is cut into block_begin / symbol_step / block_end so that a kernel can
// run symbol_step under a warp-wide vote: 32 lanes that walk 32 different blocks stay converged, one symbol per lane
// and iteration, instead of drifting apart through the loop's many exits (measured: 30x).
enum : uint32_t { kRunning = 0, kDone = 1, kBad = 2 };
struct DecState {
  BitReader br;
  uint32_t z_bits;
  uint32_t pos, out_start, out_limit, n_tok;
  uint32_t hdr;      // the block's 3 header bits
  uint32_t bitpos, end_bit;
  uint32_t status;
  // the last match, as the stream bits that coded it: flat image areas repeat one {length, distance} thousands of
  // times, and a token whose bits equal the last one's needs no table look-up (2x on such streams, one compare elsewhere)
  uint32_t rep_bits, rep_nbits, rep_len_dist, rep_len;
};

// Header of the block at `bitpos`: a stored block is finished here (its bytes copied in write mode), for a Huffman
// block the tables are built into T and the state is left kRunning at its first symbol.
template <bool WRITE>
PINF_HD void block_begin(DecState& st, const uint8_t* z, uint32_t z_len, uint32_t bitpos, HuffTables& T, uint8_t* lens, uint32_t out_pos,
                         uint32_t out_limit, uint8_t* lit) {
  st.z_bits = z_len * 8u;
  st.pos = st.out_start = out_pos;
  st.out_limit = out_limit;
  st.n_tok = 0;
  st.bitpos = bitpos;
  st.end_bit = 0;
  st.status = kBad;
  st.hdr = 0;
  st.rep_nbits = 0;
  st.rep_bits = st.rep_len_dist = st.rep_len = 0;
  if (bitpos + 3 > st.z_bits) return;
  st.br.init(z, bitpos);
  st.hdr = st.br.take(3);
  const uint32_t type = st.hdr >> 1;
  if (type == kStored) {
    const uint32_t byte = (st.br.pos() + 7) >> 3;
    if (byte + 4 > z_len) return;
    const uint32_t len = z[byte] | ((uint32_t)z[byte + 1] << 8), nlen = z[byte + 2] | ((uint32_t)z[byte + 3] << 8);
    if ((len ^ 0xFFFFu) != nlen) return;
    if (byte + 4 + len > z_len || len > out_limit - out_pos) return;
    if (WRITE)
      for (uint32_t k = 0; k < len; ++k) lit[out_pos + k] = z[byte + 4 + k];
    st.pos = out_pos + len;
    st.end_bit = (byte + 4 + len) * 8u;
    st.status = kDone;
  } else if (type == kFixed || type == kDynamic) {
    int nlen = 288, ndist = 30;
    if (type == kFixed) {
      fixed_lengths(lens);
    } else if (!read_dynamic_lengths(st.br, lens, &nlen, &ndist, st.z_bits)) {
      return;
    }
    build_table(lens, nlen, T.lcount, T.lsym, T.lit, T.stride, kLBits, kLSub);
    build_table(lens + nlen, ndist, T.dcount, T.dsym, T.dist, T.stride, kDBits, kDSub);
    st.status = kRunning;
  }
}

// One literal/length symbol (and the distance behind a length).
//   WRITE = false   count: pos advances by the bytes, n_tok by the matches
//   WRITE = true    literals go to lit[]; a match becomes a token in tok[] (its bytes are filled in later, in parallel:
//                   a thread that walks a block alone must not copy 258 bytes per symbol)
template <bool WRITE>
PINF_HD void symbol_step(DecState& st, const HuffTables& T, uint8_t* lit, Token* tok) {
  BitReader& br = st.br;
  br.refill();
  if (st.rep_nbits && ((uint32_t)br.buf & ((1u << st.rep_nbits) - 1u)) == st.rep_bits) {   // the same token again
    const uint32_t len = st.rep_len;
    if (br.pos() + st.rep_nbits > st.z_bits || len > st.out_limit - st.pos || (WRITE && (st.rep_len_dist & 0x7FFFu) + 1u > st.pos)) {
      st.status = kBad;
      return;
    }
    br.skip((int)st.rep_nbits);
    if (WRITE) {
      tok[st.n_tok].pos = st.pos;
      tok[st.n_tok].len_dist = st.rep_len_dist;
    }
    ++st.n_tok;
    st.pos += len;
    return;
  }
  const uint32_t bits0 = (uint32_t)br.buf, pos0 = br.pos();
  const int sym = lookup_symbol(br, T.lit, T.stride, kLBits, T.lcount, T.lsym);
  if (sym < 256) {
    if (sym < 0 || st.pos >= st.out_limit) {
      st.status = kBad;
    } else {
      if (WRITE) lit[st.pos] = (uint8_t)sym;
      ++st.pos;
    }
    return;
  }
  if (sym == 256) {
    st.end_bit = br.pos();
    st.status = st.end_bit > st.z_bits ? kBad : kDone;
    return;
  }
  if (sym > 285) {
    st.status = kBad;
    return;
  }
  uint32_t len;
  {
    const uint32_t idx = (uint32_t)sym - 257u;
    if (idx < 8) {
      len = 3 + idx;
    } else if (idx == 28) {
      len = 258;
    } else {
      const uint32_t eb = (idx >> 2) - 1;
      len = 3 + ((4 + (idx & 3)) << eb) + br.take((int)eb);
    }
  }
  br.refill();
  const int dsym = lookup_symbol(br, T.dist, T.stride, kDBits, T.dcount, T.dsym);
  if (dsym < 0 || dsym > 29) {
    st.status = kBad;
    return;
  }
  uint32_t dist;
  if (dsym < 4) {
    dist = 1 + (uint32_t)dsym;
  } else {
    const uint32_t eb = ((uint32_t)dsym >> 1) - 1;
    dist = 1 + ((2 + ((uint32_t)dsym & 1)) << eb) + br.take((int)eb);
  }
  // pos <= out_limit always; zlib: "invalid distance too far back" (no preset dictionary) -- positions are absolute in write mode
  if (br.pos() > st.z_bits || len > st.out_limit - st.pos || (WRITE && dist > st.pos)) {
    st.status = kBad;
    return;
  }
  if (WRITE) {
    tok[st.n_tok].pos = st.pos;
    tok[st.n_tok].len_dist = (len << 15) | (dist - 1u);
  }
  ++st.n_tok;
  st.pos += len;
  const uint32_t used = br.pos() - pos0;   // bits0 holds >= 32 of them: refill() ran before
  st.rep_nbits = used <= 31u ? used : 0u;
  st.rep_bits = bits0 & ((1u << (used & 31u)) - 1u);
  st.rep_len_dist = (len << 15) | (dist - 1u);
  st.rep_len = len;
}

PINF_HD void block_end(const DecState& st, BlockRec* rec) {
  rec->bitpos = st.bitpos;
  rec->end_bit = st.end_bit;
  rec->n_out = st.pos - st.out_start;
  rec->flags = (st.hdr & 1) | ((st.hdr >> 1) << 1) | kRecValid;
  rec->n_tok = st.n_tok;
  rec->first_sub = kNoSub;
  rec->next = kNoSub;
}

// A whole block at `bitpos`, any type, by one thread on its own (the chain walker, the CPU drive of the tests).
// count mode: rec filled in.  write mode: out_pos = absolute position of the block's first byte, tok = its token slots.
template <bool WRITE>
PINF_HD int decode_block(const uint8_t* z, uint32_t z_len, uint32_t bitpos, HuffTables& T, uint8_t* lens, BlockRec* rec, uint32_t out_pos,
                         uint32_t out_limit, uint8_t* lit, Token* tok) {
  DecState st;
  block_begin<WRITE>(st, z, z_len, bitpos, T, lens, out_pos, out_limit, lit);
  while (st.status == kRunning) symbol_step<WRITE>(st, T, lit, tok);
  if (st.status != kDone) return kBlockBad;
  if (rec) block_end(st, rec);
  return kBlockOk;
}

// 64 stream bits starting at bit offset p (for the finder); the buffer is padded.
PINF_HD uint64_t bits_at(const uint8_t* z, uint32_t p) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(z);
  const uint32_t i = p >> 5, sh = p & 31;
  const uint64_t lo = (uint64_t)w[i] | ((uint64_t)w[i + 1] << 32);
  const uint64_t hi = (uint64_t)w[i + 2];
  return sh ? (lo >> sh) | (hi << (64 - sh)) : lo;
}

// Hash of a bit position for the candidate table (open addressing, linear probing; key 0 = empty, keys are bitpos + 1).
PINF_HD uint32_t hash_pos(uint32_t p) {
  p *= 0x9E3779B1u;
  return p ^ (p >> 15);
}

}  // namespace pnginf
}  // namespace stitch
