"""TEST TOOL: re-codes a baseline JPEG file as a PROGRESSIVE one with an arbitrary scan script (T.81 Annex G encoder
procedures), so that the decoders can be held to scripts PIL / libjpeg-turbo never writes: DC scans per component,
spectral selection without successive approximation, several refinement passes, restart intervals in AC scans.

The coefficients come from the oracle decoder (oracle/jpeg_dec_ref.c); quantisation tables and frame header are copied
from the source file.  Huffman tables are flat (every symbol a scan uses gets the same code length) -- legal, if not
small.  A file written here must decode, in PIL, to exactly the pixels of the file it was made from: the tests check
that before they trust it.
"""
from __future__ import annotations

import numpy as np

from oracle import jpeg_c as J

ZZ = [0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
      35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
      62, 63]


class _Bits:
    """Symbols and raw bits in emission order; Huffman codes are assigned once every symbol of the scan is known."""

    def __init__(self):
        self.items = []          # ("s", table, symbol) | ("b", value, nbits) | ("r", n)  (restart marker)

    def sym(self, table, s):
        self.items.append(("s", table, s))

    def bits(self, v, n):
        if n:
            self.items.append(("b", v & ((1 << n) - 1), n))

    def restart(self, n):
        self.items.append(("r", n))


def _flat_table(symbols):
    """Canonical code with one length for all symbols (the all-ones code stays unused): (BITS, HUFFVAL, {sym: (code, len)})."""
    syms = sorted(symbols) or [0]
    n = len(syms)
    length = max(1, int(np.ceil(np.log2(n + 1))))
    bits = [0] * 16
    bits[length - 1] = n
    return bits, syms, {s: (i, length) for i, s in enumerate(syms)}


def _nbits(v):
    return int(v).bit_length()


def _scan_blocks(planes, hs, vs, w, h, comps, interleaved):
    """Blocks of a scan in coding order: list of (component, block) with block = int array [64] in ZIG-ZAG order (a view
    into `planes`, so refinement passes see earlier writes -- not needed by an encoder, but handy), plus MCU boundaries."""
    order = []
    if interleaved:
        mcuy, mcux = planes[0].shape[0] // vs, planes[0].shape[1] // hs
        for my in range(mcuy):
            for mx in range(mcux):
                mcu = []
                for c in comps:
                    ch, cv = (hs, vs) if c == 0 else (1, 1)
                    for v in range(cv):
                        for hh in range(ch):
                            mcu.append((c, planes[c][my * cv + v, mx * ch + hh]))
                order.append(mcu)
    else:
        c = comps[0]
        cw = w if c == 0 else -(-w // hs)
        chh = h if c == 0 else -(-h // vs)
        if len(planes) == 1:
            cw, chh = w, h
        for by in range(-(-chh // 8)):
            for bx in range(-(-cw // 8)):
                order.append([(c, planes[c][by, bx])])
    return order


def encode_progressive(baseline: bytes, script, restart_interval: int = 0) -> bytes:
    """script: list of (components, ss, se, ah, al); components = tuple of frame indices (all of them, in order, or one)."""
    w, h, nc, hs, vs = J.info(baseline)
    nat = J.coefficients(baseline)
    planes = [p[..., ZZ].astype(np.int64) for p in nat]          # zig-zag order
    # header pieces of the source file: everything up to SOF0, SOF0 itself (as SOF2), no DHT
    out = bytearray(b"\xff\xd8")
    o = 2
    while True:
        assert baseline[o] == 0xFF
        m = baseline[o + 1]
        ln = (baseline[o + 2] << 8) | baseline[o + 3]
        seg = baseline[o:o + 2 + ln]
        if m == 0xDA:
            break
        if m == 0xC0:
            out += b"\xff\xc2" + seg[2:]
        elif m not in (0xC4, 0xDD):
            out += seg
        o += 2 + ln
    if restart_interval:
        out += b"\xff\xdd\x00\x04" + bytes([restart_interval >> 8, restart_interval & 255])
    for comps, ss, se, ah, al in script:
        comps = tuple(comps)
        interleaved = len(comps) > 1
        mcus = _scan_blocks(planes, hs if nc == 3 else 1, vs if nc == 3 else 1, w, h, comps, interleaved)
        B = _Bits()
        pred = {c: 0 for c in comps}
        eobrun = 0
        pending = []             # correction bits that belong behind the next EOBn symbol (refinement scans)

        def flush_eobrun():
            nonlocal eobrun, pending
            if eobrun > 0:
                nb = _nbits(eobrun) - 1
                B.sym(1, nb << 4)
                B.bits(eobrun, nb)
                eobrun = 0
            for b in pending:
                B.bits(b, 1)
            pending = []

        for mi, mcu in enumerate(mcus):
            if restart_interval and mi and mi % restart_interval == 0:
                flush_eobrun()
                B.restart((mi // restart_interval - 1) & 7)
                pred = {c: 0 for c in comps}
            for c, blk in mcu:
                if ss == 0:
                    if ah == 0:
                        v = int(blk[0]) >> al
                        d = v - pred[c]
                        pred[c] = v
                        n = _nbits(abs(d))
                        B.sym(0, n)
                        B.bits(d if d >= 0 else d - 1, n)
                    else:
                        B.bits((int(blk[0]) >> al) & 1, 1)
                    continue
                if ah == 0:
                    r = 0
                    for k in range(ss, se + 1):
                        v = int(blk[k])
                        t = abs(v) >> al
                        if t == 0:
                            r += 1
                            continue
                        flush_eobrun()
                        while r > 15:
                            B.sym(1, 0xF0)
                            r -= 16
                        n = _nbits(t)
                        B.sym(1, (r << 4) | n)
                        B.bits(t if v > 0 else ~t, n)
                        r = 0
                    if r > 0:
                        eobrun += 1
                        if eobrun == 0x7FFF:
                            flush_eobrun()
                else:
                    absv = [abs(int(blk[k])) >> al for k in range(64)]
                    eob = max([k for k in range(ss, se + 1) if absv[k] == 1], default=-1)
                    r = 0
                    br = []
                    for k in range(ss, se + 1):
                        t = absv[k]
                        if t == 0:
                            r += 1
                            continue
                        while r > 15 and k <= eob:
                            flush_eobrun()
                            B.sym(1, 0xF0)
                            r -= 16
                            for b in br:
                                B.bits(b, 1)
                            br = []
                        if t > 1:
                            br.append(t & 1)
                            continue
                        flush_eobrun()
                        B.sym(1, (r << 4) | 1)
                        B.bits(0 if int(blk[k]) < 0 else 1, 1)
                        for b in br:
                            B.bits(b, 1)
                        br = []
                        r = 0
                    if r > 0 or br:
                        eobrun += 1
                        pending += br
                        if eobrun == 0x7FFF or len(pending) > 900:
                            flush_eobrun()
        flush_eobrun()
        # tables for this scan, then the scan
        used = {0: set(), 1: set()}
        for it in B.items:
            if it[0] == "s":
                used[it[1]].add(it[2])
        codes = {}
        for cls in (0, 1):
            if (cls == 0) != (ss == 0) or (ss == 0 and ah):
                continue
            bits, vals, codes[cls] = _flat_table(used[cls])
            seg = bytes([cls << 4]) + bytes(bits) + bytes(vals)
            out += b"\xff\xc4" + bytes([(len(seg) + 2) >> 8, (len(seg) + 2) & 255]) + seg
        hdr = bytes([len(comps)]) + b"".join(bytes([c + 1, 0x00]) for c in comps) + bytes([ss, se, (ah << 4) | al])
        out += b"\xff\xda" + bytes([(len(hdr) + 2) >> 8, (len(hdr) + 2) & 255]) + hdr
        acc, cnt = 0, 0
        data = bytearray()

        def put(v, n):
            nonlocal acc, cnt
            acc = (acc << n) | v
            cnt += n
            while cnt >= 8:
                byte = (acc >> (cnt - 8)) & 255
                data.append(byte)
                if byte == 0xFF:
                    data.append(0)
                cnt -= 8
            acc &= (1 << cnt) - 1

        for it in B.items:
            if it[0] == "s":
                code, ln = codes[it[1]][it[2]]
                put(code, ln)
            elif it[0] == "b":
                put(it[1], it[2])
            else:
                if cnt:
                    put((1 << (8 - cnt)) - 1, 8 - cnt)
                data += bytes([0xFF, 0xD0 + it[1]])
        if cnt:
            put((1 << (8 - cnt)) - 1, 8 - cnt)
        out += data
    out += b"\xff\xd9"
    return bytes(out)


def component_ids_from_one(baseline: bytes) -> bool:
    """The encoder above writes scan component selectors 1, 2, 3: true when the frame uses those ids (PIL's files do)."""
    sof = baseline.index(b"\xff\xc0")
    nc = baseline[sof + 9]
    return [baseline[sof + 10 + 3 * c] for c in range(nc)] == list(range(1, nc + 1))


SCRIPTS = {
    # spectral selection only, DC interleaved
    "spectral": lambda nc: [(tuple(range(nc)), 0, 0, 0, 0)] + [((c,), 1, 63, 0, 0) for c in range(nc)],
    # DC per component, AC in three bands
    "dc_per_component": lambda nc: [((c,), 0, 0, 0, 0) for c in range(nc)] +
                                   [((c,), a, b, 0, 0) for c in range(nc) for a, b in ((1, 2), (3, 9), (10, 63))],
    # three refinement passes for DC and AC
    "deep_refinement": lambda nc: [(tuple(range(nc)), 0, 0, 0, 3)] + [((c,), 1, 63, 0, 3) for c in range(nc)] +
                                  [step for al in (2, 1, 0) for step in
                                   ([(tuple(range(nc)), 0, 0, al + 1, al)] + [((c,), 1, 63, al + 1, al) for c in range(nc)])],
    # DC scans per component with refinement, AC bands refined separately
    "mixed": lambda nc: [((c,), 0, 0, 0, 1) for c in range(nc)] + [((c,), 1, 5, 0, 2) for c in range(nc)] +
                        [((c,), 6, 63, 0, 1) for c in range(nc)] + [((c,), 1, 5, 2, 1) for c in range(nc)] +
                        [((c,), 0, 0, 1, 0) for c in range(nc)] + [((c,), 1, 5, 1, 0) for c in range(nc)] +
                        [((c,), 6, 63, 1, 0) for c in range(nc)],
}
