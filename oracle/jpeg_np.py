"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the baseline-JPEG decode that sits immediately before the predict path
(reference call site /root/reference/src/model/model.py:23, `Image.open(image_file.file)`;
the pixels come from the third-party decoder behind that call).

Third-party dependency carrying the arithmetic: Pillow 12.2.0 (unpinned in the reference's requirements.txt:8) with its
bundled libjpeg-turbo 3.1.4.1.  This file restates libjpeg-turbo's *published* default decode pipeline:

  * entropy decoding            — ITU-T T.81 Annex F.2.2 (jdhuff.c: decode_mcu_slow)
  * dequantisation + IDCT       — the accurate integer IDCT "islow" (jidctint.c: jpeg_idct_islow, CONST_BITS=13,
                                  PASS1_BITS=2), which is JDCT_DEFAULT and what Pillow uses at full scale
  * chroma upsampling           — "fancy" triangle filters (jdsample.c: h2v1_fancy_upsample / h2v2_fancy_upsample,
                                  with jdmainct.c's edge-row duplication), libjpeg's default do_fancy_upsampling=TRUE
  * colour conversion           — YCbCr -> RGB fixed-point tables (jdcolor.c: build_ycc_rgb_table, SCALEBITS=16)

Parity is PINNED: tests/test_oracle_jpeg.py checks this restatement bit-for-bit against Pillow itself
(`Image.open(...).convert('RGB')`) on generated files covering 4:4:4 / 4:2:2 / 4:2:0 / greyscale, odd sizes, restart
intervals and optimised Huffman tables, and tests/golden/jpeg_vectors.npz carries files + Pillow's pixels.

Scope (everything else is reported as unsupported and stays on the host decoder, exactly like the GPU path):
baseline / extended-sequential Huffman (SOF0/SOF1), 8-bit, one interleaved scan, 1 or 3 components,
luma sampling 1x1, 2x1 or 2x2 with 1x1 chroma, no Adobe marker, chroma width > 2 samples.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

# zigzag position -> natural (row-major) index, T.81 Figure A.6 (jutils.c: jpeg_natural_order)
ZIGZAG = np.array([
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47,
    55, 62, 63], dtype=np.int32)


class Unsupported(Exception):
    """The file is valid input for the service but outside the GPU decoder's scope (host decode handles it)."""


@dataclass
class HuffTable:
    counts: np.ndarray          # [16] number of codes of length 1..16
    values: np.ndarray          # symbols in code order
    mincode: list = field(default_factory=list)
    maxcode: list = field(default_factory=list)
    valptr: list = field(default_factory=list)

    def finish(self):
        code = 0
        k = 0
        self.mincode, self.maxcode, self.valptr = [0] * 17, [-1] * 17, [0] * 17
        for length in range(1, 17):
            n = int(self.counts[length - 1])
            self.valptr[length] = k
            self.mincode[length] = code
            if n:
                self.maxcode[length] = code + n - 1
            code = (code + n) << 1
            k += n
        return self


@dataclass
class Component:
    ident: int
    h: int
    v: int
    tq: int
    td: int = 0
    ta: int = 0


@dataclass
class Header:
    width: int = 0
    height: int = 0
    comps: list = field(default_factory=list)
    qt: dict = field(default_factory=dict)        # tq -> [64] natural order
    dc: dict = field(default_factory=dict)
    ac: dict = field(default_factory=dict)
    restart_interval: int = 0
    scan_start: int = 0
    scan_end: int = 0                             # index of the marker that ends the entropy-coded data


def parse(data: bytes) -> Header:
    """Marker walk (T.81 Annex B).  Raises Unsupported for anything outside the scope in the module docstring and
    ValueError for files that are not JPEG at all."""
    if len(data) < 4 or data[0] != 0xFF or data[1] != 0xD8:
        raise ValueError("not a JPEG (no SOI)")
    hd = Header()
    i = 2
    seen_sof = False
    while True:
        if i + 4 > len(data):
            raise ValueError("truncated before SOS")
        if data[i] != 0xFF:
            raise ValueError("marker expected")
        while data[i + 1] == 0xFF:      # fill bytes
            i += 1
        m = data[i + 1]
        i += 2
        if m == 0xD8 or (0xD0 <= m <= 0xD7) or m == 0x01:
            continue
        seg_len = (data[i] << 8) | data[i + 1]
        seg = data[i + 2:i + seg_len]
        if len(seg) != seg_len - 2:
            raise ValueError("truncated segment")
        if m == 0xDB:                   # DQT
            p = 0
            while p < len(seg):
                pq, tq = seg[p] >> 4, seg[p] & 15
                p += 1
                if pq != 0:
                    raise Unsupported("16-bit quantisation table")
                tab = np.zeros(64, np.int32)
                tab[ZIGZAG] = np.frombuffer(seg[p:p + 64], np.uint8)
                hd.qt[tq] = tab
                p += 64
        elif m in (0xC0, 0xC1):         # SOF0 / SOF1
            if seen_sof:
                raise Unsupported("multiple frames")
            seen_sof = True
            if seg[0] != 8:
                raise Unsupported("sample precision != 8")
            hd.height = (seg[1] << 8) | seg[2]
            hd.width = (seg[3] << 8) | seg[4]
            nf = seg[5]
            if nf not in (1, 3):
                raise Unsupported("component count")
            for c in range(nf):
                cid, hv, tq = seg[6 + 3 * c], seg[7 + 3 * c], seg[8 + 3 * c]
                hd.comps.append(Component(cid, hv >> 4, hv & 15, tq))
        elif 0xC2 <= m <= 0xCF and m not in (0xC4, 0xC8, 0xCC):
            raise Unsupported("progressive / lossless / arithmetic frame")
        elif m == 0xCC:
            raise Unsupported("arithmetic conditioning")
        elif m == 0xC4:                 # DHT
            p = 0
            while p < len(seg):
                tc, th = seg[p] >> 4, seg[p] & 15
                counts = np.frombuffer(seg[p + 1:p + 17], np.uint8).astype(np.int32)
                n = int(counts.sum())
                vals = np.frombuffer(seg[p + 17:p + 17 + n], np.uint8).astype(np.int32)
                (hd.dc if tc == 0 else hd.ac)[th] = HuffTable(counts, vals).finish()
                p += 17 + n
        elif m == 0xDD:                 # DRI
            hd.restart_interval = (seg[0] << 8) | seg[1]
        elif m == 0xEE and seg[:5] == b"Adobe":
            raise Unsupported("Adobe colour transform marker")
        elif m == 0xDA:                 # SOS
            if not seen_sof:
                raise ValueError("SOS before SOF")
            ns = seg[0]
            if ns != len(hd.comps):
                raise Unsupported("non-interleaved / multiple scans")
            for c in range(ns):
                cs, tt = seg[1 + 2 * c], seg[2 + 2 * c]
                if cs != hd.comps[c].ident:
                    raise Unsupported("scan component order")
                hd.comps[c].td, hd.comps[c].ta = tt >> 4, tt & 15
            ss, se, ahal = seg[1 + 2 * ns], seg[2 + 2 * ns], seg[3 + 2 * ns]
            if ss != 0 or se != 63 or ahal != 0:
                raise Unsupported("spectral selection / successive approximation")
            hd.scan_start = i + seg_len
            break
        i += seg_len
    # end of the entropy-coded segment: first marker that is neither a stuffed zero nor RSTn
    j = hd.scan_start
    n = len(data)
    while True:
        j = data.find(b"\xff", j)
        if j < 0 or j + 1 >= n:
            raise ValueError("no EOI")
        nxt = data[j + 1]
        if nxt == 0x00 or (0xD0 <= nxt <= 0xD7) or nxt == 0xFF:
            j += 1 if nxt == 0xFF else 2
            continue
        break
    hd.scan_end = j
    _check_scope(hd)
    return hd


def _check_scope(hd: Header) -> None:
    if hd.width < 1 or hd.height < 1:
        raise ValueError("empty frame")
    for c in hd.comps:
        if c.tq not in hd.qt:
            raise ValueError("missing quantisation table")
        if c.td not in hd.dc or c.ta not in hd.ac:
            raise ValueError("missing Huffman table")
    if len(hd.comps) == 1:
        hd.comps[0].h = hd.comps[0].v = 1           # a single-component scan is never interleaved (T.81 A.2.2)
        return
    ids = tuple(c.ident for c in hd.comps)
    if ids == (82, 71, 66):
        raise Unsupported("RGB component ids")
    y, cb, cr = hd.comps
    if (cb.h, cb.v, cr.h, cr.v) != (1, 1, 1, 1) or (y.h, y.v) not in ((1, 1), (2, 1), (2, 2)):
        raise Unsupported("sampling factors")
    if y.h == 2 and (hd.width + 1) // 2 <= 2:
        raise Unsupported("chroma narrower than 3 samples (libjpeg switches to replication)")


def geometry(hd: Header):
    """MCU grid and per-component block grids (jdinput.c: initial_setup / per_scan_setup)."""
    hmax = max(c.h for c in hd.comps)
    vmax = max(c.v for c in hd.comps)
    mcux = (hd.width + 8 * hmax - 1) // (8 * hmax)
    mcuy = (hd.height + 8 * vmax - 1) // (8 * vmax)
    return hmax, vmax, mcux, mcuy


class _Bits:
    """MSB-first bit reader over one restart segment with FF00 un-stuffing (jdhuff.c: jpeg_fill_bit_buffer).
    Past the end of the segment it supplies zero bits, as libjpeg does after its 'premature end' warning."""

    def __init__(self, data: bytes, pos: int, end: int):
        self.d, self.p, self.end = data, pos, end
        self.acc = 0
        self.n = 0
        self.overrun = False

    def _fill(self, need: int):
        while self.n < need:
            if self.p < self.end:
                b = self.d[self.p]
                self.p += 1
                if b == 0xFF and self.p < self.end and self.d[self.p] == 0x00:
                    self.p += 1
            else:
                b = 0
                self.overrun = True
            self.acc = ((self.acc << 8) | b) & 0xFFFFFFFFFFFF
            self.n += 8

    def get(self, nbits: int) -> int:
        if nbits == 0:
            return 0
        self._fill(nbits)
        self.n -= nbits
        return (self.acc >> self.n) & ((1 << nbits) - 1)

    def symbol(self, t: HuffTable) -> int:
        code = 0
        for length in range(1, 17):
            code = (code << 1) | self.get(1)
            if t.maxcode[length] >= 0 and code <= t.maxcode[length]:
                return int(t.values[t.valptr[length] + code - t.mincode[length]])
        return 0            # invalid code: libjpeg warns and returns 0


def _extend(v: int, s: int) -> int:
    return v - (1 << s) + 1 if v < (1 << (s - 1)) else v


def restart_segments(data: bytes, hd: Header):
    """Byte ranges [start, end) of the independently decodable pieces of the entropy-coded data."""
    segs = []
    start = hd.scan_start
    j = start
    while True:
        j = data.find(b"\xff", j, hd.scan_end)
        if j < 0:
            break
        nxt = data[j + 1]
        if 0xD0 <= nxt <= 0xD7:
            segs.append((start, j))
            start = j + 2
            j += 2
        else:
            j += 2 if nxt == 0x00 else 1
    segs.append((start, hd.scan_end))
    return segs


def entropy_decode(data: bytes, hd: Header):
    """Returns one int16 array [blocks_y, blocks_x, 64] per component, coefficients in natural order, NOT dequantised
    (jdhuff.c: decode_mcu_slow; jdcoefct.c block placement)."""
    hmax, vmax, mcux, mcuy = geometry(hd)
    planes = [np.zeros((mcuy * c.v, mcux * c.h, 64), np.int16) for c in hd.comps]
    segs = restart_segments(data, hd)
    ri = hd.restart_interval if hd.restart_interval > 0 else mcux * mcuy
    total = mcux * mcuy
    overrun = False
    for si, (s0, s1) in enumerate(segs):
        first = si * ri
        if first >= total:
            break
        br = _Bits(data, s0, s1)
        pred = [0] * len(hd.comps)
        for mcu in range(first, min(first + ri, total)):
            my, mx = divmod(mcu, mcux)
            for ci, c in enumerate(hd.comps):
                dct, act = hd.dc[c.td], hd.ac[c.ta]
                for by in range(c.v):
                    for bx in range(c.h):
                        blk = planes[ci][my * c.v + by, mx * c.h + bx]
                        s = br.symbol(dct)
                        diff = _extend(br.get(s), s) if s else 0
                        pred[ci] += diff
                        blk[0] = np.int16(((pred[ci] + 32768) & 0xFFFF) - 32768)
                        k = 1
                        while k < 64:
                            rs = br.symbol(act)
                            r, s = rs >> 4, rs & 15
                            if s:
                                k += r
                                v = _extend(br.get(s), s)
                                blk[ZIGZAG[min(k, 63)]] = v
                                k += 1
                            else:
                                if r != 15:
                                    break
                                k += 16
        overrun = overrun or br.overrun
    return planes, overrun


# ---------------------------------------------------------------- jidctint.c: jpeg_idct_islow
_C = dict(f0_298=2446, f0_390=3196, f0_541=4433, f0_765=6270, f0_899=7373, f1_175=9633, f1_501=12299, f1_847=15137,
          f1_961=16069, f2_053=16819, f2_562=20995, f3_072=25172)


def _idct_1d(x0, x1, x2, x3, x4, x5, x6, x7):
    """The shared even/odd butterfly; inputs int64 arrays, returns the eight un-descaled outputs."""
    z1 = (x2 + x6) * _C["f0_541"]
    tmp2 = z1 + x6 * (-_C["f1_847"])
    tmp3 = z1 + x2 * _C["f0_765"]
    tmp0 = (x0 + x4) << 13
    tmp1 = (x0 - x4) << 13
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = x7, x5, x3, x1
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * _C["f1_175"]
    t0 = t0 * _C["f0_298"]
    t1 = t1 * _C["f2_053"]
    t2 = t2 * _C["f3_072"]
    t3 = t3 * _C["f1_501"]
    z1 = z1 * (-_C["f0_899"])
    z2 = z2 * (-_C["f2_562"])
    z3 = z3 * (-_C["f1_961"]) + z5
    z4 = z4 * (-_C["f0_390"]) + z5
    t0 = t0 + z1 + z3
    t1 = t1 + z2 + z4
    t2 = t2 + z2 + z3
    t3 = t3 + z1 + z4
    return (tmp10 + t3, tmp11 + t2, tmp12 + t1, tmp13 + t0, tmp13 - t0, tmp12 - t1, tmp11 - t2, tmp10 - t3)


def idct_islow(coef: np.ndarray, quant: np.ndarray) -> np.ndarray:
    """coef [..., 64] int16 natural order, quant [64] -> samples [..., 8, 8] uint8.
    Pass 1 works on columns (descale by CONST_BITS - PASS1_BITS = 11), pass 2 on rows (descale by 18), then
    + 128 through libjpeg's range-limit table, which wraps its index modulo 1024 before clamping."""
    x = (coef.astype(np.int64) * quant.astype(np.int64)).reshape(coef.shape[:-1] + (8, 8))
    cols = _idct_1d(*[x[..., r, :] for r in range(8)])
    ws = np.stack([(c + (1 << 10)) >> 11 for c in cols], axis=-2)          # [..., 8 rows, 8 cols]
    rows = _idct_1d(*[ws[..., :, c] for c in range(8)])
    out = np.stack([(r + (1 << 17)) >> 18 for r in rows], axis=-1)          # [..., 8 rows, 8 cols]
    idx = out & 1023
    signed = np.where(idx < 512, idx, idx - 1024)
    return np.clip(signed + 128, 0, 255).astype(np.uint8)


def component_planes(planes, hd: Header):
    """IDCT every block and lay the samples out as one [blocks_y*8, blocks_x*8] uint8 plane per component."""
    out = []
    for blocks, c in zip(planes, hd.comps):
        s = idct_islow(blocks, hd.qt[c.tq])                     # [by, bx, 8, 8]
        by, bx = s.shape[:2]
        out.append(s.transpose(0, 2, 1, 3).reshape(by * 8, bx * 8))
    return out


# ---------------------------------------------------------------- jdsample.c fancy upsampling
def _h2v1_fancy(p: np.ndarray, w_ds: int) -> np.ndarray:
    """[rows, >= w_ds] -> [rows, 2*w_ds]; 3/4 - 1/4 triangle, rounding biases 1 (even outputs) and 2 (odd)."""
    a = p[:, :w_ds].astype(np.int32)
    out = np.empty((a.shape[0], 2 * w_ds), np.int32)
    out[:, 0] = a[:, 0]
    out[:, 2::2] = (3 * a[:, 1:] + a[:, :-1] + 1) >> 2
    out[:, 1:-1:2] = (3 * a[:, :-1] + a[:, 1:] + 2) >> 2
    out[:, -1] = a[:, -1]
    return out.astype(np.uint8)


def _h2v2_fancy(p: np.ndarray, w_ds: int, h_ds: int) -> np.ndarray:
    """[>= h_ds, >= w_ds] -> [2*h_ds, 2*w_ds].  Vertical 3:1 blend with the nearer row first (rows outside
    [0, h_ds) are duplicates of the edge rows, jdmainct.c), then horizontal 3:1 with biases 8 / 7, >> 4."""
    a = p[:h_ds, :w_ds].astype(np.int32)
    up = np.vstack([a[:1], a[:-1]])          # row above (row 0 duplicated)
    dn = np.vstack([a[1:], a[-1:]])          # row below (last real row duplicated)
    out = np.empty((2 * h_ds, 2 * w_ds), np.int32)
    for v, other in ((0, up), (1, dn)):
        s = 3 * a + other                    # column sums
        o = np.empty((h_ds, 2 * w_ds), np.int32)
        o[:, 0] = (s[:, 0] * 4 + 8) >> 4
        o[:, 2::2] = (3 * s[:, 1:] + s[:, :-1] + 8) >> 4
        o[:, 1:-1:2] = (3 * s[:, :-1] + s[:, 1:] + 7) >> 4
        o[:, -1] = (s[:, -1] * 4 + 7) >> 4
        out[v::2] = o
    return out.astype(np.uint8)


# ---------------------------------------------------------------- jdcolor.c YCbCr -> RGB
def _fix(x: float) -> int:
    return int(x * 65536 + 0.5)


def ycc_to_rgb(y: np.ndarray, cb: np.ndarray, cr: np.ndarray) -> np.ndarray:
    yy = y.astype(np.int32)
    b = cb.astype(np.int32) - 128
    r = cr.astype(np.int32) - 128
    cr_r = (_fix(1.40200) * r + 32768) >> 16
    cb_b = (_fix(1.77200) * b + 32768) >> 16
    g_off = ((-_fix(0.34414)) * b + 32768 + (-_fix(0.71414)) * r) >> 16
    rgb = np.stack([yy + cr_r, yy + g_off, yy + cb_b], axis=-1)
    return np.clip(rgb, 0, 255).astype(np.uint8)


def decode(data: bytes) -> np.ndarray:
    """JPEG bytes -> uint8 [H, W, 3] RGB, the array `np.asarray(Image.open(f).convert('RGB'))` yields."""
    hd = parse(data)
    planes, overrun = entropy_decode(data, hd)
    if overrun:
        raise ValueError("entropy-coded data ended early")
    comp = component_planes(planes, hd)
    H, W = hd.height, hd.width
    if len(comp) == 1:
        g = comp[0][:H, :W]
        return np.stack([g, g, g], axis=-1)
    yh, yv = hd.comps[0].h, hd.comps[0].v
    w_ds, h_ds = (W + yh - 1) // yh, (H + yv - 1) // yv
    if (yh, yv) == (1, 1):
        cb, cr = comp[1], comp[2]
    elif (yh, yv) == (2, 1):
        cb, cr = _h2v1_fancy(comp[1], w_ds), _h2v1_fancy(comp[2], w_ds)
    else:
        cb, cr = _h2v2_fancy(comp[1], w_ds, h_ds), _h2v2_fancy(comp[2], w_ds, h_ds)
    return ycc_to_rgb(comp[0][:H, :W], cb[:H, :W], cr[:H, :W])
