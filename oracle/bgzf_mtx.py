"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's bgzf-indexed
Matrix Market reader (pure Python/numpy; small cases only).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module.  Parity pin: tests/golden/reader_*.npz + *.mtx.gz hold the
outputs of the reference reader itself (oracle/_ref/ref_reader_harness, which
compiles mmutil_index.hh / mmutil_bgzf_util.hh / strbuf.hh / bgzf.c verbatim);
tests/test_oracle_golden.py checks every function here against them.

The bgzf container format itself comes from tabix (vendored in the reference,
src/ext/tabix/bgzf.c, include/ext/tabix/bgzf.h).
"""
from __future__ import annotations

import ctypes
import ctypes.util
import gzip
import struct
import zlib
from typing import Dict, List, Sequence, Tuple

import numpy as np

MISSING_POS = 0  # mmutil_bgzf_util.hh:17
BGZF_BLOCK_SIZE = 0xFF00  # bgzf.h:35 (payload per block used by the writer)

_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
_libm.powf.restype = ctypes.c_float
_libm.powf.argtypes = [ctypes.c_float, ctypes.c_float]

_EOF_BLOCK = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


# --------------------------------------------------------------------------
# bgzf container
# --------------------------------------------------------------------------
def bgzf_compress_block(payload: bytes, level: int = -1) -> bytes:
    """One BGZF member (bgzf.c:193-267): gzip header with the 'BC' extra
    subfield holding BSIZE-1, raw deflate, crc32, isize."""
    co = zlib.compressobj(level, zlib.DEFLATED, -15, 8, zlib.Z_DEFAULT_STRATEGY)
    body = co.compress(payload) + co.flush()
    bsize = 18 + len(body) + 8
    hdr = b"\x1f\x8b\x08\x04" + b"\x00\x00\x00\x00" + b"\x00\xff" + struct.pack("<H", 6)
    hdr += b"BC" + struct.pack("<H", 2) + struct.pack("<H", bsize - 1)
    tail = struct.pack("<II", zlib.crc32(payload) & 0xFFFFFFFF, len(payload) & 0xFFFFFFFF)
    return hdr + body + tail


def bgzf_write(path: str, data: bytes, block_size: int = BGZF_BLOCK_SIZE) -> None:
    """Write `data` as a bgzf file (blocks of `block_size` payload bytes + EOF
    marker block, bgzf.c:487-535)."""
    with open(path, "wb") as f:
        for off in range(0, len(data), block_size):
            f.write(bgzf_compress_block(data[off:off + block_size]))
        f.write(_EOF_BLOCK)


def bgzf_is_bgzf(path: str) -> bool:
    """bgzf.c:581-593: first 16 bytes equal the magic (incl. xlen=6, 'BC', slen=2)."""
    with open(path, "rb") as f:
        b = f.read(16)
    return len(b) == 16 and b == b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00\x42\x43\x02\x00"


class BgzfReader:
    """bgzf_open/seek/tell/getline semantics of bgzf.c:385-424,559-579,626-665."""

    def __init__(self, path: str):
        with open(path, "rb") as f:
            self.raw = f.read()
        self.file_pos = 0          # _bgzf_tell of the underlying file
        self.block_address = 0
        self.block_offset = 0
        self.block_length = 0
        self.block = b""

    def seek(self, voffset: int) -> None:               # bgzf.c:559-579
        self.file_pos = voffset >> 16
        self.block_address = voffset >> 16
        self.block_offset = voffset & 0xFFFF
        self.block_length = 0

    def tell(self) -> int:                              # bgzf.h:124
        return (self.block_address << 16) | (self.block_offset & 0xFFFF)

    def _read_block(self) -> int:                       # bgzf.c:385-424
        addr = self.file_pos
        hdr = self.raw[addr:addr + 18]
        if len(hdr) == 0:
            self.block_length = 0
            return 0
        if len(hdr) != 18 or hdr[:4] != b"\x1f\x8b\x08\x04" or hdr[12:14] != b"BC":
            return -1
        bsize = struct.unpack("<H", hdr[16:18])[0] + 1
        comp = self.raw[addr + 18:addr + bsize - 8]
        payload = zlib.decompress(comp, -15)
        self.file_pos = addr + bsize
        if self.block_length != 0:
            self.block_offset = 0  # not reset if this read follows a seek
        self.block_address = addr
        self.block_length = len(payload)
        self.block = payload
        return 0

    def getline(self) -> Tuple[int, bytes]:             # bgzf.c:626-665
        """Returns (status, line); status < 0 at EOF/error like bgzf_getline."""
        out = bytearray()
        state = 0
        while state == 0:
            if self.block_offset >= self.block_length:
                if self._read_block() != 0:
                    state = -2
                    break
                if self.block_length == 0:
                    state = -1
                    break
            l = self.block.find(b"\n", self.block_offset)
            if l >= 0:
                state = 1
            else:
                l = self.block_length
            out += self.block[self.block_offset:l]
            self.block_offset = l + 1
            if self.block_offset >= self.block_length:
                self.block_address = self.file_pos
                self.block_offset = 0
                self.block_length = 0
        if len(out) == 0 and state < 0:
            return state, b""
        return len(out), bytes(out)


# --------------------------------------------------------------------------
# text -> numbers  (include/utils/strbuf.hh)
# --------------------------------------------------------------------------
def take_uint64(tok: bytes) -> int:
    """strbuf.hh:133-165 (_take_int<ptrdiff_t>): optional '-', leading digits only."""
    i, neg, x = 0, False, 0
    if tok[:1] == b"-":
        neg, i = True, 1
    while i < len(tok) and 48 <= tok[i] <= 57:
        x = x * 10 + (tok[i] - 48)
        i += 1
    return -x if neg else x


def take_float(tok: bytes) -> np.float32:
    """strbuf.hh:72-131.  `float r`; every `r*10.0 + digit` is evaluated in
    double and rounded to float; the fraction is f / pow(10.f, int) in double;
    the exponent multiplies by powf(10.f, e) in float."""
    f32, f64 = np.float32, np.float64
    i, n = 0, len(tok)
    r = f32(0.0)
    neg = False
    if i < n and tok[i] == 45:
        neg, i = True, i + 1
    while i < n and 48 <= tok[i] <= 57:
        r = f32(f64(r) * 10.0 + (tok[i] - 48))
        i += 1
    if i < n and tok[i] == 46:
        f, k = f32(0.0), 0
        i += 1
        while i < n and 48 <= tok[i] <= 57:
            f = f32(f64(f) * 10.0 + (tok[i] - 48))
            i += 1
            k += 1
        r = f32(f64(r) + f64(f) / np.power(f64(10.0), k))
    if i < n and tok[i] in (101, 69):
        i += 1
        e, neg_e = f32(0.0), False
        if i < n and tok[i] == 45:
            neg_e, i = True, i + 1
        elif i < n and tok[i] == 43:
            i += 1
        while i < n and 48 <= tok[i] <= 57:
            e = f32(f64(e) * 10.0 + (tok[i] - 48))
            i += 1
        if neg_e:
            e = f32(-e)
        with np.errstate(over="ignore"):
            r = f32(r * f32(_libm.powf(10.0, float(e))))  # std::pow(float, float) == glibc powf
    return f32(-r) if neg else r


def _isspace(ch: int) -> bool:
    return ch in (32, 9, 10, 11, 12, 13)


def split_triplet_line(line: bytes):
    """Field splitter of mmutil_bgzf_util.hh:111-127 (quirks included): returns
    (num_cols, [tok0, tok1, tok2]) where tok may be b'' (see module tests)."""
    toks: List[bytes] = [b"", b"", b""]
    buf = bytearray()
    num_cols = 0
    word = False  # S_WORD
    for ch in line:
        if _isspace(ch) and len(buf) > 0:
            if num_cols < 3:
                toks[num_cols] = bytes(buf)
            buf.clear()
            word = False
            num_cols += 1
        else:
            if not _isspace(ch):
                buf.append(ch)
            word = True
    if word:
        if num_cols < 3:
            toks[num_cols] = bytes(buf)
        num_cols += 1
    return num_cols, toks


# --------------------------------------------------------------------------
# visitors  (include/mmutil_bgzf_util.hh)
# --------------------------------------------------------------------------
def peek_bgzf_header(path: str) -> Tuple[int, int, int]:
    """mmutil_bgzf_util.hh:155-251 -> (rows, cols, nnz)."""
    fp = BgzfReader(path)
    while True:
        st, line = fp.getline()
        if st < 0:
            raise RuntimeError("Failed to read the header")
        if len(line) < 1 or line[:1] == b"%":
            continue
        n, toks = split_triplet_line(line)
        if n == 3:
            return take_uint64(toks[0]), take_uint64(toks[1]), take_uint64(toks[2])


def visit_bgzf_block(path: str, beg_pos: int, end_pos: int):
    """mmutil_bgzf_util.hh:53-151 -> list of (row0, col0, float32 weight)."""
    fp = BgzfReader(path)
    fp.seek(beg_pos)
    out = []
    while True:
        st, line = fp.getline()
        if st < 0:
            break
        if len(line) < 1 or line[:1] == b"%":
            continue
        n, toks = split_triplet_line(line)
        if n < 3:
            continue
        out.append((take_uint64(toks[0]) - 1, take_uint64(toks[1]) - 1, take_float(toks[2])))
        if end_pos != 0 and fp.tell() >= end_pos:
            break
    return out


def build_index(path: str) -> List[Tuple[int, int]]:
    """mmutil_index.hh:38-107,138-190 -> [(col0, voffset)], file order.
    Raises like the reference fails: unsorted columns / empty trailing columns."""
    if not bgzf_is_bgzf(path):
        raise RuntimeError("This file is not bgzipped")
    fp = BgzfReader(path)
    max_col = None
    while max_col is None:
        st, line = fp.getline()
        if st < 0:
            raise RuntimeError("Failed to read the header")
        if len(line) < 1 or line[:1] == b"%":
            continue
        n, toks = split_triplet_line(line)
        if n == 3:
            max_col = take_uint64(toks[1])
    first_off = last_off = fp.tell()          # eval_after_header :58-59
    col2file: List[Tuple[int, int]] = []
    lineno, last_col = 0, 0
    while True:
        st, line = fp.getline()
        if st < 0:
            break
        if len(line) < 1 or line[:1] == b"%":
            continue
        n, toks = split_triplet_line(line)
        if n < 3:
            continue
        col = take_uint64(toks[1]) - 1
        if lineno == 0:                       # :68-71
            last_col = col
            col2file.append((col, first_off))
        if col != last_col:                   # :73-82
            if not col > last_col:
                raise RuntimeError("MTX must be sorted by columns")
            col2file.append((col, last_off))
            last_col = col
        last_off = fp.tell()                  # :83
        lineno += 1
    last = col2file[-1][0] if col2file else 0
    if last != max_col - 1:                   # :171-179
        raise RuntimeError("Failed to index all the columns")
    return col2file


def write_index(path: str, col2file: Sequence[Tuple[int, int]]) -> None:
    """mmutil_index.hh:183-185 / io.hh:248-266: gzip text 'col voffset\\n'."""
    with gzip.open(path, "wb") as f:
        for c, o in col2file:
            f.write(f"{c} {o}\n".encode())


def read_index(path: str) -> np.ndarray:
    """mmutil_index.hh:192-228 -> int64[MaxIdx+1] with the gap-fill quirk."""
    pairs = []
    with gzip.open(path, "rb") as f:
        toks = f.read().split()
    for i in range(0, len(toks) - 1, 2):
        pairs.append((int(toks[i]), int(toks[i + 1])))
    if not pairs:
        raise RuntimeError("empty file")
    max_idx = max(max(p[0] for p in pairs), 0)
    tab = np.full(max_idx + 1, MISSING_POS, dtype=np.int64)
    for c, o in pairs:
        tab[c] = o
    for j in range(0, max_idx - 1):           # :219-224 (stops at MaxIdx-1, forward order)
        if tab[j] == MISSING_POS:
            tab[j] = tab[j + 1]
    return tab


# --------------------------------------------------------------------------
# minibatch data block  (include/mmvae_io.hh)
# --------------------------------------------------------------------------
def find_consecutive_blocks(tab: np.ndarray, subcol: Sequence[int], gap: int = 10):
    """mmvae_io.hh:150-204 -> [(lb, lb_mem, ub, ub_mem)]."""
    N = len(tab)
    s = sorted(int(v) for v in subcol)
    iv = []
    beg = end = s[0]
    for ii in s[1:]:
        if ii >= end + gap:
            iv.append((beg, end + 1))
            beg = end = ii
        else:
            end = ii
    if beg <= s[-1]:
        iv.append((beg, end + 1))
    out = []
    for lb, ub in iv:
        if lb >= N:
            continue
        ub_mem = int(tab[ub]) if ub < N else 0
        out.append((lb, int(tab[lb]), ub, ub_mem))
    return out


def read_batch_dense(path: str, tab: np.ndarray, subcol: Sequence[int], D: int) -> np.ndarray:
    """mtx_data_block_t::read (mmvae_io.hh:208-245) -> dense float32[B, D];
    duplicates replicated through `dup`, last write wins."""
    B = len(subcol)
    mem = np.zeros((B, D), dtype=np.float32)
    col2idx: Dict[int, int] = {}
    for cc in sorted(int(v) for v in subcol):
        if cc not in col2idx:
            col2idx[cc] = len(col2idx)
    dup: List[List[int]] = [[] for _ in range(B)]
    for j, cc in enumerate(subcol):
        dup[col2idx[int(cc)]].append(j)
    for lb, lb_mem, ub, ub_mem in find_consecutive_blocks(tab, subcol):
        for r, col, w in visit_bgzf_block(path, lb_mem, ub_mem):
            k = col2idx.get(col)
            if k is None:
                continue
            for j in dup[k]:
                mem[j, r] = w
    return mem
