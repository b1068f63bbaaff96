"""ctypes mirror of include/bvc_host.h: the host side of the `--rerun` drop-in (C++: csrc/host).

`Packer` turns batch-file lines into a tile exactly as the reference's window loop tokenises them
(src/BaseVarC.cpp:393-441); `format_cvg` / `format_vcf` print a result record as bt_f / WriteVcf do;
`run_basetype(argv)` is the whole `BaseVarC basetype --rerun` command in-process and `BIN_PATH` the
stand-alone executable.  Nothing here computes a statistic, and there is no CPU calling path.
"""
import ctypes as C
import os

import numpy as np

from ._abi import SITE_RESULT_DTYPE, Tile
from .caller import BvcError, load_library
from .pileup import PileupTile

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "libbvc_host.so")
BIN_PATH = os.path.join(_HERE, "bin", "BaseVarC_b200")

HOST_EXPORTS = (
    "bvc_packer_create", "bvc_packer_destroy", "bvc_packer_clear", "bvc_packer_begin_site", "bvc_packer_add_line",
    "bvc_packer_end_site", "bvc_packer_n_sites", "bvc_packer_n_entries", "bvc_packer_tile", "bvc_packer_positions",
    "bvc_packer_indel", "bvc_packer_error", "bvc_format_cvg", "bvc_format_vcf", "bvc_format_fixed", "bvc_basetype_main", "bvc_write_batch_files",
    "bvc_packed_open", "bvc_packed_next", "bvc_packed_names", "bvc_packed_error", "bvc_packed_close", "bvc_packed_is_complete",
    "bvc_write_packed_files", "bvc_fast_inflate", "bvc_fast_crc32",
)

_hlib = None


def load_host_library():
    global _hlib
    if _hlib is not None:
        return _hlib
    load_library()                       # libbvc_gpu.so first (also found through the host library's rpath)
    if not os.path.exists(HOST_LIB_PATH):
        raise BvcError("host library %s is missing: run __graft_entry__.build()" % HOST_LIB_PATH)
    L = C.CDLL(HOST_LIB_PATH)
    L.bvc_packer_create.argtypes = [C.c_int32]
    L.bvc_packer_create.restype = C.c_void_p
    L.bvc_packer_destroy.argtypes = [C.c_void_p]
    L.bvc_packer_destroy.restype = None
    L.bvc_packer_clear.argtypes = [C.c_void_p]
    L.bvc_packer_clear.restype = None
    L.bvc_packer_begin_site.argtypes = [C.c_void_p, C.c_int32, C.c_uint8]
    L.bvc_packer_begin_site.restype = None
    L.bvc_packer_add_line.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
    L.bvc_packer_end_site.argtypes = [C.c_void_p]
    L.bvc_packer_end_site.restype = C.c_int32
    L.bvc_packer_n_sites.argtypes = [C.c_void_p]
    L.bvc_packer_n_sites.restype = C.c_int32
    L.bvc_packer_n_entries.argtypes = [C.c_void_p]
    L.bvc_packer_n_entries.restype = C.c_int64
    L.bvc_packer_tile.argtypes = [C.c_void_p, C.POINTER(Tile)]
    L.bvc_packer_tile.restype = None
    L.bvc_packer_positions.argtypes = [C.c_void_p]
    L.bvc_packer_positions.restype = C.POINTER(C.c_int32)
    L.bvc_packer_indel.argtypes = [C.c_void_p, C.c_int64]
    L.bvc_packer_indel.restype = C.c_char_p
    L.bvc_packer_error.argtypes = [C.c_void_p]
    L.bvc_packer_error.restype = C.c_char_p
    L.bvc_format_cvg.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_int32, C.POINTER(Tile), C.c_int32, C.c_void_p,
                                 C.c_void_p, C.c_int32, C.c_void_p]
    L.bvc_format_cvg.restype = C.c_size_t
    L.bvc_format_vcf.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_int32, C.POINTER(Tile), C.c_int32, C.c_void_p,
                                 C.c_int32, C.c_int32, C.POINTER(C.c_char_p), C.c_void_p, C.c_void_p]
    L.bvc_format_vcf.restype = C.c_size_t
    L.bvc_format_fixed.argtypes = [C.c_char_p, C.c_size_t, C.c_double, C.c_int32]
    L.bvc_format_fixed.restype = C.c_size_t
    L.bvc_basetype_main.argtypes = [C.c_int, C.POINTER(C.c_char_p)]
    L.bvc_write_batch_files.argtypes = [C.POINTER(Tile), C.c_int32, C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_char_p]
    L.bvc_packed_open.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64)]
    L.bvc_packed_open.restype = C.c_void_p
    L.bvc_packed_next.argtypes = [C.c_void_p, C.c_void_p]
    L.bvc_packed_names.argtypes = [C.c_void_p]
    L.bvc_packed_names.restype = C.c_char_p
    L.bvc_packed_error.argtypes = [C.c_void_p]
    L.bvc_packed_error.restype = C.c_char_p
    L.bvc_packed_close.argtypes = [C.c_void_p]
    L.bvc_packed_close.restype = None
    L.bvc_packed_is_complete.argtypes = [C.c_char_p, C.c_int32]
    L.bvc_fast_inflate.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.bvc_fast_crc32.argtypes = [C.c_char_p, C.c_size_t]
    L.bvc_fast_crc32.restype = C.c_uint32
    L.bvc_write_packed_files.argtypes = [C.POINTER(Tile), C.c_int32, C.c_void_p, C.c_char_p, C.c_int32, C.c_int32, C.c_char_p]
    _hlib = L
    return L


class Packer:
    """bvc_packer: one per window; keeps the reference's stale-token state across sites and clear()."""

    def __init__(self, n_samples):
        self.lib = load_host_library()
        self.n_samples = int(n_samples)
        self.h = self.lib.bvc_packer_create(self.n_samples)
        if not self.h:
            raise BvcError("bvc_packer_create failed")

    def close(self):
        if getattr(self, "h", None):
            self.lib.bvc_packer_destroy(self.h)
            self.h = None

    __del__ = close

    def clear(self):
        self.lib.bvc_packer_clear(self.h)

    def add_site(self, pos, ref_base, lines):
        """One site from its batch-file lines (bytes, one per batch file).  Returns the entry count."""
        self.lib.bvc_packer_begin_site(self.h, int(pos), int(ref_base))
        for ln in lines:
            if self.lib.bvc_packer_add_line(self.h, ln, len(ln)) != 0:
                raise BvcError("bvc_packer_add_line: %s" % self.lib.bvc_packer_error(self.h).decode())
        return self.lib.bvc_packer_end_site(self.h)

    def ctile(self):
        t = Tile()
        self.lib.bvc_packer_tile(self.h, C.byref(t))
        return t

    def to_tile(self):
        """Copy the packed arrays out as a PileupTile (indel texts included)."""
        t = self.ctile()
        S, R = t.n_sites, self.lib.bvc_packer_n_entries(self.h)

        def arr(ptr, n, dt):
            if n == 0:
                return np.zeros(0, dt)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), (n * np.dtype(dt).itemsize,)).view(dt).copy()

        pos = np.ctypeslib.as_array(self.lib.bvc_packer_positions(self.h), (S,)).copy() if S else np.zeros(0, np.int32)
        tile = PileupTile(self.n_samples, arr(t.ref_base, S, np.uint8), arr(t.entry_off, S + 1, np.uint32),
                          arr(t.code, R, np.uint8), arr(t.qual, R, np.uint8), arr(t.mapq, R, np.uint8),
                          arr(t.rpr, R, np.uint8), arr(t.sample, R, np.uint32), pos)
        if t.packed8:
            tile.packed8 = arr(t.packed8, R, np.uint8)
            tile.qual_dict = arr(t.qual_dict, t.n_qual_dict, np.uint8)
        for e in np.nonzero((tile.code >> 4) & 1)[0]:
            s = self.lib.bvc_packer_indel(self.h, int(e))
            if s is not None:
                tile.indel_text[int(e)] = s.decode()
        return tile


def _fmt(call):
    buf = C.create_string_buffer(1 << 16)
    n = call(buf, len(buf))
    if n >= len(buf):
        buf = C.create_string_buffer(n + 1)
        call(buf, len(buf))
    return buf.value.decode()


def format_fixed(v, digits):
    """`v` as the reference's "{:.Nf}" prints it (fmt 5.3.1, tiny-value quirk included)."""
    L = load_host_library()
    b = C.create_string_buffer(96)
    n = L.bvc_format_fixed(b, 96, float(v), int(digits))
    return b.raw[:n].decode()


def format_cvg(chrom, pos, ctile, site, rec, packer=None, group_depth=None):
    L = load_host_library()
    rec = np.ascontiguousarray(rec, SITE_RESULT_DTYPE).reshape(1)
    gd = None if group_depth is None else np.ascontiguousarray(group_depth, np.int32)
    return _fmt(lambda b, n: L.bvc_format_cvg(b, n, chrom.encode(), int(pos), C.byref(ctile), int(site), rec.ctypes.data,
                                              packer.h if packer is not None else None,
                                              0 if gd is None else gd.shape[0], None if gd is None else gd.ctypes.data))


def format_vcf(chrom, pos, ctile, site, rec, n_samples, group_names=(), group_depth=None, group_af=None):
    L = load_host_library()
    rec = np.ascontiguousarray(rec, SITE_RESULT_DTYPE).reshape(1)
    names = (C.c_char_p * max(len(group_names), 1))(*[g.encode() for g in group_names])
    gd = None if group_depth is None else np.ascontiguousarray(group_depth, np.int32)
    ga = None if group_af is None else np.ascontiguousarray(group_af, np.float64)
    return _fmt(lambda b, n: L.bvc_format_vcf(b, n, chrom.encode(), int(pos), C.byref(ctile), int(site), rec.ctypes.data,
                                              int(n_samples), len(group_names), names,
                                              None if gd is None else gd.ctypes.data, None if ga is None else ga.ctypes.data))


def run_basetype(argv):
    """bvc_basetype_main in-process; argv = ["basetype", "--rerun", ...].  Returns the exit status."""
    L = load_host_library()
    arr = (C.c_char_p * (len(argv) + 1))(*[a.encode() for a in argv], None)
    return L.bvc_basetype_main(len(argv), arr)


def write_batch_files(tile, prefix, threads, batch, level=1, indel_token="+A"):
    """Fast (C++, one thread per window) writer of the reference's batch files; same files as
    batchfile.write_batch_files for tiles whose indel entries all print as `indel_token`."""
    L = load_host_library()
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    ct = tile.as_ctypes(with_sample=True)
    rc = L.bvc_write_batch_files(C.byref(ct), tile.n_samples, prefix.encode(), int(threads), int(batch), int(level),
                                 indel_token.encode())
    if rc != 0:
        raise BvcError("bvc_write_batch_files failed (%d): bad (sites, threads) partition or unwritable prefix" % rc)


def fast_inflate(raw, out_cap=65536):
    """bvc_fast_inflate on a raw deflate stream: the inflated bytes, or None when the decoder declines the stream."""
    L = load_host_library()
    out = C.create_string_buffer(max(out_cap, 1))
    n = C.c_size_t(0)
    ok = L.bvc_fast_inflate(raw, len(raw), out, out_cap, C.byref(n))
    return out.raw[:n.value] if ok else None


def fast_crc32(data):
    return int(load_host_library().bvc_fast_crc32(data, len(data)))


def write_packed_files(tile, prefix, threads, tile_sites=2048, indel_token="+A"):
    """bvc_write_packed_files: a tile as packed tile files (include/bvc_host.h), one per window; sites numbered by tile.pos."""
    L = load_host_library()
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    ct = tile.as_ctypes(with_sample=True)
    pos = np.ascontiguousarray(tile.pos, np.int32)
    rc = L.bvc_write_packed_files(C.byref(ct), tile.n_samples, pos.ctypes.data, prefix.encode(), int(threads), int(tile_sites),
                                  indel_token.encode())
    if rc != 0:
        raise BvcError("bvc_write_packed_files failed (%d)" % rc)


def read_packed_file(path):
    """Every tile of a packed tile file: (sample names, [(positions, PileupTile, {entry: indel text})])."""
    L = load_host_library()
    n, nt, ms, me = C.c_int32(0), C.c_int32(0), C.c_int32(0), C.c_int64(0)
    r = L.bvc_packed_open(path.encode(), C.byref(n), C.byref(nt), C.byref(ms), C.byref(me))
    if not r:
        raise BvcError("cannot open packed tile file %s" % path)
    names = L.bvc_packed_names(r).decode()
    pk = Packer(n.value)
    out = []
    try:
        while True:
            rc = L.bvc_packed_next(r, pk.h)
            if rc == 0:
                break
            if rc < 0:
                raise BvcError(L.bvc_packed_error(r).decode())
            t = pk.to_tile()
            out.append((np.array(t.pos, copy=True), t, dict(t.indel_text)))
    finally:
        L.bvc_packed_close(r)
        pk.close()
    return names, out
