"""Host-side mirror of the C-ABI (include/bvc_gpu.h) through ctypes.

`BaseTypeCaller.call_sites(tile)` is the tile-granular replacement of the reference's per-site
`bt_f(...)` call (src/BaseVarC.cpp:442-443): same inputs (the site's AlleleInfo entries), same
numeric outputs (what bt_f / WriteVcf format into the CVG / VCF text).  There is NO CPU fallback:
a missing library or a missing GPU raises `BvcError`.
"""
import ctypes as C
import os

import numpy as np

from ._abi import CODED_CHUNK, CodedTile, SITE_CORE_DTYPE, SITE_RESULT_DTYPE, Params, Tile, default_params

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbvc_gpu.so")

EXPORTS = (
    "bvc_device_count", "bvc_default_params", "bvc_create", "bvc_destroy", "bvc_last_error", "bvc_enqueue", "bvc_wait",
    "bvc_enqueue_split", "bvc_wait_split",
    "bvc_fetch_groups", "bvc_host_alloc", "bvc_host_free", "bvc_host_register", "bvc_host_unregister", "bvc_resident_upload", "bvc_resident_run",
    "bvc_resident_download", "bvc_kernel_launch_count", "bvc_measure_fp64_peak", "bvc_kernel_name", "bvc_transfer_bytes", "bvc_work_counters",
    "bvc_shard_unique_id", "bvc_shard_init", "bvc_shard_call", "bvc_shard_upload", "bvc_shard_run", "bvc_shard_init_group",
    "bvc_shard_call_group", "bvc_shard_stats", "bvc_shard_phase_name", "bvc_call_sites_noalloc",
    "bvc_coded_words", "bvc_encode_codes", "bvc_decode_codes_host", "bvc_enqueue_split_coded",
)
NOALLOC_MAX_SITES = 65536       # include/bvc_gpu.h BVC_NOALLOC_*
NOALLOC_MAX_SAMPLES = 16384
SHARD_ID_BYTES = 128
SHARD_N_PHASES = 9


class BvcError(RuntimeError):
    pass


_lib = None


def load_library():
    """dlopen basevarc_b200/libbvc_gpu.so (built in-tree by `make -C basevarc_b200/csrc`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise BvcError("CUDA extension %s is missing: run __graft_entry__.build() (no CPU fallback exists)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.bvc_default_params.argtypes = [C.POINTER(Params), C.c_int32, C.c_double]
    L.bvc_default_params.restype = None
    L.bvc_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(Params)]
    L.bvc_destroy.argtypes = [C.c_void_p]
    L.bvc_destroy.restype = None
    L.bvc_last_error.argtypes = [C.c_void_p]
    L.bvc_last_error.restype = C.c_char_p
    L.bvc_enqueue.argtypes = [C.c_void_p, C.POINTER(Tile), C.c_void_p, C.POINTER(C.c_uint64)]
    L.bvc_wait.argtypes = [C.c_void_p, C.c_uint64]
    L.bvc_enqueue_split.argtypes = [C.c_void_p, C.POINTER(Tile), C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]
    L.bvc_wait_split.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_int32)]
    L.bvc_fetch_groups.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    L.bvc_host_alloc.argtypes = [C.c_size_t]
    L.bvc_host_alloc.restype = C.c_void_p
    L.bvc_host_free.argtypes = [C.c_void_p]
    L.bvc_host_free.restype = None
    L.bvc_resident_upload.argtypes = [C.c_void_p, C.POINTER(Tile)]
    L.bvc_resident_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_int)]
    L.bvc_resident_download.argtypes = [C.c_void_p, C.c_void_p]
    L.bvc_kernel_launch_count.argtypes = [C.c_void_p]
    L.bvc_kernel_launch_count.restype = C.c_int64
    L.bvc_measure_fp64_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
    L.bvc_kernel_name.argtypes = [C.c_int]
    L.bvc_kernel_name.restype = C.c_char_p
    L.bvc_transfer_bytes.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.bvc_transfer_bytes.restype = None
    L.bvc_work_counters.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    L.bvc_shard_unique_id.argtypes = [C.c_void_p]
    L.bvc_shard_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
    L.bvc_shard_call.argtypes = [C.c_void_p, C.POINTER(Tile), C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.bvc_shard_upload.argtypes = [C.c_void_p, C.POINTER(Tile)]
    L.bvc_shard_run.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.bvc_shard_init_group.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int]
    L.bvc_shard_call_group.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.POINTER(Tile)), C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.bvc_shard_stats.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
    L.bvc_shard_phase_name.argtypes = [C.c_int]
    L.bvc_shard_phase_name.restype = C.c_char_p
    L.bvc_coded_words.argtypes = [C.c_int64, C.c_int]
    L.bvc_coded_words.restype = C.c_size_t
    L.bvc_encode_codes.argtypes = [C.POINTER(Tile), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(CodedTile)]
    L.bvc_decode_codes_host.argtypes = [C.POINTER(CodedTile), C.POINTER(Tile), C.c_void_p]
    L.bvc_enqueue_split_coded.argtypes = [C.c_void_p, C.POINTER(Tile), C.POINTER(CodedTile), C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]
    L.bvc_call_sites_noalloc.argtypes = [C.c_int, C.POINTER(Params), C.POINTER(Tile), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    _lib = L
    return L


def call_sites_noalloc(params, n_sites, n_entries, d_ref_base, d_entry_off, d_code, d_qual, d_mapq, d_rpr, d_out, d_status=None,
                       stream=None, device=0):
    """bvc_call_sites_noalloc (include/bvc_gpu.h): every d_* is a DEVICE address (int); enqueues on `stream` and returns."""
    L = load_library()
    t = Tile()
    t.n_sites = int(n_sites)
    t.ref_base, t.entry_off, t.code, t.qual, t.mapq, t.rpr = d_ref_base, d_entry_off, d_code, d_qual, d_mapq, d_rpr
    rc = L.bvc_call_sites_noalloc(int(device), C.byref(params), C.byref(t), int(n_entries), d_out, d_status, stream)
    if rc != 0:
        raise BvcError("bvc_call_sites_noalloc failed (%d): %s" % (rc, L.bvc_last_error(None).decode()))


def measure_fp64_peak(device=0):
    L = load_library()
    t = C.c_double(0.0)
    rc = L.bvc_measure_fp64_peak(device, C.byref(t))
    if rc != 0:
        raise BvcError("bvc_measure_fp64_peak: %s" % L.bvc_last_error(None).decode())
    return t.value


class CodedArrays:
    """A tile's packed8 bytes in the coded wire form (include/bvc_gpu.h): the bvc_coded_tile + the arrays it points into.
    `alloc(shape, dtype)` supplies the arrays (numpy by default; bench.py passes a pinned allocator)."""

    def __init__(self, tile, alloc=None):
        L = load_library()
        assert tile.packed8 is not None, "tile.pack8() first"
        n = tile.n_entries
        mk = alloc or (lambda shape, dtype: np.zeros(shape, dtype))
        self._keep = [mk((max(int(L.bvc_coded_words(n, 6)), 1),), np.uint32), mk((n // 8 + 16,), np.uint8), mk((n // CODED_CHUNK + 2,), np.uint32)]
        arrs = [getattr(k, "array", k) for k in self._keep]
        self.words, self.escapes, self.chunk_escapes = arrs
        self.ct = CodedTile()
        self._tile_ct = tile.as_ctypes(with_sample=False)
        self._tile = tile
        rc = L.bvc_encode_codes(C.byref(self._tile_ct), self.words.ctypes.data, self.escapes.ctypes.data, self.chunk_escapes.ctypes.data, C.byref(self.ct))
        if rc != 0:
            raise BvcError("bvc_encode_codes failed (%d)" % rc)
        self.n = n

    @property
    def bits(self):
        return int(self.ct.bits)

    @property
    def wire_bytes(self):
        """Bytes this form puts on the wire for the (code, qual) information of the tile."""
        L = load_library()
        return int(L.bvc_coded_words(self.n, self.bits)) * 4 + int(self.ct.n_escape) + (self.n // CODED_CHUNK + 2) * 4 if self.bits else self.n

    def decode_host(self):
        out = np.zeros(self.n, np.uint8)
        rc = load_library().bvc_decode_codes_host(C.byref(self.ct), C.byref(self._tile_ct), out.ctypes.data)
        if rc != 0:
            raise BvcError("bvc_decode_codes_host failed (%d)" % rc)
        return out


class PinnedArray:
    """A numpy view over cudaHostAlloc memory (freed with the object)."""

    def __init__(self, shape, dtype):
        L = load_library()
        self.dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * self.dtype.itemsize
        self.ptr = L.bvc_host_alloc(max(n, 1))
        if not self.ptr:
            raise BvcError("bvc_host_alloc(%d) failed" % n)
        buf = (C.c_char * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(shape))).reshape(shape)

    def __del__(self):
        try:
            if getattr(self, "ptr", None):
                load_library().bvc_host_free(self.ptr)
                self.ptr = None
        except Exception:
            pass


class BaseTypeCaller:
    """One context = one GPU.  Not thread-safe (as the C ABI)."""

    def __init__(self, n_samples, maf=0.001, device=0, max_sites=1 << 16, max_entries=None, sample_group=None,
                 n_groups=0, lrt_threshold=None, em_epsilon=None):
        self.lib = load_library()
        p = default_params(n_samples, maf, max_sites, 0)
        if lrt_threshold is not None:              # the reference's constants (src/BaseType.h:9, src/BaseType.cpp:45): overridden by tests only
            p.lrt_threshold = float(lrt_threshold)
        if em_epsilon is not None:
            p.em_epsilon = float(em_epsilon)
        p.max_entries = int(max_entries if max_entries is not None else min(max_sites * n_samples, (1 << 32) - 1))
        self._sg = None
        if n_groups > 0:
            self._sg = np.ascontiguousarray(np.asarray(sample_group, np.int32))
            assert self._sg.size == n_samples
            p.n_groups = int(n_groups)
            p.sample_group = self._sg.ctypes.data_as(C.POINTER(C.c_int32))
        self.params = p
        self.ctx = C.c_void_p()
        rc = self.lib.bvc_create(C.byref(self.ctx), int(device), C.byref(p))
        if rc != 0:
            raise BvcError("bvc_create failed (%d): %s" % (rc, self.lib.bvc_last_error(None).decode()))

    def close(self):
        if getattr(self, "ctx", None) and self.ctx.value:
            self.lib.bvc_destroy(self.ctx)
            self.ctx = C.c_void_p()

    __del__ = close

    def _check(self, rc, what):
        if rc != 0:
            raise BvcError("%s failed (%d): %s" % (what, rc, self.lib.bvc_last_error(self.ctx).decode()))

    # -- reference-facing call: host buffers in, host records out ------------------------------------
    def enqueue(self, tile, out):
        ct = tile.as_ctypes(with_sample=self.params.n_groups > 0, drop_code_qual=True)
        tk = C.c_uint64(0)
        self._check(self.lib.bvc_enqueue(self.ctx, C.byref(ct), out.ctypes.data, C.byref(tk)), "bvc_enqueue")
        return tk.value

    def wait(self, ticket):
        self._check(self.lib.bvc_wait(self.ctx, C.c_uint64(ticket)), "bvc_wait")

    def enqueue_split(self, tile, core, vcf):
        """bvc_enqueue_split: `core` [n_sites] SITE_CORE_DTYPE, `vcf` [cap] SITE_RESULT_DTYPE (both must outlive wait_split)."""
        ct = tile.as_ctypes(with_sample=self.params.n_groups > 0, drop_code_qual=True)
        tk = C.c_uint64(0)
        self._check(self.lib.bvc_enqueue_split(self.ctx, C.byref(ct), core.ctypes.data, vcf.ctypes.data if vcf.shape[0] else None,
                                               int(vcf.shape[0]), C.byref(tk)), "bvc_enqueue_split")
        return tk.value

    def enqueue_split_coded(self, tile, coded, core, vcf):
        """bvc_enqueue_split_coded: `tile` carries qual_dict (tile.pack8() done) and `coded` = CodedArrays(tile)."""
        ct = tile.as_ctypes(with_sample=False, drop_code_qual=True)
        ct.packed8 = None
        tk = C.c_uint64(0)
        self._check(self.lib.bvc_enqueue_split_coded(self.ctx, C.byref(ct), C.byref(coded.ct), core.ctypes.data, vcf.ctypes.data if vcf.shape[0] else None,
                                                     int(vcf.shape[0]), C.byref(tk)), "bvc_enqueue_split_coded")
        return tk.value

    def wait_split(self, ticket):
        """Number of LRT-positive sites of the tile = valid entries of the vcf array handed to enqueue_split."""
        n = C.c_int32(0)
        self._check(self.lib.bvc_wait_split(self.ctx, C.c_uint64(ticket), C.byref(n)), "bvc_wait_split")
        return n.value

    def call_sites_split(self, tile, vcf_cap=None):
        """(core records, full records of the LRT-positive sites indexed by core['vcf_index'])."""
        core = np.zeros(tile.n_sites, SITE_CORE_DTYPE)
        vcf = np.zeros(tile.n_sites if vcf_cap is None else vcf_cap, SITE_RESULT_DTYPE)
        n = self.wait_split(self.enqueue_split(tile, core, vcf))
        return core, vcf[:n]

    def call_sites_with_groups(self, tile):
        """(records, group_depth[S][G][4], group_af[S][G][4]) -- bt_f with a --group file (src/BaseVarC.cpp:617-661)."""
        out = np.zeros(tile.n_sites, SITE_RESULT_DTYPE)
        G = self.params.n_groups
        gd = np.zeros((tile.n_sites, G, 4), np.int32)
        ga = np.zeros((tile.n_sites, G, 4), np.float64)
        tk = self.enqueue(tile, out)
        self.wait(tk)
        self._check(self.lib.bvc_fetch_groups(self.ctx, C.c_uint64(tk), gd.ctypes.data, ga.ctypes.data), "bvc_fetch_groups")
        return out, gd, ga

    def call_sites(self, tile, out=None):
        if out is None:
            out = np.zeros(tile.n_sites, SITE_RESULT_DTYPE)
        self.wait(self.enqueue(tile, out))
        return out

    # -- device-resident path (kernel-only timing) ---------------------------------------------------
    def resident_upload(self, tile):
        ct = tile.as_ctypes(with_sample=self.params.n_groups > 0, drop_code_qual=True)
        self._check(self.lib.bvc_resident_upload(self.ctx, C.byref(ct)), "bvc_resident_upload")
        self._resident_sites = tile.n_sites

    def resident_run(self, iters=1, flush_l2=True, per_kernel=False):
        """Mean device ms per pass; with per_kernel=True also {kernel name: ms per pass} (events between launches)."""
        ms = C.c_float(0)
        per = (C.c_float * 16)()
        nk = C.c_int(0)
        self._check(self.lib.bvc_resident_run(self.ctx, int(iters), int(bool(flush_l2)), C.byref(ms),
                                              per if per_kernel else None, C.byref(nk)), "bvc_resident_run")
        names = [self.lib.bvc_kernel_name(i).decode() for i in range(nk.value)]
        return ms.value, ({n: per[i] for i, n in enumerate(names)} if per_kernel else {})

    def work_counters(self):
        out = (C.c_uint64 * 4)()
        self._check(self.lib.bvc_work_counters(self.ctx, out), "bvc_work_counters")
        return dict(em_class_passes=out[0], em_long_class_passes=out[1], loglik_class_evals=out[2], em_fits=out[3])

    @property
    def transfer_bytes(self):
        h, d = C.c_int64(0), C.c_int64(0)
        self.lib.bvc_transfer_bytes(self.ctx, C.byref(h), C.byref(d))
        return h.value, d.value

    def resident_download(self):
        out = np.zeros(self._resident_sites, SITE_RESULT_DTYPE)
        self._check(self.lib.bvc_resident_download(self.ctx, out.ctypes.data), "bvc_resident_download")
        return out

    @property
    def kernel_launches(self):
        return int(self.lib.bvc_kernel_launch_count(self.ctx))


# ---- sample-sharded mode (include/bvc_gpu.h "Sample-sharded mode") ------------------------------------
def shard_unique_id():
    """ncclGetUniqueId as bytes (rank 0 calls it; broadcast the bytes to the other ranks)."""
    L = load_library()
    buf = C.create_string_buffer(SHARD_ID_BYTES)
    rc = L.bvc_shard_unique_id(buf)
    if rc != 0:
        raise BvcError("bvc_shard_unique_id failed (%d): %s" % (rc, L.bvc_last_error(None).decode()))
    return buf.raw


class ShardedCaller(BaseTypeCaller):
    """Rank `rank` of `world`: holds the entries of its own sample range for every site of a tile and returns the
    records of its site slice.  n_samples is the size of the WHOLE cohort.  Every call is collective (NCCL)."""

    def __init__(self, n_samples, rank, world, unique_id, qual_lo=1, qual_hi=63, **kw):
        super().__init__(n_samples, **kw)
        self.rank, self.world = int(rank), int(world)
        assert len(unique_id) == SHARD_ID_BYTES
        self._id = C.create_string_buffer(bytes(unique_id), SHARD_ID_BYTES)
        self._check(self.lib.bvc_shard_init(self.ctx, self.rank, self.world, self._id, int(qual_lo), int(qual_hi)), "bvc_shard_init")

    def shard_call(self, tile, out=None):
        """(first_site, records of the owned slice) through host buffers."""
        if out is None:
            out = np.zeros((tile.n_sites + self.world - 1) // self.world, SITE_RESULT_DTYPE)
        ct = tile.as_ctypes(with_sample=False, drop_code_qual=True)
        self._shard_sites = tile.n_sites
        first, n = C.c_int32(0), C.c_int32(0)
        self._check(self.lib.bvc_shard_call(self.ctx, C.byref(ct), out.ctypes.data, C.byref(first), C.byref(n)), "bvc_shard_call")
        return first.value, out[:n.value]

    def shard_upload(self, tile):
        ct = tile.as_ctypes(with_sample=False, drop_code_qual=True)
        self._check(self.lib.bvc_shard_upload(self.ctx, C.byref(ct)), "bvc_shard_upload")
        self._shard_sites = tile.n_sites

    def shard_run(self, out=None):
        if out is None:
            out = np.zeros((self._shard_sites + self.world - 1) // self.world, SITE_RESULT_DTYPE)
        first, n = C.c_int32(0), C.c_int32(0)
        self._check(self.lib.bvc_shard_run(self.ctx, out.ctypes.data, C.byref(first), C.byref(n)), "bvc_shard_run")
        return first.value, out[:n.value]

    def shard_stats(self):
        return _shard_stats(self.lib, self.ctx)


def _shard_stats(lib, ctx):
    ms = (C.c_float * SHARD_N_PHASES)()
    xb, npos = C.c_int64(0), C.c_int32(0)
    rc = lib.bvc_shard_stats(ctx, ms, C.byref(xb), C.byref(npos))
    if rc != 0:
        raise BvcError("bvc_shard_stats failed (%d)" % rc)
    return dict(phase_ms={lib.bvc_shard_phase_name(i).decode(): float(ms[i]) for i in range(SHARD_N_PHASES)},
                exchange_bytes=xb.value, n_positive=npos.value)


class ShardGroup:
    """W contexts of THIS process on one device with the in-process exchange (no NCCL): the same kernels and phases
    as the NCCL path, used to parity-test sample sharding on a single GPU."""

    def __init__(self, n_samples, world, qual_lo=1, qual_hi=63, **kw):
        self.world = int(world)
        self.callers = [BaseTypeCaller(n_samples, **kw) for _ in range(self.world)]
        self.lib = self.callers[0].lib
        self._ctxs = (C.c_void_p * self.world)(*[c.ctx for c in self.callers])
        rc = self.lib.bvc_shard_init_group(self._ctxs, self.world, int(qual_lo), int(qual_hi))
        if rc != 0:
            raise BvcError("bvc_shard_init_group failed (%d): %s" % (rc, self._errors()))

    def _errors(self):
        return " | ".join(self.lib.bvc_last_error(c.ctx).decode() for c in self.callers)

    def call_sites(self, tiles):
        """tiles[r] = rank r's local tile (same sites, its samples' entries) -> the records of ALL sites, in order."""
        assert len(tiles) == self.world
        S = tiles[0].n_sites
        sr = (S + self.world - 1) // self.world
        outs = [np.zeros(sr, SITE_RESULT_DTYPE) for _ in range(self.world)]
        cts = [t.as_ctypes(with_sample=False, drop_code_qual=True) for t in tiles]
        tp = (C.POINTER(Tile) * self.world)(*[C.pointer(ct) for ct in cts])
        op = (C.c_void_p * self.world)(*[o.ctypes.data for o in outs])
        first = (C.c_int32 * self.world)()
        n = (C.c_int32 * self.world)()
        rc = self.lib.bvc_shard_call_group(self._ctxs, self.world, tp, op, first, n)
        if rc != 0:
            raise BvcError("bvc_shard_call_group failed (%d): %s" % (rc, self._errors()))
        assert [first[r] for r in range(self.world)] == [r * sr for r in range(self.world)]
        return np.concatenate([outs[r][:n[r]] for r in range(self.world)])

    def stats(self, rank=0):
        return _shard_stats(self.lib, self.callers[rank].ctx)

    @property
    def kernel_launches(self):
        return sum(c.kernel_launches for c in self.callers)

    def close(self):
        for c in self.callers:
            c.close()
