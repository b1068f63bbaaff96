"""ctypes / numpy mirrors of include/bvc_gpu.h (the C-ABI boundary).

Nothing here computes: it only describes memory layouts.  `SITE_RESULT_DTYPE` is the 176-byte
`bvc_site_result` record; `Params` / `Tile` are the two argument structs.
"""
import ctypes as C

import numpy as np

ABI_VERSION = 3

OK, ERR_ARG, ERR_CUDA, ERR_CAPACITY, ERR_STATE = 0, -1, -2, -3, -4

FLAG_EM_DELTA, FLAG_LRT_CHI, FLAG_MIN_AF, FLAG_ARGMIN_TIE, FLAG_QUAL60 = 1, 2, 4, 8, 16


def make_code(base, strand, indel):
    """BVC_MAKE_CODE: base:3 | strand:1 | is_indel:1 (AlleleInfo, src/BamProcess.h:30-39)."""
    return (np.asarray(base, np.uint8) & 7) | ((np.asarray(strand, np.uint8) & 1) << 3) | (
        (np.asarray(indel, np.uint8) & 1) << 4)


class Params(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("n_samples", C.c_int32),
        ("min_af", C.c_double),
        ("lrt_threshold", C.c_double),
        ("em_epsilon", C.c_double),
        ("em_max_iter", C.c_int32),
        ("n_groups", C.c_int32),
        ("sample_group", C.POINTER(C.c_int32)),
        ("max_sites", C.c_int32),
        ("reserved0", C.c_int32),
        ("max_entries", C.c_int64),
    ]


class Tile(C.Structure):
    _fields_ = [
        ("n_sites", C.c_int32),
        ("reserved0", C.c_int32),
        ("ref_base", C.c_void_p),
        ("entry_off", C.c_void_p),
        ("code", C.c_void_p),
        ("qual", C.c_void_p),
        ("mapq", C.c_void_p),
        ("rpr", C.c_void_p),
        ("sample", C.c_void_p),
        ("packed8", C.c_void_p),
        ("qual_dict", C.c_void_p),
        ("n_qual_dict", C.c_int32),
        ("reserved1", C.c_int32),
    ]


class CodedTile(C.Structure):
    """bvc_coded_tile (include/bvc_gpu.h, coded wire form)."""
    _fields_ = [
        ("bits", C.c_int32),
        ("n_codes", C.c_int32),
        ("code_byte", C.c_uint8 * 64),
        ("words", C.c_void_p),
        ("escapes", C.c_void_p),
        ("chunk_escapes", C.c_void_p),
        ("n_escape", C.c_int64),
    ]


CODED_CHUNK = 8192

SITE_RESULT_DTYPE = np.dtype(
    [
        ("cvg_fs", "<f8"),
        ("cvg_sor", "<f8"),
        ("af_lrt", "<f8", (4,)),
        ("chi", "<f8"),
        ("var_qual", "<f8"),
        ("vcf_fs", "<f8"),
        ("vcf_sor", "<f8"),
        ("ranksum", "<f8", (3,)),
        ("n_entries", "<i4"),
        ("depth", "<i4", (4,)),
        ("cvg_strand", "<i4", (4,)),
        ("vcf_strand", "<i4", (4,)),
        ("em_fits", "<i4"),
        ("em_passes", "<i4"),
        ("flags", "<u4"),
        ("cvg_alt_base", "i1"),
        ("lrt_ok", "i1"),
        ("n_alt", "i1"),
        ("alt_bases", "i1", (3,)),
        ("pad", "i1", (2,)),
    ],
    align=False,
)
assert SITE_RESULT_DTYPE.itemsize == 176, SITE_RESULT_DTYPE.itemsize

# bvc_site_core (include/bvc_gpu.h): the 72-byte part of a record every site has -- bvc_enqueue_split
SITE_CORE_DTYPE = np.dtype(
    [
        ("cvg_fs", "<f8"),
        ("cvg_sor", "<f8"),
        ("n_entries", "<i4"),
        ("depth", "<i4", (4,)),
        ("cvg_strand", "<i4", (4,)),
        ("em_fits", "<i4"),
        ("em_passes", "<i4"),
        ("flags", "<u4"),
        ("cvg_alt_base", "i1"),
        ("lrt_ok", "i1"),
        ("n_alt", "i1"),
        ("reserved", "i1"),
        ("vcf_index", "<i4"),
    ],
    align=False,
)
assert SITE_CORE_DTYPE.itemsize == 72, SITE_CORE_DTYPE.itemsize

# integer / index fields that must be bit-exact against the reference, and the FP64 fields held to 1e-9
EXACT_FIELDS = ("n_entries", "depth", "cvg_strand", "cvg_alt_base", "lrt_ok", "n_alt", "alt_bases")
EXACT_FIELDS_IF_OK = ("vcf_strand",)
FLOAT_FIELDS = ("cvg_fs", "cvg_sor")
FLOAT_FIELDS_IF_OK = ("af_lrt", "chi", "var_qual", "vcf_fs", "vcf_sor", "ranksum")


def default_params(n_samples, maf=0.001, max_sites=0, max_entries=0):
    """bvc_default_params in Python: min_af = min(100/N, 0.001, maf) (src/BaseVarC.cpp:541-543)."""
    p = Params()
    p.abi_version = ABI_VERSION
    p.n_samples = int(n_samples)
    min_af = 100.0 / n_samples
    if min_af > 0.001:
        min_af = 0.001
    if maf < min_af:
        min_af = maf
    p.min_af = min_af
    p.lrt_threshold = 24.0
    p.em_epsilon = 0.001
    p.em_max_iter = 100
    p.n_groups = 0
    p.sample_group = None
    p.max_sites = int(max_sites)
    p.max_entries = int(max_entries)
    return p
