"""oracle/oracle_py.py -- TEST INFRASTRUCTURE ONLY.

ctypes bindings of (a) oracle/_ref/libbvc_oracle.so, this repo's per-read C restatement of the
reference's calling core (basetype_oracle.c), and (b) oracle/_ref/libref_oracle.so, the UNMODIFIED
reference objects behind ref_shim.cpp (present only where /root/reference was available at build
time; it travels to the GPU box as a built file).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from basevarc_b200._abi import SITE_RESULT_DTYPE, Params  # noqa: E402

PORT_SO = os.path.join(HERE, "_ref", "libbvc_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "BaseVarC_ref")


def build(ref=True):
    """make -C oracle [port|all]; the reference part is attempted only when /root/reference exists."""
    target = "all" if (ref and os.path.isdir("/root/reference/src")) else "port"
    subprocess.run(["make", "-s", "-C", HERE, target], check=True)


class OracleLRT(C.Structure):
    _fields_ = [("ok", C.c_int32), ("n_alt", C.c_int32), ("alt_bases", C.c_int32 * 4), ("depth", C.c_int32 * 4),
                ("af_lrt", C.c_double * 4), ("var_qual", C.c_double), ("depth_total", C.c_double),
                ("chi", C.c_double), ("em_fits", C.c_int32), ("em_passes", C.c_int32)]


class RefLRT(C.Structure):
    _fields_ = [("ok", C.c_int32), ("n_alt", C.c_int32), ("alt_bases", C.c_int32 * 4), ("depth", C.c_int32 * 4),
                ("af_lrt", C.c_double * 4), ("var_qual", C.c_double), ("depth_total", C.c_double)]


_port = None
_ref = None


def port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            build(ref=False)
        L = C.CDLL(PORT_SO)
        L.bvc_oracle_chisf.restype = C.c_double
        L.bvc_oracle_chisf.argtypes = [C.c_double, C.c_double]
        L.bvc_oracle_normsf.restype = C.c_double
        L.bvc_oracle_normsf.argtypes = [C.c_double]
        L.bvc_oracle_fisher.restype = C.c_double
        L.bvc_oracle_fisher.argtypes = [C.c_int] * 4
        L.bvc_oracle_ranksum.restype = C.c_double
        L.bvc_oracle_ranksum.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
        L.bvc_oracle_lrt_reads.restype = C.c_int
        L.bvc_oracle_lrt_reads.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_double,
                                           C.c_void_p, C.c_int32, C.POINTER(OracleLRT)]
        L.bvc_oracle_call_sites.restype = C.c_int
        L.bvc_oracle_call_sites.argtypes = [C.POINTER(Params), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.bvc_oracle_call_range.restype = C.c_int
        L.bvc_oracle_call_range.argtypes = [C.POINTER(Params), C.c_void_p, C.c_int32, C.c_int32,
                                            C.c_void_p, C.c_void_p, C.c_void_p]
        _port = L
    return _port


def ref():
    """The unmodified reference objects, or None when they were not built."""
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            return None
        L = C.CDLL(REF_SO)
        L.ref_chisf.restype = C.c_double
        L.ref_chisf.argtypes = [C.c_double, C.c_double]
        L.ref_normsf.restype = C.c_double
        L.ref_normsf.argtypes = [C.c_double]
        L.ref_fisher.restype = C.c_double
        L.ref_fisher.argtypes = [C.c_int] * 4
        L.ref_ranksum.restype = C.c_double
        L.ref_ranksum.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
        L.ref_lrt.restype = C.c_int
        L.ref_lrt.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_double,
                              C.c_void_p, C.c_int32, C.POINTER(RefLRT)]
        L.ref_writevcf.restype = C.c_int
        _ref = L
    return _ref


def _i8(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int8))


def port_lrt(bases, quals, ref_base, min_af, base_comb=None):
    b, q = _i8(bases), _i8(quals)
    out = OracleLRT()
    bc = None if base_comb is None else _i8(base_comb)
    port().bvc_oracle_lrt_reads(b.size, b.ctypes.data, q.ctypes.data, int(ref_base), float(min_af),
                                None if bc is None else bc.ctypes.data, 0 if bc is None else bc.size, C.byref(out))
    return out


def ref_lrt(bases, quals, ref_base, min_af, base_comb=None):
    b, q = _i8(bases), _i8(quals)
    out = RefLRT()
    bc = None if base_comb is None else _i8(base_comb)
    ref().ref_lrt(b.size, b.ctypes.data, q.ctypes.data, int(ref_base), float(min_af),
                  None if bc is None else bc.ctypes.data, 0 if bc is None else bc.size, C.byref(out))
    return out


def _f8(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def port_ranksum(x, y):
    x, y = _f8(x), _f8(y)
    return port().bvc_oracle_ranksum(x.ctypes.data, x.size, y.ctypes.data, y.size)


def ref_ranksum(x, y):
    x, y = _f8(x), _f8(y)
    return ref().ref_ranksum(x.ctypes.data, x.size, y.ctypes.data, y.size)


def call_sites(tile, params, s0=0, s1=None, want_groups=False):
    """Run the C restatement over sites [s0, s1) of a PileupTile; returns the record array (and group outputs)."""
    s1 = tile.n_sites if s1 is None else s1
    out = np.zeros(tile.n_sites, SITE_RESULT_DTYPE)
    ct = tile.as_ctypes()
    gd = ga = None
    if want_groups and params.n_groups > 0:
        gd = np.zeros((tile.n_sites, params.n_groups, 4), np.int32)
        ga = np.zeros((tile.n_sites, params.n_groups, 4), np.float64)
    rc = port().bvc_oracle_call_range(C.byref(params), C.addressof(ct), s0, s1, out.ctypes.data,
                                      None if gd is None else gd.ctypes.data, None if ga is None else ga.ctypes.data)
    if rc != 0:
        raise ValueError("oracle rejected the tile: rc=%d" % rc)
    return (out[s0:s1], gd, ga) if want_groups else out[s0:s1]
