"""CPU: the whole-buffer DEFLATE decoder and the carry-less-multiply CRC-32 of the BGZF reader (csrc/host/fast_inflate.cc,
SURVEY 8(f) N1) against zlib: every block type, every compression level and strategy, sizes around the BGZF block limit, the
reference's own batch files (tests/golden/c1), and corrupted streams (declined or decoded, never a crash, never too long)."""
import glob
import os
import struct
import zlib

import numpy as np
import pytest

from basevarc_b200 import host

HERE = os.path.dirname(os.path.abspath(__file__))
MAXB = 65536


def _raw(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, mem=8):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, mem, strategy)
    return c.compress(data) + c.flush()


def _payloads():
    rng = np.random.default_rng(7)
    text = b" ".join(b"%d,%d,%d,%d,%d" % (rng.integers(0, 4), rng.integers(1, 60), rng.integers(2, 41), rng.integers(1, 101), rng.integers(0, 2))
                     if rng.random() < 0.4 else b"." for _ in range(9000))
    yield "empty", b""
    yield "one", b"A"
    yield "run", b"z" * 65280                                   # distance 1, length 258 chains
    yield "period3", b"abc" * 21000                             # distances 2..7
    yield "period7", b"abcdefg" * 9000
    yield "text", text[:65280]
    yield "random", rng.integers(0, 256, 65280, dtype=np.uint8).tobytes()      # stored blocks at every level
    yield "lowent", rng.integers(0, 3, 65280, dtype=np.uint8).tobytes()
    yield "full", (text * 2)[:MAXB]
    for n in (1, 2, 3, 7, 8, 9, 257, 258, 259, 4095, 65279):
        yield "text%d" % n, text[:n]


@pytest.mark.parametrize("level", [0, 1, 2, 4, 6, 9])
def test_matches_zlib_at_every_level(level):
    for name, data in _payloads():
        got = host.fast_inflate(_raw(data, level), MAXB)
        assert got == data, (name, level, None if got is None else len(got))


@pytest.mark.parametrize("strategy", [zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FILTERED])
def test_matches_zlib_for_every_strategy(strategy):
    for name, data in _payloads():
        for mem in (1, 9):                                       # memLevel 1: many small blocks per stream
            got = host.fast_inflate(_raw(data, 6, strategy, mem), MAXB)
            assert got == data, (name, strategy, mem)


def test_multi_block_streams_mixing_types():
    rng = np.random.default_rng(3)
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    parts, raw = [], b""
    for i in range(40):
        p = (b"chr17\t%d\tA\t.\t" % i) * int(rng.integers(1, 60)) if i % 3 else rng.integers(0, 256, int(rng.integers(1, 900)), dtype=np.uint8).tobytes()
        parts.append(p)
        raw += c.compress(p) + c.flush(zlib.Z_FULL_FLUSH if i % 2 else zlib.Z_SYNC_FLUSH)      # empty stored blocks between
    raw += c.flush()
    data = b"".join(parts)
    assert host.fast_inflate(raw, MAXB) == data


def test_output_limit_and_truncation_are_declined():
    data = b"0,40,30,12,1 . " * 3000
    raw = _raw(data, 6)
    assert host.fast_inflate(raw, len(data)) == data
    assert host.fast_inflate(raw, len(data) - 1) is None         # one byte more than out_cap
    for cut in (1, 2, 5, len(raw) // 2, len(raw) - 1):
        assert host.fast_inflate(raw[:cut], MAXB) is None, cut     # the bits a cut stream still needs lie outside the input


def test_corrupted_streams_never_crash_or_overrun():
    rng = np.random.default_rng(11)
    data = (b"2,27,28,65,0 . . 3,35,35,32,1 " * 2200)[:65000]
    for level in (1, 6):
        raw = bytearray(_raw(data, level))
        for _ in range(1500):
            bad = bytearray(raw)
            for _k in range(int(rng.integers(1, 4))):
                bad[int(rng.integers(0, len(bad)))] ^= 1 << int(rng.integers(0, 8))
            got = host.fast_inflate(bytes(bad), MAXB)
            assert got is None or len(got) <= MAXB
    for _ in range(500):                                         # pure noise as a stream
        got = host.fast_inflate(rng.integers(0, 256, int(rng.integers(1, 400)), dtype=np.uint8).tobytes(), MAXB)
        assert got is None or len(got) <= MAXB


def test_crc32_equals_zlib():
    rng = np.random.default_rng(5)
    buf = rng.integers(0, 256, 70000, dtype=np.uint8).tobytes()
    for n in list(range(0, 200)) + [255, 256, 1023, 4096, 65279, 65280, 65536, 70000]:
        for off in (0, 1, 5):
            d = buf[off:off + n]
            assert host.fast_crc32(d) == zlib.crc32(d), (n, off)


def test_every_block_of_the_reference_batch_files():
    """The 40 batch files the reference's load phase wrote for its own test/ example (htslib's deflate)."""
    files = sorted(glob.glob(os.path.join(HERE, "golden", "c1", "**", "batch.*"), recursive=True))
    assert files
    n_blocks = 0
    for path in files:
        blob = open(path, "rb").read()
        pos = 0
        while pos < len(blob):
            assert blob[pos:pos + 4] == b"\x1f\x8b\x08\x04"
            xlen = struct.unpack_from("<H", blob, pos + 10)[0]
            bsize = struct.unpack_from("<H", blob, pos + 16)[0] + 1        # the BC sub-field is the only one htslib writes
            raw = blob[pos + 12 + xlen:pos + bsize - 8]
            crc, isize = struct.unpack_from("<II", blob, pos + bsize - 8)
            want = zlib.decompress(raw, -15)
            assert len(want) == isize
            assert host.fast_inflate(raw, MAXB) == want
            assert host.fast_crc32(want) == crc
            pos += bsize
            n_blocks += 1
    assert n_blocks >= len(files)
