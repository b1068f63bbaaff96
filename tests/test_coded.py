"""The coded wire form (include/bvc_gpu.h: W = 4 / 5 / 6 bits per entry, bytes relative to the site's reference base, escapes in
a side array).  CPU: bvc_encode_codes -> bvc_decode_codes_host (the unit decoder the device kernel shares, csrc/bvc_codes.h) gives
the packed8 bytes back, and an independent numpy reading of the bit stream agrees.  GPU: bvc_enqueue_split_coded returns the very
records of the packed8 call."""
import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200._abi import CODED_CHUNK
from basevarc_b200.pileup import synth_tile


def _numpy_decode(c, tile):
    """Bit stream -> bytes without the library: entry e is W bits at bit e*W of the little-endian stream."""
    W, n = c.bits, tile.n_entries
    bits = np.unpackbits(c.words[:(n + 31) // 32 * W].view(np.uint8), bitorder="little")
    codes = (bits[:((n + 31) // 32) * 32 * W].reshape(-1, W) * (1 << np.arange(W))).sum(1)[:n]
    lut = np.frombuffer(bytes(c.ct.code_byte), np.uint8)
    rel = lut[np.minimum(codes, 63)]
    ref = np.repeat(tile.ref_base.astype(np.uint8) & 3, np.diff(tile.entry_off.astype(np.int64)))
    out = (rel & 0x3F) | ((((rel >> 6) + ref) & 3) << 6)
    esc = codes == (1 << W) - 1
    assert int(esc.sum()) == int(c.ct.n_escape)
    out[esc] = c.escapes[:int(esc.sum())]
    # the per-chunk offsets the kernel's blocks start from
    want = np.concatenate([[0], np.cumsum(np.add.reduceat(esc.astype(np.int64), np.arange(0, n, CODED_CHUNK)))])
    assert np.array_equal(c.chunk_escapes[:want.size], want.astype(np.uint32))
    return out.astype(np.uint8)


def _same_split(core, vcf, n, core0, vcf0):
    """Two split results are the same records: the compaction order of the LRT-positive sites (an atomic counter) may differ."""
    for f in core.dtype.names:
        if f != "vcf_index" and core[f].tobytes() != core0[f].tobytes():
            return False
    pos = np.nonzero(core["lrt_ok"] == 1)[0]
    if n != vcf0.shape[0] or pos.size != n:
        return False
    return vcf[core["vcf_index"][pos]].tobytes() == vcf0[core0["vcf_index"][pos]].tobytes()


def _tiles():
    yield "lowpass", synth_tile(1000, 3000, 0.5, "lowpass", seed=1), 6
    yield "highpoly+indels", synth_tile(200, 5000, 1.0, "highpoly", seed=2, indel_frac=0.05), 6
    yield "ragged", synth_tile(60, 9000, 0.05, "lowpass", seed=7, indel_frac=0.2), 6
    t = synth_tile(1000, 3000, 0.5, "lowpass", seed=4)
    t.qual = np.random.default_rng(1).choice(np.array([2, 12, 23, 37], np.uint8), t.qual.size)
    yield "binned quals", t, 4
    t = synth_tile(500, 2000, 0.7, "lowpass", seed=5)
    t.qual = np.random.default_rng(2).choice(np.array([3, 5, 9, 12, 14, 17, 20, 25, 28, 31, 36, 40], np.uint8), t.qual.size)
    yield "12 quals", t, 5


def test_encode_decode_round_trip():
    for name, tile, want_bits in _tiles():
        assert tile.pack8()
        c = B.CodedArrays(tile)
        assert c.bits == want_bits, (name, c.bits)
        assert c.wire_bytes < 0.8 * tile.n_entries
        assert np.array_equal(c.decode_host(), tile.packed8), name
        assert np.array_equal(_numpy_decode(c, tile), tile.packed8), name


def test_not_coded_when_it_does_not_pay():
    small = synth_tile(50, 40, 0.5, "lowpass", seed=3)
    assert small.pack8() and B.CodedArrays(small).bits == 0            # too small for a kernel
    wild = synth_tile(300, 3000, 1.0, "highpoly", seed=9)
    rng = np.random.default_rng(3)
    wild.qual = rng.integers(1, 31, wild.qual.size).astype(np.uint8)   # 30 quals x 2 strands x 4 bases, nearly flat
    wild.code = ((wild.code & 0x18) | rng.integers(0, 4, wild.code.size).astype(np.uint8)).astype(np.uint8)
    assert wild.pack8() and B.CodedArrays(wild).bits == 0              # > n/8 escapes at 6 bits


@pytest.mark.gpu
def test_coded_call_returns_the_packed8_records():
    for name, tile, _ in _tiles():
        assert tile.pack8()
        c = B.CodedArrays(tile)
        caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
        h0 = caller.transfer_bytes[0]
        core0, vcf0 = caller.call_sites_split(tile)
        h1 = caller.transfer_bytes[0]
        core = np.zeros(tile.n_sites, B.SITE_CORE_DTYPE)
        vcf = np.zeros(tile.n_sites, B.SITE_RESULT_DTYPE)
        n = caller.wait_split(caller.enqueue_split_coded(tile, c, core, vcf))
        h2 = caller.transfer_bytes[0]
        caller.close()
        assert _same_split(core, vcf, n, core0, vcf0), name
        # the entries' bytes on the wire: packed8 sends n, the coded form what CodedArrays.wire_bytes says (offsets, reference bases,
        # mapq / rpr travel as before)
        assert abs((h1 - h0) - (h2 - h1) - (tile.n_entries - c.wire_bytes)) <= 64, (name, h1 - h0, h2 - h1, c.wire_bytes)
        assert c.wire_bytes < 0.8 * tile.n_entries


@pytest.mark.gpu
def test_coded_call_with_a_qual_0_tile_expands_after_decoding():
    tile = synth_tile(64, 3000, 1.5, "highpoly", seed=13, indel_frac=0.05)
    q = tile.qual.copy(); q[::17] = 0; tile.qual = q
    assert tile.pack8()
    c = B.CodedArrays(tile)
    assert c.bits > 0
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    core0, vcf0 = caller.call_sites_split(tile)
    core = np.zeros(tile.n_sites, B.SITE_CORE_DTYPE)
    vcf = np.zeros(tile.n_sites, B.SITE_RESULT_DTYPE)
    n = caller.wait_split(caller.enqueue_split_coded(tile, c, core, vcf))
    caller.close()
    assert _same_split(core, vcf, n, core0, vcf0)
