"""-m gpu: bvc_call_sites_noalloc (include/bvc_gpu.h), the allocation-free entry point, through the C ABI.

Collected LAST on purpose.  The entry point has no counterpart in the reference; its records
are held to the existing entry point's (bvc_enqueue + bvc_wait): bitwise to the same kernel behind a context
(BVC_LEGACY_KERNEL=1) and within the suite's parity tolerance (tests/parity.py) to the default pipeline, on two calls in a row
that share the library's static scratch array.  The caller's device memory comes from torch here; the entry point itself
allocates nothing.
"""

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import caller as bc
from basevarc_b200._abi import SITE_RESULT_DTYPE
from basevarc_b200.pileup import synth_tile
from parity import compare_records

pytestmark = pytest.mark.gpu

REC = SITE_RESULT_DTYPE.itemsize


def _existing(tile, monkeypatch, legacy):
    """Records of the existing entry point; legacy=True: the one-warp-per-site kernel for every site."""
    if legacy:
        monkeypatch.setenv("BVC_LEGACY_KERNEL", "1")
    else:
        monkeypatch.delenv("BVC_LEGACY_KERNEL", raising=False)
    c = B.BaseTypeCaller(tile.n_samples, max_sites=max(tile.n_sites, 1), max_entries=max(tile.n_entries, 1))
    try:
        return c.call_sites(tile)
    finally:
        c.close()
        monkeypatch.delenv("BVC_LEGACY_KERNEL", raising=False)


class _DeviceTile:
    """The tile's arrays and an output / status buffer in device memory owned by the CALLER (torch)."""

    def __init__(self, torch, tile, fill=0xEE):
        dev = torch.device("cuda:0")
        self.n_sites, self.n_entries, self.n_samples = tile.n_sites, tile.n_entries, tile.n_samples
        up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt)).to(dev)
        self.ref = up(tile.ref_base, np.uint8)
        self.off = up(tile.entry_off.view(np.int32), np.int32)
        self.code, self.qual, self.mapq, self.rpr = (up(getattr(tile, k), np.uint8) for k in ("code", "qual", "mapq", "rpr"))
        self.out = torch.full((max(tile.n_sites, 1) * REC,), fill, dtype=torch.uint8, device=dev)
        self.status = torch.full((1,), 77, dtype=torch.int32, device=dev)
        self.params = B.default_params(tile.n_samples, 0.001)

    def call(self, stream):
        bc.call_sites_noalloc(self.params, self.n_sites, self.n_entries, self.ref.data_ptr(), self.off.data_ptr(), self.code.data_ptr(),
                              self.qual.data_ptr(), self.mapq.data_ptr(), self.rpr.data_ptr(), self.out.data_ptr(),
                              self.status.data_ptr(), stream)

    def records(self):
        return np.frombuffer(self.out.cpu().numpy().tobytes(), SITE_RESULT_DTYPE, count=self.n_sites)


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    return torch


def test_noalloc_matches_existing_entry_point_twice_in_a_row(torch_cuda, monkeypatch):
    torch = torch_cuda
    tiles = [synth_tile(500, 3000, 0.8, "highpoly", seed=21, indel_frac=0.02),
             synth_tile(200, 1100, 0.5, "lowpass", seed=22)]          # the second call is smaller: nothing of the first may leak
    want_same_kernel = [_existing(t, monkeypatch, legacy=True) for t in tiles]
    want_pipeline = [_existing(t, monkeypatch, legacy=False) for t in tiles]
    assert sum(int(w["lrt_ok"].sum()) for w in want_pipeline) > 0      # not vacuous
    stream = torch.cuda.current_stream().cuda_stream
    # set-up call (n_sites == 0: the address of the static array + the tables, once per device), then a warm-up so that the
    # kernels' code is loaded before free memory is sampled
    bc.call_sites_noalloc(B.default_params(10, 0.001), 0, 0, None, None, None, None, None, None, None, None, stream)
    warm = _DeviceTile(torch, tiles[1])
    warm.call(stream)
    torch.cuda.synchronize()
    for rnd in range(2):
        dts = [_DeviceTile(torch, t) for t in tiles]
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        for d in dts:                                                  # two calls in a row, same stream, no synchronisation between
            d.call(stream)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before, "device memory was allocated inside bvc_call_sites_noalloc"
        for d, same, pipe in zip(dts, want_same_kernel, want_pipeline):
            assert int(d.status.item()) == 0
            got = d.records()
            assert got.tobytes() == same.tobytes(), "round %d: differs from the same kernel behind bvc_enqueue" % rnd
            nbad, lines = compare_records(got, pipe)
            assert nbad == 0, "\n".join(lines)


def test_noalloc_rejects_beyond_the_static_array(torch_cuda):
    torch = torch_cuda
    p = B.default_params(100, 0.001)
    with pytest.raises(bc.BvcError, match=r"beyond the static scratch array.*max %d" % bc.NOALLOC_MAX_SITES):
        bc.call_sites_noalloc(p, bc.NOALLOC_MAX_SITES + 1, 10, 1, 1, 1, 1, 1, 1, 1)
    big = B.default_params(bc.NOALLOC_MAX_SAMPLES + 1, 0.001)
    with pytest.raises(bc.BvcError, match=r"n_samples %d \(max %d\)" % (bc.NOALLOC_MAX_SAMPLES + 1, bc.NOALLOC_MAX_SAMPLES)):
        bc.call_sites_noalloc(big, 4, 10, 1, 1, 1, 1, 1, 1, 1)
    with pytest.raises(bc.BvcError, match="2\\^32"):
        bc.call_sites_noalloc(p, 4, 1 << 32, 1, 1, 1, 1, 1, 1, 1)


def test_noalloc_device_checks_write_status_and_no_record(torch_cuda):
    torch = torch_cuda
    stream = torch.cuda.current_stream().cuda_stream
    base = synth_tile(64, 300, 1.0, "highpoly", seed=23)
    cases = []
    t = synth_tile(64, 300, 1.0, "highpoly", seed=23); t.qual = t.qual.copy(); t.qual[5] = 200
    cases.append((t, 1))
    t = synth_tile(64, 300, 1.0, "highpoly", seed=23); t.code = t.code.copy(); t.code[7] = 6       # non-indel base 6
    cases.append((t, 2))
    t = synth_tile(64, 300, 1.0, "highpoly", seed=23); t.entry_off = t.entry_off.copy(); t.entry_off[10] = t.entry_off[12] + 1
    cases.append((t, 4))
    for tile, bit in cases:
        d = _DeviceTile(torch, tile, fill=0xEE)
        d.call(stream)
        torch.cuda.synchronize()
        assert int(d.status.item()) & bit, (int(d.status.item()), bit)
        assert bool((d.out == 0xEE).all()), "a record was written although the tile was rejected"
    good = _DeviceTile(torch, base)                                     # and the next call is not affected by the rejected ones
    good.call(stream)
    torch.cuda.synchronize()
    assert int(good.status.item()) == 0 and int(good.records()["n_entries"].sum()) == base.n_entries
