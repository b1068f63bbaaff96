"""Packed tile files (include/bvc_host.h): the binary form of the reference's batch files.
CPU: `BaseVarC_b200 basetype --pack` on the reference's own C1 batch files (no GPU needed) and bvc_write_packed_files on
synthetic tiles give back, tile by tile, exactly what the text packer produces.  GPU: `--rerun` from the packed files prints
the reference's text."""
import os
import shutil
import subprocess

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import batchfile, host
from basevarc_b200.pileup import synth_tile
from test_c1 import C1, BATCH, N, T, _ref_text, _sites, _window


def _c1_workdir(tmp_path):
    length, lo, seq = _window()
    full = bytearray(b"N" * length)
    full[lo:lo + len(seq)] = seq.encode()
    with open(tmp_path / "chr17.fa", "wb") as f:
        f.write(b">chr17\n")
        for i in range(0, length, 60):
            f.write(full[i:i + 60]); f.write(b"\n")
    for t in range(T):
        shutil.copytree(os.path.join(C1, "test.out.tmp.thread.%d" % t), tmp_path / ("test.out.tmp.thread.%d" % t))
    shutil.copy(os.path.join(C1, "bam.list"), tmp_path / "bam.list")
    with open(os.path.join(C1, "ARGS")) as f:
        return f.read().split() + ["-i", "bam.list", "-r", "chr17.fa", "-o", "test.out"]


def _concat(tiles):
    pos = np.concatenate([p for p, _, _ in tiles])
    out = {k: np.concatenate([getattr(t, k) for _, t, _ in tiles]) for k in ("ref_base", "code", "qual", "mapq", "rpr", "sample")}
    out["n"] = np.concatenate([np.diff(t.entry_off.astype(np.int64)) for _, t, _ in tiles])
    ind, base = {}, 0
    for _, t, d in tiles:
        ind.update({base + k: v for k, v in d.items()})
        base += t.n_entries
    return pos, out, ind


def test_pack_cli_on_c1_equals_the_text_packer(tmp_path):
    args = _c1_workdir(tmp_path)
    r = subprocess.run([host.BIN_PATH, "basetype", "--pack", "--tile_sites", "700"] + args, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "basetype packing done" in r.stdout
    pv = _sites()
    window = len(pv) % T + len(pv) // T
    n_indel = 0
    for t in range(T):
        path = str(tmp_path / ("test.out.tmp.thread.%d" % t) / "packed.bvc")
        assert host.load_host_library().bvc_packed_is_complete(path.encode(), N) == 1
        names, tiles = host.read_packed_file(path)
        assert len(names.split("\t")) == N and len(tiles) > 5 and all(tl.n_sites <= 700 for _, tl, _ in tiles)
        pos, got, ind = _concat(tiles)
        # the text packer over the same window (one stale-token state per window: quirk Q1)
        mine = pv[t * window:] if t == T - 1 else pv[t * window:(t + 1) * window]
        files = [batchfile.read_bgzf_text(os.path.join(C1, "test.out.tmp.thread.%d" % t, "batch.%d" % j)).encode().split(b"\n")
                 for j in range(N // BATCH)]
        pk = host.Packer(N)
        for i, (p, rb) in enumerate(mine):
            pk.add_site(p, rb, [f[i + 1] for f in files])
        want = pk.to_tile()
        pk.close()
        keep = np.diff(want.entry_off.astype(np.int64)) > 0            # sites without entries print nothing: not stored
        assert np.array_equal(pos, np.asarray(want.pos)[keep])
        assert np.array_equal(got["ref_base"], want.ref_base[keep]) and np.array_equal(got["n"], np.diff(want.entry_off.astype(np.int64))[keep])
        for k in ("code", "qual", "mapq", "rpr", "sample"):
            assert np.array_equal(got[k], getattr(want, k)), k
        assert ind == want.indel_text
        n_indel += len(ind)
    assert n_indel == 57
    # C1 has 39 distinct quals (> 30): its tiles may be stored in either form; a cut-short file never validates
    path = str(tmp_path / "test.out.tmp.thread.0" / "packed.bvc")
    with open(path, "r+b") as f:
        f.seek(16); f.write(b"\0\0\0\0")                                  # n_tiles = 0 = "writer never finished"
    assert host.load_host_library().bvc_packed_is_complete(path.encode(), N) == 0


@pytest.mark.parametrize("quals,indel", [((20, 40), 0.02), ((2, 93), 0.0)])
def test_write_packed_files_roundtrip(tmp_path, quals, indel):
    rng = np.random.default_rng(3)
    tile = synth_tile(300, 2500, 0.8, "highpoly", seed=8, indel_frac=indel)
    tile.qual = rng.integers(quals[0], quals[1] + 1, tile.qual.size).astype(np.uint8)
    tile.pos = (np.arange(tile.n_sites) * 3 + 1000).astype(np.int32)
    prefix = str(tmp_path / "w")
    host.write_packed_files(tile, prefix, threads=3, tile_sites=400, indel_token="-AC")
    tiles = []
    for t in range(3):
        _, ts = host.read_packed_file(prefix + ".tmp.thread.%d/packed.bvc" % t)
        tiles += ts
    pos, got, ind = _concat(tiles)
    keep = np.diff(tile.entry_off.astype(np.int64)) > 0
    assert np.array_equal(pos, tile.pos[keep]) and np.array_equal(got["n"], np.diff(tile.entry_off.astype(np.int64))[keep])
    is_indel = ((tile.code >> 4) & 1) == 1
    for k in ("code", "mapq", "rpr", "sample"):
        assert np.array_equal(got[k], getattr(tile, k)), k
    assert np.array_equal(got["qual"], tile.qual)                          # indel entries keep their stale qual (side table)
    assert sorted(ind) == np.nonzero(is_indel)[0].tolist() and set(ind.values()) <= {"-AC"}
    packed_form = [tl.packed8 is not None for _, tl, _ in tiles]
    assert all(packed_form) if quals == (20, 40) else not any(packed_form)


@pytest.mark.gpu
def test_rerun_from_packed_files_prints_the_reference_text(tmp_path):
    from test_gpu_rerun import _compare_text
    args = _c1_workdir(tmp_path)
    r = subprocess.run([host.BIN_PATH, "basetype", "--pack"] + args, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    for t in range(T):                                                     # the text files are no longer needed
        for j in range(N // BATCH):
            os.remove(tmp_path / ("test.out.tmp.thread.%d" % t) / ("batch.%d" % j))
    r = subprocess.run([host.BIN_PATH, "basetype", "--rerun", "--keep_tmp"] + args, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "reading packed tile files" in r.stderr
    cvg = batchfile.read_bgzf_text(str(tmp_path / "test.out.cvg.gz"))
    vcf = batchfile.read_bgzf_text(str(tmp_path / "test.out.vcf.gz"))
    _compare_text(cvg, vcf, _ref_text("cvg"), _ref_text("vcf"), N)
    assert len([l for l in vcf.splitlines() if not l.startswith("#")]) == 76
    # --no_packed insists on the text files (now gone): the command refuses, pointing at --load
    r = subprocess.run([host.BIN_PATH, "basetype", "--rerun", "--no_packed"] + args, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode != 0 and "batch files are complete" in r.stderr


def test_inflate_pool_gives_the_same_packed_files(tmp_path):
    """BGZF blocks inflated ahead by a pool of threads, by the caller's thread, and by zlib only."""
    outs = []
    for name, env in (("pool", {"BVC_INFLATE_THREADS": "3"}), ("inline", {"BVC_INFLATE_THREADS": "0"}),
                      ("zlib", {"BVC_INFLATE_THREADS": "2", "BVC_ZLIB_INFLATE": "1"})):
        d = tmp_path / name
        d.mkdir()
        args = _c1_workdir(d)
        r = subprocess.run([host.BIN_PATH, "basetype", "--pack", "-v", "--tile_sites", "500"] + args, cwd=d, env=dict(os.environ, **env),
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        assert ("inflate pool : %s threads" % env["BVC_INFLATE_THREADS"]) in r.stdout
        outs.append([open(d / ("test.out.tmp.thread.%d" % t) / "packed.bvc", "rb").read() for t in range(T)])
    assert outs[0] == outs[1] == outs[2] and all(len(b) > 1000 for b in outs[0])


@pytest.mark.parametrize("threads", ["0", "3"])
def test_corrupt_batch_file_fails_the_command_with_and_without_the_pool(tmp_path, threads):
    """A flipped byte inside a BGZF block (CRC mismatch after zlib's second opinion) and a block cut short: the command stops with
    the reader's message -- with the block pool too, where the failure surfaces on a pool thread."""
    for kind in ("flip", "cut"):
        d = tmp_path / kind
        d.mkdir()
        args = _c1_workdir(d)
        victim = d / "test.out.tmp.thread.1" / "batch.3"
        blob = bytearray(open(victim, "rb").read())
        if kind == "flip":
            blob[len(blob) // 2] ^= 0x10
        else:
            blob = blob[:len(blob) // 2]
        open(victim, "wb").write(bytes(blob))
        r = subprocess.run([host.BIN_PATH, "basetype", "--pack"] + args, cwd=d, env=dict(os.environ, BVC_INFLATE_THREADS=threads),
                           capture_output=True, text=True, timeout=120)
        assert r.returncode != 0, kind
        assert "corrupt batch file" in r.stderr or "complete" in r.stderr, r.stderr[-500:]
        assert not os.path.exists(d / "test.out.tmp.thread.1" / "packed.bvc")      # nothing half-written passes for a packed file


def test_lazy_read_of_packed_tiles_equals_the_full_read(tmp_path):
    """The dispatcher reads one-byte tiles lazily (indel entries only; code / qual / sample of a site on demand for the VCF
    columns): tools/ubench/packed_lazy_test.cc holds PackedIo / Packer to the full read on C1's files and on a synthetic
    window with many indel entries, N-base ones included."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "packed_lazy_test")
    cc = subprocess.run(["g++", "-O1", "-std=c++17", "-I" + os.path.join(root, "include"), "-I" + os.path.join(root, "basevarc_b200", "csrc", "host"),
                         os.path.join(root, "tools", "ubench", "packed_lazy_test.cc"), os.path.join(root, "basevarc_b200", "csrc", "host", "packed_io.cc"),
                         os.path.join(root, "basevarc_b200", "csrc", "host", "packer.cc"), "-o", exe], capture_output=True, text=True, timeout=300)
    assert cc.returncode == 0, cc.stderr[-2000:]
    args = _c1_workdir(tmp_path)
    r = subprocess.run([host.BIN_PATH, "basetype", "--pack", "--tile_sites", "600"] + args, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    files = [str(tmp_path / ("test.out.tmp.thread.%d" % t) / "packed.bvc") for t in range(T)]
    tile = synth_tile(300, 2500, 1.2, "highpoly", seed=5, indel_frac=0.08)
    c = tile.code.copy()
    ind = np.nonzero((c & 16) != 0)[0]
    c[ind[::5]] = (c[ind[::5]] & 0x18) | 4
    tile.code = c
    host.write_packed_files(tile, str(tmp_path / "syn"), 2, tile_sites=400)
    files += [str(tmp_path / ("syn.tmp.thread.%d" % k) / "packed.bvc") for k in range(2)]
    r = subprocess.run([exe] + files, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "lazy == full" in r.stdout, r.stdout + r.stderr
