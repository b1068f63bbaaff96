"""The reference's on-disk hand-off between `--load` and `--rerun`: BGZF text batch files.

Writer side only (fixtures, benchmarks, the CPU-baseline run of the unmodified reference binary):
  {prefix}.tmp.thread.{t}/batch.{j}   t < T threads, j < ceil(N/B) sample batches
  line 1 = sample names, each followed by a tab; then one line per site of thread t's window with one
  token per sample: "." | "b,mapq,qual,rpr,strand" | "+SEQ" | "-SEQ", each followed by one space
(src/BaseVarC.cpp:467-534 writes them, :316-465 reads them; window formula :399-403,497,523).
The product-side reader/packer is the C++ dispatcher (basevarc_b200/csrc/host), not this module.
"""
import os
import struct
import zlib

import numpy as np

_BGZF_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


class BgzfWriter:
    """Minimal BGZF (blocked gzip, 'BC' extra field, <= 64 KiB blocks, EOF marker) writer."""

    def __init__(self, path, level=1, block=0xff00):
        self.f = open(path, "wb")
        self.buf = bytearray()
        self.level = level
        self.block = block

    def write(self, data):
        self.buf += data
        while len(self.buf) >= self.block:
            self._flush_block(bytes(self.buf[:self.block]))
            del self.buf[:self.block]

    def _flush_block(self, raw):
        c = zlib.compressobj(self.level, zlib.DEFLATED, -15)
        comp = c.compress(raw) + c.flush()
        bsize = len(comp) + 25
        self.f.write(b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" + struct.pack("<H", bsize))
        self.f.write(comp)
        self.f.write(struct.pack("<II", zlib.crc32(raw) & 0xffffffff, len(raw)))

    def close(self):
        if self.buf:
            self._flush_block(bytes(self.buf))
            self.buf = bytearray()
        self.f.write(_BGZF_EOF)
        self.f.close()


def read_bgzf_text(path):
    """Decompress a BGZF/gzip file fully (concatenated members) -> str.  Test/fixture helper."""
    import gzip
    with gzip.open(path, "rb") as f:
        return f.read().decode()


def window_bounds(n_sites, threads):
    """The reference's site partition (src/BaseVarC.cpp:399-403): window = P%T + P/T."""
    w = n_sites % threads + n_sites // threads
    b = [min(t * w, n_sites) for t in range(threads)] + [n_sites]
    if threads > 1 and (threads - 1) * w > n_sites:
        raise ValueError("the reference indexes past its site list for this (sites, threads) shape")
    return b


def tile_tokens(tile, default_indel="+A"):
    """Per-entry token strings of a PileupTile (indel entries use tile.indel_text or `default_indel`)."""
    base = tile.code & 7
    strand = (tile.code >> 3) & 1
    indel = (tile.code >> 4) & 1
    toks = np.char.add(np.char.add(np.char.add(np.char.add(
        base.astype("U1"), ","), np.char.add(tile.mapq.astype("U3"), ",")),
        np.char.add(tile.qual.astype("U3"), ",")), np.char.add(np.char.add(tile.rpr.astype("U3"), ","), strand.astype("U1")))
    toks = toks.astype(object)
    for k in np.nonzero(indel)[0]:
        toks[k] = tile.indel_text.get(int(k), default_indel)
    return toks


def write_batch_files(tile, prefix, threads, batch, sample_names=None, level=1):
    """Write the tile as the reference's batch files.  Returns (paths[t][j], sample_names)."""
    N, S = tile.n_samples, tile.n_sites
    names = sample_names or ["S%d" % i for i in range(N)]
    nb = 1 + (N - 1) // batch
    bounds = window_bounds(S, threads)
    toks = tile_tokens(tile)
    off = tile.entry_off.astype(np.int64)
    paths = []
    for t in range(threads):
        d = "%s.tmp.thread.%d" % (prefix, t)
        os.makedirs(d, exist_ok=True)
        ws = [BgzfWriter(os.path.join(d, "batch.%d" % j), level) for j in range(nb)]
        for j, w in enumerate(ws):
            w.write(("".join(n + "\t" for n in names[j * batch:(j + 1) * batch]) + "\n").encode())
        for s in range(bounds[t], bounds[t + 1]):
            row = np.full(N, ".", dtype=object)
            e0, e1 = off[s], off[s + 1]
            row[tile.sample[e0:e1]] = toks[e0:e1]
            for j, w in enumerate(ws):
                w.write((" ".join(row[j * batch:(j + 1) * batch]) + " \n").encode())
        for w in ws:
            w.close()
        paths.append([os.path.join(d, "batch.%d" % j) for j in range(nb)])
    return paths, names


def write_fasta(path, chrom, ref_base, first_pos, pad=1100):
    """A plain FASTA whose bases at 1-based positions first_pos.. are the tile's reference bases (the
    rest 'N'), long enough for the reference's +1000 fetch (src/BaseVarC.cpp:194-196)."""
    seq = np.full(first_pos - 1 + ref_base.size + pad, ord("N"), np.uint8)
    seq[first_pos - 1:first_pos - 1 + ref_base.size] = np.frombuffer(b"ACGT", np.uint8)[ref_base]
    with open(path, "w") as f:
        f.write(">%s\n" % chrom)
        s = seq.tobytes().decode()
        for i in range(0, len(s), 60):
            f.write(s[i:i + 60] + "\n")
    fai = path + ".fai"
    if os.path.exists(fai):
        os.remove(fai)


def region_for(chrom, first_pos, n_sites):
    """`--region` string whose processed sites are exactly first_pos .. first_pos+n_sites-1
    (the reference drops the last two positions: SURVEY S1)."""
    return "%s:%d-%d" % (chrom, first_pos, first_pos + n_sites + 1)
