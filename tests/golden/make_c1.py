#!/usr/bin/env python
"""Builds the C1 golden (BASELINE.json configs[0]: the reference's own test/ example) -- run HERE, where /root/reference exists.

The reference ships test/test.sh + 100 BAMs but not test/data/chr17.fa.gz (only its .fai/.gzi).  Every read carries MD:Z,
so the reference bases are recoverable wherever a read aligns; uncovered positions become N, which changes no
output row (a site without entries prints nothing, src/BaseVarC.cpp:442).  Steps:
  1. minimal BAM reader (BGZF = multi-member gzip) -> chr17 bases over the covered window from CIGAR + MD
  2. chr17.fa (full-length, N-filled) in a scratch dir; test.sh's exact arguments with the UNMODIFIED reference binary
     (oracle/_ref/BaseVarC_ref), --keep_tmp so the batch files written by its load phase survive
  3. fixture: tests/golden/c1/  batch files as written by the reference (BGZF), the window's bases, bam.list,
     and the reference's decompressed .cvg / .vcf text (gzip)
usage: python tests/golden/make_c1.py
"""
import gzip
import os
import shutil
import struct
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden", "c1")
REGION = "chr17:41197700-41276155"
ARGS = ["-q", "20", "-t", "4", "-b", "10"]
SEQ = "=ACMGRSVTWYHKDBN"


def bam_records(path):
    with gzip.open(path, "rb") as f:
        data = f.read()
    assert data[:4] == b"BAM\1"
    l_text, = struct.unpack_from("<i", data, 4)
    p = 8 + l_text
    n_ref, = struct.unpack_from("<i", data, p); p += 4
    refs = []
    for _ in range(n_ref):
        l_name, = struct.unpack_from("<i", data, p); p += 4
        name = data[p:p + l_name - 1].decode(); p += l_name
        l_ref, = struct.unpack_from("<i", data, p); p += 4
        refs.append((name, l_ref))
    while p < len(data):
        bs, = struct.unpack_from("<i", data, p); p += 4
        rec = data[p:p + bs]; p += bs
        ref_id, pos, l_rn, mapq, _bin, n_cig, flag, l_seq = struct.unpack_from("<iiBBHHHi", rec, 0)
        q = 32 + l_rn
        cigar = [(c & 15, c >> 4) for c in struct.unpack_from("<%dI" % n_cig, rec, q)]; q += 4 * n_cig
        sb = rec[q:q + (l_seq + 1) // 2]; q += (l_seq + 1) // 2
        seq = "".join(SEQ[(sb[i >> 1] >> (4 if i % 2 == 0 else 0)) & 15] for i in range(l_seq))
        q += l_seq
        md = None
        while q < len(rec):                                  # tags
            tag, typ = rec[q:q + 2], chr(rec[q + 2]); q += 3
            if typ == "Z":
                e = rec.index(b"\0", q)
                if tag == b"MD":
                    md = rec[q:e].decode()
                q = e + 1
            elif typ in "AcC":
                q += 1
            elif typ in "sS":
                q += 2
            elif typ in "iIf":
                q += 4
            elif typ == "H":
                q = rec.index(b"\0", q) + 1
            elif typ == "B":
                sub = chr(rec[q]); n, = struct.unpack_from("<i", rec, q + 1)
                q += 5 + n * {"c": 1, "C": 1, "s": 2, "S": 2, "i": 4, "I": 4, "f": 4}[sub]
            else:
                raise ValueError(typ)
        yield refs[ref_id][0] if ref_id >= 0 else None, pos, flag, cigar, seq, md, refs


def ref_bases_of(pos, cigar, seq, md):
    """(position, base) pairs of the reference this read pins, from CIGAR + MD."""
    qi, aligned = 0, []                                      # per ref-consuming column: query base or None (deletion)
    for op, n in cigar:
        if op in (0, 7, 8):
            aligned += list(seq[qi:qi + n]); qi += n
        elif op in (1, 4):
            qi += n
        elif op == 2:
            aligned += [None] * n
        elif op == 3:
            aligned += ["skip"] * n
    out, i, k = [], 0, 0
    cols = [c for c in aligned if c != "skip"]               # MD does not describe N (ref-skip) columns
    j = 0
    while j < len(md):
        if md[j].isdigit():
            e = j
            while e < len(md) and md[e].isdigit():
                e += 1
            for _ in range(int(md[j:e])):
                out.append(cols[k]); k += 1
            j = e
        elif md[j] == "^":
            j += 1
            while j < len(md) and md[j].isalpha():
                out.append(md[j]); k += 1; j += 1
        else:
            out.append(md[j]); k += 1; j += 1
    assert k == len(cols), (k, len(cols), md)
    res, o = [], iter(out)
    for c in aligned:
        if c != "skip":
            res.append((pos + i, next(o).upper()))
        i += 1
    return res


def main():
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "BaseVarC_ref")
    assert os.path.exists(ref_bin), "make -C oracle ref first"
    with open(os.path.join(REF, "test", "bam.list")) as f:
        bams = [l.strip() for l in f if l.strip()]
    known, conflicts, length = {}, 0, None
    for b in bams:
        for chrom, pos, flag, cigar, seq, md, refs in bam_records(os.path.join(REF, "test", b)):
            if chrom != "chr17" or flag & 4 or md is None:
                continue
            length = dict(refs)["chr17"]
            for p, base in ref_bases_of(pos, cigar, seq, md):
                if known.setdefault(p, base) != base:
                    conflicts += 1
    lo, hi = min(known), max(known) + 1
    print("reference bases recovered: %d positions in [%d, %d), %d conflicts" % (len(known), lo, hi, conflicts))
    assert conflicts == 0
    window = "".join(known.get(p, "N") for p in range(lo, hi))

    work = tempfile.mkdtemp(prefix="bvc_c1_")
    try:
        full = bytearray(b"N" * length)
        full[lo:hi] = window.encode()
        with open(os.path.join(work, "chr17.fa"), "wb") as f:
            f.write(b">chr17\n")
            for i in range(0, length, 60):
                f.write(full[i:i + 60]); f.write(b"\n")
        os.symlink(os.path.join(REF, "test", "data"), os.path.join(work, "data"))
        shutil.copy(os.path.join(REF, "test", "bam.list"), os.path.join(work, "bam.list"))
        cmd = [ref_bin, "basetype", "--rerun", "--keep_tmp"] + ARGS + ["-i", "bam.list", "-s", REGION, "-r", "chr17.fa", "-o", "test.out"]
        r = subprocess.run(cmd, cwd=work, capture_output=True, text=True)
        assert r.returncode == 0 and "basetype done" in r.stdout, r.stderr[-2000:]
        if os.path.isdir(OUT):
            shutil.rmtree(OUT)
        os.makedirs(OUT)
        for t in range(4):
            d = "test.out.tmp.thread.%d" % t
            shutil.copytree(os.path.join(work, d), os.path.join(OUT, d))
        for ext in ("cvg", "vcf"):
            with gzip.open(os.path.join(work, "test.out.%s.gz" % ext), "rb") as f:
                txt = f.read()
            with gzip.open(os.path.join(OUT, "ref.%s.txt.gz" % ext), "wb", compresslevel=9) as f:
                f.write(txt)
            print(ext, txt.count(b"\n"), "lines")
        shutil.copy(os.path.join(REF, "test", "bam.list"), os.path.join(OUT, "bam.list"))
        with gzip.open(os.path.join(OUT, "chr17_window.txt.gz"), "wt", compresslevel=9) as f:
            f.write("%d %d %d\n%s\n" % (length, lo, hi, window))
        with open(os.path.join(OUT, "ARGS"), "w") as f:
            f.write(" ".join(ARGS + ["-s", REGION]) + "\n")
    finally:
        shutil.rmtree(work, ignore_errors=True)
    print("fixture bytes:", sum(os.path.getsize(os.path.join(dp, fn)) for dp, _, fns in os.walk(OUT) for fn in fns))


if __name__ == "__main__":
    main()
