#!/usr/bin/env python
"""Regenerates the golden fixtures of this directory FROM THE REFERENCE ITSELF (run in the build
container, where /root/reference exists and `make -C oracle ref` has built oracle/_ref):

  leaf_kats.json      known answers of the unmodified reference objects (libref_oracle.so): bt_fisher_exact,
                      chisf, normsf, RankSumTest, BaseType::LRT (full double precision, float.hex()).
  synth_ref_*.npz/.txt.gz
                      a small synthetic pileup tile + the decompressed .cvg/.vcf text the unmodified
                      reference BINARY (`BaseVarC_ref basetype --rerun`) printed for it.
The reference ships no golden vectors or known-answer tests for this path, so
these are the pins.  Usage: python tests/golden/make_golden.py
"""
import gzip
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import basevarc_b200 as B  # noqa: E402
import oracle_py as O  # noqa: E402
import refrun  # noqa: E402


def hx(x):
    return float(x).hex() if x == x else "nan"


def leaf_kats():
    R = O.ref()
    rng = np.random.default_rng(20261104)
    out = {"fisher": [], "chisf": [], "normsf": [], "ranksum": [], "lrt": []}
    tabs = [(7, 2, 0, 5), (149, 154, 39, 45), (0, 0, 0, 0), (1, 0, 0, 0), (10, 10, 10, 10), (3, 0, 0, 3), (200, 180, 1, 0),
            (5000, 4500, 30, 2), (11, 22, 33, 44), (0, 12, 12, 0)]
    tabs += [tuple(int(v) for v in rng.integers(0, rng.choice([4, 40, 400, 4000]), 4)) for _ in range(300)]
    for t in tabs:
        out["fisher"].append([list(t), hx(R.ref_fisher(*t))])
    for x in [0.0, 1e-300, 1e-9, 0.5, 1.0, 2.0, 2.0000001, 3.84, 23.9999, 24.0, 60.0, 700.0, 1500.0, 1e4] + rng.uniform(0, 120, 200).tolist():
        out["chisf"].append([hx(x), hx(R.ref_chisf(x, 1.0))])
    for x in [0.0, 0.1, 1.0, 5.0, 7.07, 7.08, 26.0, 27.0, 40.0] + rng.uniform(0, 30, 100).tolist():
        out["normsf"].append([hx(x), hx(R.ref_normsf(x))])
    for i in range(200):
        n1, n2 = (int(v) for v in rng.integers(0 if i % 10 == 0 else 1, 60, 2))
        x = rng.integers(0, int(rng.choice([3, 60, 256])), n1).tolist()
        y = rng.integers(0, 60, n2).tolist()
        out["ranksum"].append([x, y, hx(O.ref_ranksum(x, y))])
    out["ranksum"].append([[30, 30, 31, 35], [20, 30, 36], hx(O.ref_ranksum([30, 30, 31, 35], [20, 30, 36]))])
    # LRT: the survey's smoke cases + random pileups incl. qual 0 (NaN paths) and restricted base_comb (group path)
    cases = [([0] * 30 + [2] * 10 + [1], [30] * 40 + [20], 0, None), ([3] * 20, [30] * 20, 3, None),
             ([1] * 12, [30] * 12, 3, None), ([1], [30], 3, None)]
    for i in range(400):
        n = int(rng.choice([1, 2, 5, 12, 40, 400, 2500]))
        ref = int(rng.integers(0, 4))
        probs = rng.dirichlet([6, 0.4, 0.3, 0.1]) if i % 3 else rng.dirichlet([1, 1, 1, 1])
        bases = ((rng.choice(4, n, p=probs) + ref) % 4).tolist()
        quals = rng.integers(0 if i % 9 == 0 else 1, 42, n).tolist()
        comb = None
        if i % 5 == 0:
            alts = [b for b in range(4) if b != ref][: int(rng.integers(1, 4))]
            comb = [ref] + alts
        cases.append((bases, quals, ref, comb))
    for bases, quals, ref, comb in cases:
        r = O.ref_lrt(bases, quals, ref, 0.001, comb)
        out["lrt"].append(dict(bases="".join(map(str, bases)), quals=quals, ref=ref, comb=comb, ok=r.ok, n_alt=r.n_alt,
                               alt=list(r.alt_bases)[: r.n_alt], depth=list(r.depth), af=[hx(v) for v in r.af_lrt],
                               var_qual=hx(r.var_qual)))
    with open(os.path.join(HERE, "leaf_kats.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("leaf_kats.json:", {k: len(v) for k, v in out.items()})


def group_of(i):
    """The fixed sample -> population assignment of the *_groups fixture (every 4th sample has no group)."""
    return ["EAS", "AFR", "EUR", None][i % 4]


def synth_ref(name, N, S, depth, model, seed, threads, batch, indel_frac, groups=False):
    tile = B.synth_tile(N, S, depth, model, seed=seed, first_pos=5001, indel_frac=indel_frac)
    with tempfile.TemporaryDirectory() as d:
        extra = []
        if groups:
            gf = os.path.join(d, "groups.txt")
            with open(gf, "w") as f:
                f.write("".join("S%d %s\n" % (i, group_of(i)) for i in range(N) if group_of(i)))
            extra = ["-g", gf]
        cvg, vcf = refrun.run_reference(tile, d, threads=threads, batch=batch, extra=extra)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), n_samples=N, ref_base=tile.ref_base, entry_off=tile.entry_off,
                        code=tile.code, qual=tile.qual, mapq=tile.mapq, rpr=tile.rpr, sample=tile.sample, pos=tile.pos)
    for kind, text in (("cvg", cvg), ("vcf", vcf)):
        with gzip.open(os.path.join(HERE, "%s.%s.txt.gz" % (name, kind)), "wt", compresslevel=9) as f:
            f.write(text)
    print(name, "sites", S, "cvg rows", cvg.count("\n") - 3, "vcf records", sum(1 for l in vcf.splitlines() if not l.startswith("#")))


if __name__ == "__main__":
    if O.ref() is None or not refrun.have_ref_bin():
        sys.exit("build the reference first: make -C oracle ref")
    leaf_kats()
    synth_ref("synth_ref_lowpass", 120, 2500, 0.6, "lowpass", 101, 3, 16, 0.01)
    synth_ref("synth_ref_highpoly", 48, 1500, 2.0, "highpoly", 102, 2, 10, 0.03)
    synth_ref("synth_ref_groups", 60, 1200, 3.0, "highpoly", 103, 2, 16, 0.02, groups=True)
