"""Run the UNMODIFIED reference binary (oracle/_ref/BaseVarC_ref) on batch files written from a tile,
and parse its .cvg / .vcf text back into numbers (tests + fixture generation only)."""
import os
import subprocess

import numpy as np

from basevarc_b200 import batchfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "BaseVarC_ref")


def have_ref_bin():
    return os.path.exists(REF_BIN)


def run_reference(tile, workdir, threads=2, batch=10, chrom="chrS", extra=(), keep_tmp=False):
    """Returns (cvg_text, vcf_text) of `BaseVarC_ref basetype --rerun` on the tile."""
    os.makedirs(workdir, exist_ok=True)
    prefix = os.path.join(workdir, "out")
    first = int(tile.pos[0])
    batchfile.write_batch_files(tile, prefix, threads, batch)
    fa = os.path.join(workdir, "ref.fa")
    batchfile.write_fasta(fa, chrom, tile.ref_base, first)
    bl = os.path.join(workdir, "bam.list")
    with open(bl, "w") as f:
        f.write("".join("s%d.bam\n" % i for i in range(tile.n_samples)))
    cmd = [REF_BIN, "basetype", "--rerun", "-t", str(threads), "-b", str(batch), "-i", bl,
           "-s", batchfile.region_for(chrom, first, tile.n_sites), "-r", fa, "-o", prefix] + list(extra)
    if keep_tmp:
        cmd.insert(3, "--keep_tmp")
    r = subprocess.run(cmd, cwd=workdir, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("reference failed: %s\n%s" % (r.returncode, r.stderr[-2000:]))
    return batchfile.read_bgzf_text(prefix + ".cvg.gz"), batchfile.read_bgzf_text(prefix + ".vcf.gz")


def parse_cvg(text):
    """pos -> dict(depth[4], dep, fs, sor, strand[4], ref, indels, groups[...])."""
    out = {}
    for ln in text.splitlines():
        if ln.startswith("#"):
            continue
        f = ln.split("\t")
        out[int(f[1])] = dict(ref=f[2], dep=int(f[3]), depth=[int(x) for x in f[4:8]], indels=f[8],
                              fs=float(f[9]), sor=float(f[10]), strand=[int(x) for x in f[11].split(",")],
                              groups=f[12:])
    return out


def parse_vcf(text):
    """pos -> dict(ref, alt[list], qual, filter, info{...}, samples[list])."""
    out = {}
    for ln in text.splitlines():
        if ln.startswith("#"):
            continue
        f = ln.split("\t")
        info = dict(kv.split("=", 1) for kv in f[7].split(";"))
        out[int(f[1])] = dict(ref=f[3], alt=f[4].split(","), qual=float(f[5]), filter=f[6], info=info,
                              samples=f[9:])
    return out


def load_golden_tile(name):
    """tests/golden/<name>.npz -> (PileupTile, cvg_text, vcf_text) written by tests/golden/make_golden.py."""
    import gzip
    from basevarc_b200.pileup import PileupTile
    g = os.path.join(ROOT, "tests", "golden")
    z = np.load(os.path.join(g, name + ".npz"))
    tile = PileupTile(int(z["n_samples"]), z["ref_base"], z["entry_off"], z["code"], z["qual"], z["mapq"], z["rpr"],
                      z["sample"], z["pos"])
    texts = []
    for kind in ("cvg", "vcf"):
        with gzip.open(os.path.join(g, "%s.%s.txt.gz" % (name, kind)), "rt") as f:
            texts.append(f.read())
    return tile, texts[0], texts[1]


def _feq(text_val, x, digits):
    """A printed `{:.<digits>f}` value against a double: equal after rounding, +-1 unit in the last digit."""
    t = float(text_val)
    if t != t or x != x:
        return t != t and x != x
    return abs(t - x) <= 1.01 * 10.0 ** (-digits) * max(1.0, abs(t)) * 0.5 + 1e-12 or ("%.*f" % (digits, x)) == text_val


def compare_records_to_text(tile, rec, cvg_text, vcf_text):
    """Record array (oracle or GPU) vs the reference binary's printed output.  Integers and the emitted
    site set / ALT alleles exactly; floats to the printed digits.  Returns a list of mismatch strings."""
    C, V = parse_cvg(cvg_text), parse_vcf(vcf_text)
    bad = []
    n_cvg = 0
    for s in range(tile.n_sites):
        p, r = int(tile.pos[s]), rec[s]
        if r["n_entries"] == 0:
            if p in C:
                bad.append("site %d: reference printed a CVG row for an empty site" % p)
            continue
        n_cvg += 1
        if p not in C:
            bad.append("site %d: missing in reference CVG" % p)
            continue
        c = C[p]
        ok = c["depth"] == r["depth"].tolist() and c["dep"] == int(r["depth"].sum())
        ok &= c["strand"] == r["cvg_strand"].tolist()
        ok &= _feq(c["fs"], r["cvg_fs"], 3) and _feq(c["sor"], r["cvg_sor"], 3)
        ok &= (p in V) == bool(r["lrt_ok"])
        if ok and p in V:
            v = V[p]
            alts = r["alt_bases"][: r["n_alt"]]
            ok &= v["alt"] == ["ACGT"[b] for b in alts]
            ok &= _feq(v["qual"], r["var_qual"], 2)
            ok &= v["filter"] == ("." if r["var_qual"] > 60 else "LowQual")
            ok &= all(_feq(a, r["af_lrt"][b], 6) for a, b in zip(v["info"]["CM_AF"].split(","), alts))
            ok &= v["info"]["CM_AC"] == ",".join(str(r["depth"][b]) for b in alts)
            for k, i in (("MQRankSum", 0), ("BaseQRankSum", 1), ("ReadPosRankSum", 2)):
                ok &= _feq(v["info"][k], r["ranksum"][i], 3)
            ok &= v["info"]["SB_REF"] == "%d,%d" % tuple(r["vcf_strand"][:2])
            ok &= v["info"]["SB_ALT"] == "%d,%d" % tuple(r["vcf_strand"][2:])
            ok &= _feq(v["info"]["FS"], r["vcf_fs"], 3) and _feq(v["info"]["SOR"], r["vcf_sor"], 3)
            ad = sum(int(r["depth"][b]) for b in alts)
            ok &= _feq(v["info"]["QD"], r["var_qual"] / ad, 3)
        if not ok:
            bad.append("site %d differs: ref cvg=%r vcf=%r rec=%r" % (p, c, V.get(p, {}).get("info"), r))
    if n_cvg != len(C):
        bad.append("CVG row count: reference %d vs records %d" % (len(C), n_cvg))
    if len(V) != int(rec["lrt_ok"].sum()):
        bad.append("VCF record count: reference %d vs records %d" % (len(V), int(rec["lrt_ok"].sum())))
    return bad
