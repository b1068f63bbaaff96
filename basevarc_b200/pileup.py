"""Packed pileup tiles: the structure-of-arrays input of the calling core, and the synthetic
low-pass cohorts BASELINE.json names (generator spec: SURVEY.md section 8(d)).

A tile is CSR over the covered (site, sample) pairs -- the reference keeps at most one read per
sample per site (src/BamProcess.cpp:76,81) -- with one byte each for code / qual / mapq / rpr
(AlleleInfo, src/BamProcess.h:30-39) and the global sample index per entry.
"""
from dataclasses import dataclass, field

import numpy as np

from ._abi import Tile, make_code

# name -> (n_samples, sites of the full configuration, mean depth c, site model)
CONFIGS = {
    "C2": dict(n_samples=1000, n_sites=1_000_000, depth=0.5, model="lowpass"),
    "C3": dict(n_samples=10_000, n_sites=10_000_000, depth=0.3, model="lowpass"),
    "C4": dict(n_samples=100_000, n_sites=1_000_000, depth=0.1, model="lowpass"),
    "C5": dict(n_samples=2000, n_sites=1_000_000, depth=5.0, model="highpoly"),
}


@dataclass
class PileupTile:
    n_samples: int
    ref_base: np.ndarray      # [S] uint8 0..3
    entry_off: np.ndarray     # [S+1] uint32
    code: np.ndarray          # [R] uint8
    qual: np.ndarray          # [R] uint8
    mapq: np.ndarray          # [R] uint8
    rpr: np.ndarray           # [R] uint8
    sample: np.ndarray        # [R] uint32
    pos: np.ndarray = None    # [S] int32 1-based positions (host only)
    indel_text: dict = field(default_factory=dict)   # entry index -> indel token (host only)
    packed8: np.ndarray = None     # [R] uint8 compact wire form of (code, qual), see pack8()
    qual_dict: np.ndarray = None   # [<= 30] uint8

    @property
    def n_sites(self):
        return int(self.ref_base.shape[0])

    @property
    def n_entries(self):
        return int(self.code.shape[0])

    def as_ctypes(self, with_sample=True, drop_code_qual=False):
        """The bvc_tile struct over these arrays (which must outlive the call)."""
        t = Tile()
        t.n_sites = self.n_sites
        for name in ("ref_base", "entry_off", "code", "qual", "mapq", "rpr"):
            a = getattr(self, name)
            assert a.flags["C_CONTIGUOUS"]
            setattr(t, name, a.ctypes.data)
        t.sample = self.sample.ctypes.data if (with_sample and self.sample is not None) else None
        if self.packed8 is not None:
            assert self.packed8.flags["C_CONTIGUOUS"] and self.packed8.shape[0] == self.n_entries
            t.packed8 = self.packed8.ctypes.data
            t.qual_dict = self.qual_dict.ctypes.data
            t.n_qual_dict = int(self.qual_dict.shape[0])
            if drop_code_qual:
                t.code = None
                t.qual = None
        return t

    def pack8(self):
        """Adds the compact wire form (include/bvc_gpu.h BVC_PACK8: base:2 | qi:5 | strand:1 + a dictionary of the tile's
        distinct quals) when the tile qualifies: <= 30 distinct quals among the counted entries and no non-indel entry with
        a base code > 3.  Returns True if packed8 / qual_dict were set."""
        base, strand, indel = self.code & 7, (self.code >> 3) & 1, (self.code >> 4) & 1
        counted = indel == 0
        if (base[counted] > 3).any() or (base[indel == 1] > 4).any():
            return False
        quals = np.unique(self.qual[counted])
        if quals.size > 30:
            return False
        if quals.size == 0:
            quals = np.array([0], np.uint8)
        lut = np.zeros(256, np.uint8)
        lut[quals] = np.arange(quals.size, dtype=np.uint8)
        qi = lut[self.qual].astype(np.uint8)
        qi[indel == 1] = np.where(base[indel == 1] == 4, 30, 31).astype(np.uint8)
        self.packed8 = np.ascontiguousarray(((base & 3) << 6) | (qi << 1) | strand).astype(np.uint8)
        self.qual_dict = np.ascontiguousarray(quals.astype(np.uint8))
        return True

    def slice_sites(self, s0, s1):
        e0, e1 = int(self.entry_off[s0]), int(self.entry_off[s1])
        return PileupTile(
            n_samples=self.n_samples,
            ref_base=np.ascontiguousarray(self.ref_base[s0:s1]),
            entry_off=np.ascontiguousarray(self.entry_off[s0:s1 + 1] - np.uint32(e0)),
            code=np.ascontiguousarray(self.code[e0:e1]), qual=np.ascontiguousarray(self.qual[e0:e1]),
            mapq=np.ascontiguousarray(self.mapq[e0:e1]), rpr=np.ascontiguousarray(self.rpr[e0:e1]),
            sample=np.ascontiguousarray(self.sample[e0:e1]),
            pos=None if self.pos is None else np.ascontiguousarray(self.pos[s0:s1]),
            indel_text={k - e0: v for k, v in self.indel_text.items() if e0 <= k < e1},
        )

    def shard_samples(self, world):
        """Per-rank local tiles of sample sharding: rank r keeps, for EVERY site, the entries of the samples
        [r*N/W, (r+1)*N/W) -- what a rank reads from its own range of batch files."""
        bounds = [(r * self.n_samples) // world for r in range(world + 1)]
        site_of = np.repeat(np.arange(self.n_sites, dtype=np.int64), np.diff(self.entry_off.astype(np.int64)))
        tiles = []
        for r in range(world):
            keep = (self.sample >= bounds[r]) & (self.sample < bounds[r + 1])
            off = np.zeros(self.n_sites + 1, np.int64)
            np.cumsum(np.bincount(site_of[keep], minlength=self.n_sites), out=off[1:])
            idx = np.nonzero(keep)[0]
            tiles.append(PileupTile(
                n_samples=self.n_samples, ref_base=self.ref_base, entry_off=off.astype(np.uint32),
                code=np.ascontiguousarray(self.code[idx]), qual=np.ascontiguousarray(self.qual[idx]),
                mapq=np.ascontiguousarray(self.mapq[idx]), rpr=np.ascontiguousarray(self.rpr[idx]),
                sample=np.ascontiguousarray(self.sample[idx]), pos=self.pos))
        return tiles

    def nbytes(self):
        return sum(getattr(self, k).nbytes for k in ("ref_base", "entry_off", "code", "qual", "mapq", "rpr"))

    @staticmethod
    def concat(tiles):
        offs = [np.zeros(1, np.uint32)]
        base = 0
        for t in tiles:
            offs.append(t.entry_off[1:].astype(np.uint64) + base)
            base += t.n_entries
        assert base < 2 ** 32
        cat = lambda k: np.concatenate([getattr(t, k) for t in tiles])
        return PileupTile(tiles[0].n_samples, cat("ref_base"), np.concatenate(offs).astype(np.uint32),
                          cat("code"), cat("qual"), cat("mapq"), cat("rpr"), cat("sample"),
                          None if tiles[0].pos is None else cat("pos"))


def _site_allele_model(rng, S, model):
    """Per-site reference base and allele-frequency vector over the 4 bases (rows sum to 1)."""
    ref = rng.integers(0, 4, S, dtype=np.uint8)
    af = np.zeros((S, 4), np.float64)
    af[np.arange(S), ref] = 1.0
    if model == "lowpass":
        var = rng.random(S) < (1.0 / 200.0)
        idx = np.nonzero(var)[0]
        alt_af = rng.beta(0.3, 2.0, idx.size)
        alt = (ref[idx] + rng.integers(1, 4, idx.size)) % 4
        af[idx, ref[idx]] = 1.0 - alt_af
        af[idx, alt] = alt_af
    elif model == "highpoly":
        var = rng.random(S) < (1.0 / 5.0)
        idx = np.nonzero(var)[0]
        alt_af = rng.beta(0.3, 2.0, idx.size)
        multi = rng.random(idx.size) < 0.4
        n_minor = np.where(multi, rng.integers(2, 4, idx.size), 1)       # minor alleles: 1, or 2..3
        w = rng.gamma(1.0, 1.0, (idx.size, 3))                            # Dirichlet(1,..) split
        w[np.arange(3)[None, :] >= n_minor[:, None]] = 0.0
        w /= w.sum(1, keepdims=True)
        perm = np.argsort(rng.random((idx.size, 3)), axis=1)              # which non-ref bases get mass
        af[idx, ref[idx]] = 1.0 - alt_af
        for k in range(3):
            b = (ref[idx] + 1 + perm[:, k]) % 4
            af[idx, b] += alt_af * w[:, k]
    else:
        raise ValueError(model)
    return ref, af


def synth_tile(n_samples, n_sites, depth, model="lowpass", seed=1, indel_frac=0.005,
               first_pos=1, chunk_cells=1 << 25, shard=None):
    """Deterministic synthetic pileup (numpy default_rng(seed)), spec in SURVEY.md section 8(d).

    shard = (rank, world): only the entries of the samples [rank*N/world, (rank+1)*N/world) -- one rank's local tile
    of the sample-sharded mode.  The site model (ref base, allele frequencies, coverage) comes from `seed` alone, so
    every rank sees the same sites; the reads come from a per-rank stream."""
    rng = np.random.default_rng(seed)
    S, N = int(n_sites), int(n_samples)
    ref, af = _site_allele_model(rng, S, model)
    p_cov = 1.0 - np.exp(-depth)
    site_cov = np.full(S, p_cov)
    if model == "highpoly":
        site_cov = site_cov * rng.uniform(0.1, 1.0, S)
    n_total, sample0 = N, 0
    if shard is not None:
        r, w = shard
        sample0 = (r * N) // w
        N = ((r + 1) * N) // w - sample0
        rng = np.random.default_rng([int(seed), 7919 + int(r), int(w)])
    cum_af = np.cumsum(af, axis=1)
    is_var = af.max(1) < 1.0
    perr_lut = np.power(10.0, -np.arange(256) / 10.0).astype(np.float32)
    counts = np.empty(S, np.int64)
    parts = []
    rows = max(1, chunk_cells // max(N, 1))
    for s0 in range(0, S, rows):
        s1 = min(S, s0 + rows)
        cov = rng.random((s1 - s0, N), dtype=np.float32) < site_cov[s0:s1, None].astype(np.float32)
        counts[s0:s1] = cov.sum(1)
        site_i, samp = np.nonzero(cov)
        site_i = site_i + s0
        R = site_i.size
        qual = rng.integers(20, 41, R, dtype=np.uint8)
        mapq = rng.integers(20, 61, R, dtype=np.uint8)
        rpr = rng.integers(1, 101, R, dtype=np.uint8)
        strand = rng.integers(0, 2, R, dtype=np.uint8)
        # true allele: REF except at the (rare) variant sites, where it is drawn from the site's AF vector
        true_b = ref[site_i]
        at_var = np.nonzero(is_var[site_i])[0]
        if at_var.size:
            u = rng.random(at_var.size)
            true_b[at_var] = (u[:, None] >= cum_af[site_i[at_var], :3]).sum(1).astype(np.uint8)
        flip = rng.random(R, dtype=np.float32) < perr_lut[qual]
        obs = true_b
        nf = np.nonzero(flip)[0]
        obs[nf] = (true_b[nf] + rng.integers(1, 4, nf.size, dtype=np.uint8)) % 4
        indel = (rng.random(R, dtype=np.float32) < indel_frac).astype(np.uint8)
        parts.append((obs, strand, indel, qual, mapq, rpr, (samp + sample0).astype(np.uint32)))
    cat = lambda k: np.concatenate([p[k] for p in parts]) if parts else np.zeros(0, np.uint8)
    obs, strand, indel, qual, mapq, rpr, samp = (cat(k) for k in range(7))
    if indel.size:
        indel[0] = 0           # the reference's stale fields are uninitialised before the first numeric token
    # Q1: an indel entry carries the fields of the last numeric token parsed before it (forward fill)
    R = obs.size
    idx = np.where(indel == 0, np.arange(R), 0)
    np.maximum.accumulate(idx, out=idx)
    obs, strand, qual, mapq, rpr = obs[idx], strand[idx], qual[idx], mapq[idx], rpr[idx]
    entry_off = np.zeros(S + 1, np.uint64)
    np.cumsum(counts, out=entry_off[1:])
    assert entry_off[-1] < 2 ** 32, "tile too large for uint32 CSR offsets"
    return PileupTile(
        n_samples=n_total, ref_base=ref, entry_off=entry_off.astype(np.uint32),
        code=make_code(obs, strand, indel), qual=qual, mapq=mapq, rpr=rpr, sample=samp.astype(np.uint32),
        pos=(np.arange(S, dtype=np.int32) + np.int32(first_pos)),
    )


def synth_config(name, n_sites=None, seed=None, **kw):
    cfg = CONFIGS[name]
    return synth_tile(cfg["n_samples"], n_sites or cfg["n_sites"], cfg["depth"], cfg["model"],
                      seed=seed if seed is not None else int(name[1:]), **kw)
