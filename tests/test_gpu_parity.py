"""GPU tier (B200): the CUDA path through the C ABI vs the golden fixtures, vs the CPU oracle on seeded inputs, and
size-independent properties at BASELINE.json's shapes.  Integer tallies are bit-exact; fp64 results within 1e-12
relative (the tolerance BASELINE.json:north_star states)."""
from __future__ import annotations

import numpy as np
import pytest

import helpers
from speq_b200 import capi, driver, engine, synth

pytestmark = pytest.mark.gpu
TOL = 1e-12


def _reads(seqs, quals):
    seq, q, offs = helpers.flat_reads(seqs, quals)
    return driver.Reads(seq, q, offs)


def _compare(out: dict, exp: dict, V: int):
    assert [int(x) for x in out["unique_ref"]] == exp["unique_ref"]
    assert [int(x) for x in out["total_ref"]] == exp["total_ref"]
    assert out["total_kmers"] == exp["total_kmers"]
    assert out["ambiguous"] == exp["ambiguous"]
    assert out["lookups"] == exp["lookups"]
    assert out["n"] == exp["n"]
    if exp["mode"] == "global":
        assert [int(x) for x in out["unique"]] == exp["unique"]
        assert list(out["unique_totals"]) == exp["unique_totals"]        # one division per entry: bit-exact
        assert list(out["percent_initial"]) == exp["percent_initial"]
    else:
        assert helpers.rel_close(out["unique_w"], exp["unique_w"], TOL)
        assert helpers.rel_close(out["percent_initial"], exp["percent_initial"], TOL)
    assert out["iterations"] == exp["iterations"]
    assert helpers.rel_close(out["percent_final"], exp["percent_final"], TOL)
    for a, b in zip(out["iter_tkpg"], exp["iter_tkpg"]):
        assert helpers.rel_close(a, b, 1e-11)


@pytest.mark.parametrize("name", helpers.GOLDEN_NAMES + helpers.GOLDEN_LONG_K)
@pytest.mark.parametrize("raw", [True, False])
def test_golden(lib, name, raw):
    """The rolling FM search (the path for every k) against the fixtures; the k > 31 fixtures (k = 33, 70, 127; k = 70 in local
    mode with the default phred cutoff is the reference's default invocation) exercise e-bit contraction and parent jumps at
    depths up to 126."""
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    out = driver.run_all(scaff, gs, counts, _reads(seqs, quals), k=run["k"], phred_cutoff=run.get("phred_cutoff", 30),
                         fixed_accuracy=run["fixed_accuracy"], paired=run["paired"], raw=raw, batch_reads=64,
                         algorithm=capi.ALGO_ROLLING)
    _compare(out, exp, len(counts))


@pytest.mark.parametrize("name", helpers.GOLDEN_NAMES)
@pytest.mark.parametrize("raw", [True, False])
@pytest.mark.parametrize("algorithm", [capi.ALGO_KTABLE, capi.ALGO_KPAIR])
def test_golden_ktable_algorithm(lib, name, raw, algorithm):
    """The k-mer table scan (one probe per window) against the same fixtures, all modes, with the refinement running on
    multiplicities recorded per table slot."""
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    out = driver.run_all(scaff, gs, counts, _reads(seqs, quals), k=run["k"], phred_cutoff=run.get("phred_cutoff", 30),
                         fixed_accuracy=run["fixed_accuracy"], paired=run["paired"], raw=raw, batch_reads=64,
                         algorithm=algorithm)
    _compare(out, exp, len(counts))


@pytest.mark.parametrize("name", helpers.GOLDEN_LONG_K)
def test_golden_long_k_auto(lib, name):
    """SPEQG_ALGO_AUTO at k > 31 resolves to the rolling FM search (scan32_kernel): the product path of the reference's
    default k = 70.  Same fixtures through AUTO, raw submits in small batches."""
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    idx = engine.Index.from_scaffolds(scaff, gs, len(counts))
    probe = engine.Scan(idx, k=run["k"])
    assert probe.info()["algorithm"] == capi.ALGO_ROLLING
    probe.close()
    out = driver.run_all(scaff, gs, counts, _reads(seqs, quals), k=run["k"], phred_cutoff=run.get("phred_cutoff", 30),
                         fixed_accuracy=run["fixed_accuracy"], paired=run["paired"], raw=True, batch_reads=32,
                         algorithm=capi.ALGO_AUTO, index=idx)
    _compare(out, exp, len(counts))
    idx.close()


@pytest.mark.parametrize("name", ["global_single", "global_paired", "refs_with_n", "short_k", "k70_global_single", "k33_global_paired"])
def test_golden_direct_algorithm(lib, name):
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    out = driver.run_all(scaff, gs, counts, _reads(seqs, quals), k=run["k"], fixed_accuracy=run["fixed_accuracy"],
                         paired=run["paired"], algorithm=capi.ALGO_DIRECT)
    _compare(out, exp, len(counts))


def test_reference_work_counters(lib):
    """speqg_scan_stats reproduces the oracle's reference-semantics counters (steps incl. the failing one, hits, occ)."""
    fx = helpers.load_golden("global_single")
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    idx = engine.Index.from_scaffolds(scaff, gs, len(counts))
    seq, q, offs = helpers.flat_reads(seqs, quals)
    pb = engine.pack_reads(seq, q, offs, n_only=True)
    db = engine.DeviceBatch(0, pb)
    sc = engine.Scan(idx, k=run["k"])
    st = sc.stats(db)
    assert st["lookups"] == exp["lookups"]
    assert st["steps_executed"] == exp["steps_executed"]
    assert st["kmers_hit"] == exp["kmers_hit"]
    assert st["occ"] == exp["occ"]
    assert st["occ_sectors"] == exp["occ_sectors"]
    sc.close(); db.close(); idx.close()


@pytest.mark.parametrize("algorithm", [capi.ALGO_ROLLING, capi.ALGO_KTABLE, capi.ALGO_KPAIR])
@pytest.mark.parametrize("mode", ["global", "local", "global_paired"])
def test_vs_oracle_c1_like(lib, oracle_bin, tmp_path, mode, algorithm):
    """C1-shaped input (2 variants x 3 isolates x 30 kb, 150 bp reads, k=21) at a read count the oracle finishes in seconds."""
    sh = synth.CONFIGS["C1"]
    refs = synth.make_refs(sh, synth.SEED0 + 1)
    R = 6000
    seq, qual = synth.make_reads(sh, refs, R, synth.SEED0 + 101, local_quals=(mode == "local"), n_rate=0.001)
    seq, qual = seq.numpy(), qual.numpy()
    synth.write_fasta(str(tmp_path / "refs.fa"), refs)
    synth.write_groups(str(tmp_path / "groups.txt"), sh)
    paired = mode.endswith("paired")
    if paired:
        synth.write_fastq(str(tmp_path / "r1.fq"), seq[0::2], qual[0::2])
        synth.write_fastq(str(tmp_path / "r2.fq"), seq[1::2], qual[1::2])
    else:
        synth.write_fastq(str(tmp_path / "r1.fq"), seq, qual)
    acc = 0.0 if mode == "local" else sh.accuracy
    cutoff = 19 if mode == "local" else 30
    exp = helpers.run_oracle(oracle_bin, tmp_path / "refs.fa", tmp_path / "groups.txt", tmp_path / "r1.fq",
                             reads2=(tmp_path / "r2.fq") if paired else None, k=sh.k, fixed_accuracy=acc, phred_cutoff=cutoff, threads=8)
    offs = (np.arange(R + 1, dtype=np.uint64) * sh.L)
    out = driver.run_all(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), [sh.I] * sh.V,
                         driver.Reads(seq.reshape(-1), qual.reshape(-1), offs), k=sh.k, phred_cutoff=cutoff, fixed_accuracy=acc,
                         paired=paired, batch_reads=2048, algorithm=algorithm)
    _compare(out, exp, sh.V)


@pytest.mark.parametrize("algorithm", [capi.ALGO_ROLLING, capi.ALGO_KTABLE, capi.ALGO_KPAIR])
def test_edge_cases(lib, algorithm):
    scaff = [engine.ascii_to_ranks("ACGTTGCAAGGCTTAACCGGTTAGCATCGA"), engine.ascii_to_ranks("ACGTTGCATGGCTTAACCGGTAAGCATCGA")]
    idx = engine.Index.from_scaffolds(scaff, [0, 1], 2)
    sc = engine.Scan(idx, k=8, algorithm=algorithm)
    # no reads at all; then empty reads, reads shorter than k, read exactly k long, all-N read
    r = sc.finish()
    assert r.total_kmers == 0 and r.lookups == 0 and r.reads == 0
    seqs = ["", "ACG", "ACGTTGCA", "NNNNNNNNNNNN", "ACGTTGCANGGCTTAACC"]
    quals = ["", "III", "IIIIIIII", "IIIIIIIIIIII", "IIIIIIIIIIIIIIIIII"]
    seq, q, offs = helpers.flat_reads(seqs, quals)
    sc.submit_raw(seq, q, offs, len(seqs))
    r = sc.finish()
    assert r.reads == 5 and r.lookups == 0 + 0 + 1 + 5 + 11
    assert r.total_kmers == 1 + 0 + (1 + 2)       # windows without N: ACGTTGCA | [0..8) and [9..17),[10..18) of the last read
    # ACGTTGCA occurs in both variants -> not unique; GGCTTAAC / GCTTAACC occur in both as well
    assert list(r.unique) == [0, 0]
    sc.reset()
    seq, q, offs = helpers.flat_reads(["AAGGCTTAACCGGTTAGC"], ["I" * 18])
    sc.submit_raw(seq, q, offs, 1)
    r = sc.finish()
    assert r.total_kmers == 11 and int(r.unique[0]) > 0 and int(r.unique[1]) == 0
    # illegal letter -> parse_error surfaced at finish (device packer)
    sc.reset()
    seq, q, offs = helpers.flat_reads(["ACGTTGCAXGGCTTAACC"], ["I" * 18])
    sc.submit_raw(seq, q, offs, 1)
    with pytest.raises(capi.SpeqError) as ei:
        sc.finish()
    assert "parse_error" in ei.value.message
    # paired mode needs an even number of reads; local mode needs qualities
    sp = engine.Scan(idx, k=8, paired=True, algorithm=algorithm)
    with pytest.raises(capi.SpeqError):
        sp.submit_raw(seq, q, offs, 1)
    sl = engine.Scan(idx, k=8, mode=capi.MODE_LOCAL, algorithm=algorithm)
    with pytest.raises(capi.SpeqError):
        sl.submit_raw(seq, None, offs, 1)
    with pytest.raises(capi.SpeqError):
        engine.Scan(idx, k=0)
    with pytest.raises(capi.SpeqError):
        engine.Scan(idx, k=128)
    with pytest.raises(capi.SpeqError):
        engine.Scan(idx, k=32, algorithm=capi.ALGO_KTABLE)      # the table covers k <= 31; AUTO falls back to rolling
    engine.Scan(idx, k=32, algorithm=capi.ALGO_AUTO).close()
    with pytest.raises(capi.SpeqError):
        sc.refine(np.ones(2), np.ones(2, np.int32))     # opened without record_multiplicity -> ESTATE
    for s in (sc, sp, sl):
        s.close()
    idx.close()


def _scan_tallies(idx, k, seq, qual, offs, **kw):
    sc = engine.Scan(idx, k=k, **kw)
    sc.submit_raw(seq, qual, offs, offs.size - 1)
    r = sc.finish()
    sc.close()
    return r


@pytest.mark.parametrize("cfg,R", [("C1", 100_000), ("C2", 400_000), ("C3", 200_000), ("C4", 200_000), ("C5", 300)])
def test_properties_at_config_shapes(lib, cfg, R):
    """Size-independent properties on the full-size indexes of BASELINE.json's configs (read count bounded for test time):
    rolling == direct; shard additivity; lookups = R (L-k+1); error-free reads from a scaffold always hit."""
    import torch
    sh = synth.CONFIGS[cfg]
    dev = "cuda"
    refs = synth.make_refs(sh, synth.SEED0 + int(cfg[1]), device=dev)
    idx = engine.Index.from_scaffolds(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), sh.V)
    seq, qual = synth.make_reads(sh, refs, R, 77, n_rate=0.0005)
    seq, qual = seq.cpu().numpy().reshape(-1), qual.cpu().numpy().reshape(-1)
    offs = np.arange(R + 1, dtype=np.uint64) * sh.L
    a = _scan_tallies(idx, sh.k, seq, qual, offs)
    probe = engine.Scan(idx, k=sh.k)
    # AUTO, global mode: the k-mer table in walk form, also where it cannot stay in L2 (C4: positions in a side array + filter)
    assert probe.info()["algorithm"] == capi.ALGO_KTABLE
    probe.close()
    b = _scan_tallies(idx, sh.k, seq, qual, offs, algorithm=capi.ALGO_DIRECT)
    assert a.lookups == R * sh.windows_per_read == b.lookups
    assert a.total_kmers == b.total_kmers and a.ambiguous == b.ambiguous
    assert np.array_equal(a.unique, b.unique)
    assert a.unique.sum() > 0
    c = _scan_tallies(idx, sh.k, seq, qual, offs, algorithm=capi.ALGO_ROLLING64)
    assert np.array_equal(a.unique, c.unique) and a.total_kmers == c.total_kmers and a.ambiguous == c.ambiguous
    # a ran the default (AUTO: the k-mer table for k <= 31); the rolling FM search and the explicit table must agree
    for alg in (capi.ALGO_ROLLING, capi.ALGO_KTABLE, capi.ALGO_KPAIR):
        d = _scan_tallies(idx, sh.k, seq, qual, offs, algorithm=alg)
        assert np.array_equal(a.unique, d.unique) and a.total_kmers == d.total_kmers and a.ambiguous == d.ambiguous
        assert d.lookups == a.lookups and d.reads == a.reads
    h = R // 2
    p1 = _scan_tallies(idx, sh.k, seq, qual, offs[:h + 1])
    p2 = _scan_tallies(idx, sh.k, seq, qual, offs[h:])
    assert np.array_equal(p1.unique + p2.unique, a.unique)
    assert p1.total_kmers + p2.total_kmers == a.total_kmers and p1.ambiguous + p2.ambiguous == a.ambiguous
    # error-free reads: every window occurs, so total == lookups and refine's recorded windows == total
    import dataclasses
    sh0 = dataclasses.replace(sh, e=0.0)
    n0 = max(64, R // 8)
    s0, q0 = synth.make_reads(sh0, refs, n0, 78)
    sc = engine.Scan(idx, k=sh.k, record=True, algorithm=capi.ALGO_ROLLING)
    sc.submit_raw(s0.cpu().numpy().reshape(-1), q0.cpu().numpy().reshape(-1), np.arange(n0 + 1, dtype=np.uint64) * sh.L, n0)
    r0 = sc.finish()
    assert r0.total_kmers == r0.lookups == n0 * sh.windows_per_read
    # refinement estimate conserves mass: sum_g tkpg[g] == number of recorded (hit) windows
    tk = sc.refine(np.full(sh.V, 100.0 / sh.V), np.full(sh.V, sh.I, np.int32))
    assert abs(tk.sum() - r0.total_kmers) <= 1e-9 * r0.total_kmers
    # the same through the k-mer table: multiplicities recorded per slot, folded to interval starts before the estimate
    for alg in (capi.ALGO_KTABLE, capi.ALGO_KPAIR):
        sk = engine.Scan(idx, k=sh.k, record=True, algorithm=alg)
        sk.submit_raw(s0.cpu().numpy().reshape(-1), q0.cpu().numpy().reshape(-1), np.arange(n0 + 1, dtype=np.uint64) * sh.L, n0)
        rk = sk.finish()
        assert rk.total_kmers == r0.total_kmers and np.array_equal(rk.unique, r0.unique)
        tkk = sk.refine(np.full(sh.V, 100.0 / sh.V), np.full(sh.V, sh.I, np.int32))
        assert helpers.rel_close(tkk, tk, 1e-11)
        assert sk.multiplicity_count() == sc.multiplicity_count()       # the same set of SA intervals was recorded
        sk.close()
    sc.close()
    idx.close()
    del refs
    torch.cuda.empty_cache()


def test_backcheck_idempotent_and_index_roundtrip(lib, tmp_path):
    sh = synth.CONFIGS["C1"]
    refs = synth.make_refs(sh, 5)
    idx = engine.Index.from_scaffolds(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), sh.V)
    u1, t1 = engine.backcheck(idx, sh.k)
    u2, t2 = engine.backcheck(idx, sh.k)
    assert np.array_equal(u1, u2) and np.array_equal(t1, t2)
    assert list(t1) == [2 * sh.I * (sh.G - sh.k + 1)] * sh.V
    idx.save(str(tmp_path / "c1.idx"))
    idx2 = engine.Index.load(str(tmp_path / "c1.idx"))
    u3, t3 = engine.backcheck(idx2, sh.k)
    assert np.array_equal(u1, u3) and np.array_equal(t1, t3)
    idx.close(); idx2.close()


@pytest.mark.gpu
def test_many_variants_u16_groups(lib):
    """More than 255 variants: the scan kernel reads u16 groups from the separate array instead of the bytes inside its
    blocks.  rolling == direct == rolling64."""
    rng = np.random.default_rng(5)
    V = 300
    anc = rng.integers(0, 4, 400)
    scaff = []
    for v in range(V):
        s = anc.copy()
        pos = rng.integers(0, s.size, 12)
        s[pos] = rng.integers(0, 4, 12)
        scaff.append(np.where(s == 3, 4, s).astype(np.uint8))     # dna5 ranks: A0 C1 G2 N3 T4
    idx = engine.Index.from_scaffolds(scaff, list(range(V)), V)
    R, L = 3000, 80
    src = rng.integers(0, V, R)
    st = rng.integers(0, anc.size - L, R)
    letters = np.frombuffer(b"ACGNT", np.uint8)
    seq = np.concatenate([letters[scaff[int(v)][int(p):int(p) + L]] for v, p in zip(src, st)])
    flip = rng.random(seq.size) < 0.01
    seq[flip] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(flip.sum()))]
    qual = np.full(seq.size, ord("I"), np.uint8)
    offs = np.arange(R + 1, dtype=np.uint64) * L
    a = _scan_tallies(idx, 15, seq, qual, offs)
    b = _scan_tallies(idx, 15, seq, qual, offs, algorithm=capi.ALGO_DIRECT)
    c = _scan_tallies(idx, 15, seq, qual, offs, algorithm=capi.ALGO_ROLLING64)
    d = _scan_tallies(idx, 15, seq, qual, offs, algorithm=capi.ALGO_KTABLE)
    e = _scan_tallies(idx, 15, seq, qual, offs, algorithm=capi.ALGO_ROLLING)
    with pytest.raises(capi.SpeqError):
        engine.Scan(idx, k=15, algorithm=capi.ALGO_KPAIR)          # one byte per k-mer: at most 254 variants
    assert a.unique.sum() > 0
    for o in (b, c, d, e):
        assert np.array_equal(a.unique, o.unique) and a.total_kmers == o.total_kmers and a.ambiguous == o.ambiguous
    idx.close()


@pytest.mark.parametrize("k", [1, 2, 16, 24, 31])
def test_ktable_k_range(lib, k):
    """The k-mer table at the ends of its k range (k = 31 stores 45 remainder bits in a table of >= 2^17 buckets; k = 1, 2
    have fewer k-mers than one bucket has slots) against the rolling FM search and the from-the-root search, with Ns in
    the references and in the reads, in single and paired mode."""
    rng = np.random.default_rng(100 + k)
    V, G = 4, 3000
    anc = rng.integers(0, 4, G)
    scaff = []
    for v in range(V):
        s = anc.copy()
        pos = rng.integers(0, G, 60)
        s[pos] = rng.integers(0, 4, 60)
        s = np.where(s == 3, 4, s).astype(np.uint8)              # dna5 ranks: A0 C1 G2 N3 T4
        s[rng.integers(0, G, 3)] = 3
        scaff.append(s)
    idx = engine.Index.from_scaffolds(scaff, list(range(V)), V)
    R, L = 4000, 100
    src, st = rng.integers(0, V, R), rng.integers(0, G - L, R)
    letters = np.frombuffer(b"ACGNT", np.uint8)
    seq = np.concatenate([letters[scaff[int(v)][int(p):int(p) + L]] for v, p in zip(src, st)])
    flip = rng.random(seq.size) < 0.01
    seq[flip] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(flip.sum()))]
    qual = np.where(rng.random(seq.size) < 0.01, ord("#"), ord("I")).astype(np.uint8)
    offs = np.arange(R + 1, dtype=np.uint64) * L
    for paired in (False, True):
        p = _scan_tallies(idx, k, seq, qual, offs, algorithm=capi.ALGO_KPAIR, paired=paired)
        a = _scan_tallies(idx, k, seq, qual, offs, algorithm=capi.ALGO_KTABLE, paired=paired)
        assert np.array_equal(a.unique, p.unique) and a.total_kmers == p.total_kmers and a.ambiguous == p.ambiguous
        assert p.probes > 0 and a.probes > 0        # (the walking table scan probes for a fraction of the windows only)
        b = _scan_tallies(idx, k, seq, qual, offs, algorithm=capi.ALGO_ROLLING, paired=paired)
        c = _scan_tallies(idx, k, seq, qual, offs, algorithm=capi.ALGO_DIRECT, paired=paired)
        assert a.total_kmers > 0
        for o in (b, c):
            assert np.array_equal(a.unique, o.unique) and a.total_kmers == o.total_kmers and a.ambiguous == o.ambiguous
            assert a.lookups == o.lookups
    idx.close()


@pytest.mark.parametrize("mode", ["global", "local", "record"])
def test_long_reads_warp_per_read(lib, mode):
    """Batches whose reads average >= 1024 bases are scanned by a warp per read (scanh.cu, WPR): each lane takes a share of
    the windows and the read's ambiguity state is merged across the warp.  Against the rolling FM search (a lane per read),
    with Ns, low qualities, reads shorter than k and reads of fewer windows than a warp has lanes mixed in, single and
    paired."""
    rng = np.random.default_rng(41)
    V, G, k = 5, 40_000, 15
    anc = rng.integers(0, 4, G)
    scaff = []
    for v in range(V):
        s = anc.copy()
        pos = rng.integers(0, G, 400)
        s[pos] = rng.integers(0, 4, 400)
        s = np.where(s == 3, 4, s).astype(np.uint8)              # dna5 ranks: A0 C1 G2 N3 T4
        s[rng.integers(0, G, 5)] = 3
        scaff.append(s)
    idx = engine.Index.from_scaffolds(scaff, list(range(V)), V)
    R = 600
    lens = rng.integers(1500, 9000, R)
    lens[rng.integers(0, R, 20)] = rng.integers(0, 40, 20)       # empty reads, reads below k, reads of a few windows
    lens[5] = k; lens[6] = k + 31; lens[7] = k + 32
    assert lens.sum() >= 1024 * R
    src, st = rng.integers(0, V, R), np.array([rng.integers(0, G - int(n) + 1) for n in lens])
    letters = np.frombuffer(b"ACGNT", np.uint8)
    # odd reads are chimeras of two variants: ambiguity has to be seen across lanes
    parts = []
    for i, (v, p, n) in enumerate(zip(src, st, lens)):
        r = scaff[int(v)][int(p):int(p) + int(n)].copy()
        if i % 2 and n > 200:
            w = scaff[int((v + 1) % V)]
            r[int(n) // 2:] = w[int(p) + int(n) // 2:int(p) + int(n)]
        parts.append(letters[r])
    seq = np.concatenate(parts)
    flip = rng.random(seq.size) < 0.005
    seq[flip] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(flip.sum()))]
    qual = np.where(rng.random(seq.size) < 0.003, ord("#"), ord("I")).astype(np.uint8)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    kw = dict(mode=capi.MODE_LOCAL if mode == "local" else capi.MODE_GLOBAL, record=mode == "record")
    pct, cnt = np.full(V, 100.0 / V), np.full(V, 1, np.int32)

    def run(algorithm, paired):
        sc = engine.Scan(idx, k=k, algorithm=algorithm, paired=paired, **kw)
        sc.submit_raw(seq, qual, offs, R)
        r = sc.finish()
        extra = (sc.multiplicity_count(), sc.refine(pct, cnt)) if mode == "record" else None
        sc.close()
        return r, extra

    for paired in (False, True):
        b, eb = run(capi.ALGO_ROLLING, paired)
        for alg in (capi.ALGO_KTABLE, capi.ALGO_KPAIR):                 # both table kernels have the warp-per-read form
            a, ea = run(alg, paired)
            assert a.total_kmers > 0 and 0 < a.ambiguous < (R // 2 if paired else R)
            assert a.total_kmers == b.total_kmers and a.ambiguous == b.ambiguous and a.lookups == b.lookups
            assert a.reads == b.reads
            if mode == "local":
                assert helpers.rel_close(a.unique_w, b.unique_w, 1e-12)
            else:
                assert np.array_equal(a.unique, b.unique)
            if mode == "record":
                assert ea[0] == eb[0]
                assert helpers.rel_close(ea[1], eb[1], 1e-11)
    idx.close()


@pytest.mark.parametrize("algorithm", [capi.ALGO_ROLLING, capi.ALGO_KTABLE, capi.ALGO_KPAIR])
@pytest.mark.parametrize("paired", [False, True])
def test_fusion_counts(lib, algorithm, paired):
    """speqg_scan_fusion_counts: ambiguous units by (smallest, largest) unique variant.  Checked against the same scan run on
    every unit alone (the variants with a non-zero unique tally are the unit's set); long reads take the warp-per-read form."""
    rng = np.random.default_rng(5)
    V, G, k = 4, 6000, 15
    scaff = [np.where((s := rng.integers(0, 4, G)) == 3, 4, s).astype(np.uint8) for _ in range(V)]   # unrelated variants
    idx = engine.Index.from_scaffolds(scaff, list(range(V)), V)
    letters = np.frombuffer(b"ACGNT", np.uint8)
    for L, R in ((120, 240), (3000, 40)):
        parts = []
        for i in range(R):
            a, b, c = rng.integers(0, V, 3)
            p = int(rng.integers(0, G - L))
            r = scaff[int(a)][p:p + L].copy()
            if i % 3 == 1:                                            # chimera of two variants
                r[L // 2:] = scaff[int(b)][p + L // 2:p + L]
            if i % 3 == 2:                                            # three variants
                r[L // 3:2 * L // 3] = scaff[int(b)][p + L // 3:p + 2 * L // 3]
                r[2 * L // 3:] = scaff[int(c)][p + 2 * L // 3:p + L]
            parts.append(letters[r])
        seq = np.concatenate(parts)
        qual = np.full(seq.size, ord("I"), np.uint8)
        offs = np.arange(R + 1, dtype=np.uint64) * L
        sc = engine.Scan(idx, k=k, algorithm=algorithm, paired=paired)
        sc.submit_raw(seq, qual, offs, R)
        r = sc.finish()
        fus = sc.fusion_counts()
        assert int(fus.sum()) == r.ambiguous > 0
        assert np.all(np.tril(fus) == 0)                              # smallest < largest
        sets = sc.fusion_sets()
        assert sum(sets.values()) == r.ambiguous and any(len(t) >= 3 for t in sets)      # units with three variants keep all three
        exp = np.zeros((V, V), np.uint64)
        exp_sets = {}
        step = 2 if paired else 1
        for u in range(0, R, step):
            sc.reset()
            sc.submit_raw(seq, qual, offs[u:u + step + 1], step)
            ru = sc.finish()
            nz = np.nonzero(ru.unique)[0]
            if nz.size > 1:
                exp[nz[0], nz[-1]] += 1
                t = tuple(int(g) for g in nz)
                exp_sets[t] = exp_sets.get(t, 0) + 1
        assert np.array_equal(fus, exp)
        assert sets == exp_sets
        sc.close()
    idx.close()


def test_device_counters_reduce_in_place(lib):
    """speqg_scan_device_counters: the counters a multi-GPU driver sums across ranks in place before finish().  One GPU: the
    views hold what finish() returns, and an in-place update on the scan's stream (here: doubling, what an all-reduce over two
    identical ranks yields) is what finish() then reports."""
    import torch
    fx = helpers.load_golden("global_single")
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    V = len(counts)
    idx = engine.Index.from_scaffolds(scaff, gs, V)
    seq, qual, offs = helpers.flat_reads(seqs, quals)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        sc = engine.Scan(idx, k=fx["run"]["k"], phred_cutoff=fx["run"].get("phred_cutoff", 30), stream=st.cuda_stream)
        sc.submit_raw(seq, qual, offs, offs.size - 1)
        ci, cf = sc.device_counters()
        assert ci.dtype == torch.int64 and ci.numel() == 2 * V + 8 and cf.dtype == torch.float64 and cf.numel() == V
        ci.mul_(2)
        r = sc.finish()
    exp = fx["expected"]
    assert [int(x) for x in r.unique] == [2 * x for x in exp["unique"]]
    assert r.total_kmers == 2 * exp["total_kmers"] and r.ambiguous == 2 * exp["ambiguous"] and r.lookups == 2 * exp["lookups"]
    assert [int(x) for x in ci[:V].cpu()] == [2 * x for x in exp["unique"]]
    sc.close()
    idx.close()


@pytest.mark.parametrize("case", ["random", "repeats", "collection"])
def test_device_suffix_array_matches_host(lib, case, tmp_path):
    """SPEQG_BUILD_DEVICE_SA: prefix doubling on the GPU yields the suffix array SA-IS builds on the host (it is unique), hence
    the same index arrays; inputs with long repeats need many doubling rounds."""
    rng = np.random.default_rng(7)
    if case == "random":
        texts = [rng.integers(0, 5, 5000).astype(np.uint8), rng.integers(0, 5, 1).astype(np.uint8), np.zeros(0, np.uint8)]
    elif case == "repeats":
        unit = rng.integers(0, 5, 37).astype(np.uint8)
        texts = [np.tile(unit, 300), np.zeros(4000, np.uint8), np.tile(unit, 150)[5:]]
    else:
        sh = synth.CONFIGS["C1"]
        refs = synth.make_refs(sh, 9)
        texts = []
        for s in synth.refs_to_ranks(refs):
            texts += [s, engine.revcomp_ranks(s)]
    grp = [i % 3 for i in range(len(texts))]
    a = engine.Index.from_texts(texts, grp, 3, keep_sa=True)
    b = engine.Index.from_texts(texts, grp, 3, keep_sa=True, device_sa=True)
    ea, eb = a.export(sa=True), b.export(sa=True)
    for key in ea:
        assert np.array_equal(ea[key], eb[key]), key
    # everything else the index stores (rank blocks, previous / next-smaller links, packed texts): the files are identical
    a.save(str(tmp_path / "host.idx")); b.save(str(tmp_path / "device.idx"))
    assert (tmp_path / "host.idx").read_bytes() == (tmp_path / "device.idx").read_bytes()
    a.close(); b.close()


def _fastq_text(seqs, quals, nl="\n", trailing=True):
    recs = [nl.join([f"@r{i} some id", s, "+", q]) for i, (s, q) in enumerate(zip(seqs, quals))]
    return (nl.join(recs) + (nl if trailing else "")).encode("latin-1")


@pytest.mark.parametrize("algorithm", [capi.ALGO_ROLLING, capi.ALGO_KTABLE, capi.ALGO_KPAIR])
@pytest.mark.parametrize("name", ["global_single", "local_single", "refs_with_n"])
def test_submit_fastq_device_parser(lib, name, algorithm):
    """speqg_scan_submit_fastq: the raw FASTQ bytes go to the device, where the 4-line record grammar runs as kernels (line
    starts, record checks, packer reading in place).  Same tallies as the fixture (oracle/_ref) for LF, CRLF, no trailing
    newline and several regions; irregular regions (multi-line record, blank line, length mismatch, illegal letter) are
    reported and not scanned, and scanning them through the raw-read path afterwards completes the tallies."""
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    local = exp["mode"] == "local"
    idx = engine.Index.from_scaffolds(scaff, gs, len(counts))
    sc = engine.Scan(idx, k=run["k"], phred_cutoff=run.get("phred_cutoff", 30), mode=capi.MODE_LOCAL if local else capi.MODE_GLOBAL,
                     algorithm=algorithm, record=True)

    def check(r):
        assert (r.total_kmers, r.ambiguous, r.lookups, r.reads) == (exp["total_kmers"], exp["ambiguous"], exp["lookups"], len(seqs))
        if local:
            assert helpers.rel_close(r.unique_w, exp["unique_w"], TOL)
        else:
            assert [int(x) for x in r.unique] == exp["unique"]

    for nl, trailing, pieces in (("\n", True, 1), ("\r\n", True, 1), ("\n", False, 1), ("\n", True, 5), ("\r\n", False, 3)):
        sc.reset()
        n = len(seqs)
        cuts = [n * i // pieces for i in range(pieces + 1)]
        for a, b in zip(cuts[:-1], cuts[1:]):
            reg, _ = sc.submit_fastq(_fastq_text(seqs[a:b], quals[a:b], nl, trailing))
        assert reg == pieces - 1
        assert sc.fastq_rejected() == []
        check(sc.finish())
    # the refinement runs on what the device-parsed regions recorded
    pct, cnt = np.full(len(counts), 100.0 / len(counts)), np.asarray(counts, np.int32)
    tk = sc.refine(pct, cnt)
    sc.reset()
    seq, q, offs = helpers.flat_reads(seqs, quals)
    sc.submit_raw(seq, q, offs, len(seqs))
    sc.finish()
    assert helpers.rel_close(sc.refine(pct, cnt), tk, 1e-11)
    # irregular regions: region 1 of 3 is damaged in four different ways; it is reported, the other two are scanned
    third = len(seqs) // 3
    parts = [(0, third), (third, 2 * third), (2 * third, len(seqs))]
    good = [_fastq_text(seqs[a:b], quals[a:b]) for a, b in parts]
    a, b = parts[1]
    s_mid, q_mid = seqs[a:b], quals[a:b]
    j = next(i for i, s in enumerate(s_mid) if len(s) >= 8)
    multi = _fastq_text(s_mid[:j], q_mid[:j]) + f"@m\n{s_mid[j][:4]}\n{s_mid[j][4:]}\n+\n{q_mid[j][:4]}\n{q_mid[j][4:]}\n".encode("latin-1") + \
        _fastq_text(s_mid[j + 1:], q_mid[j + 1:])
    blank = _fastq_text(s_mid[:j], q_mid[:j]) + b"\n" + _fastq_text(s_mid[j:], q_mid[j:])
    mism = _fastq_text(s_mid[:j] + [s_mid[j] + "A"] + s_mid[j + 1:], q_mid)
    illegal = _fastq_text(s_mid[:j] + ["X" + s_mid[j][1:]] + s_mid[j + 1:], q_mid)
    for bad in (multi, blank, mism, illegal):
        sc.reset()
        for piece in (good[0], bad, good[2]):
            sc.submit_fastq(piece)
        assert sc.fastq_rejected() == [1]
        # the caller's part: the rejected region again through the host grammar (here: the clean records) and the raw-read path
        seq, q, offs = helpers.flat_reads(s_mid, q_mid)
        sc.submit_raw(seq, q, offs, len(s_mid))
        check(sc.finish())
    # an empty region and a region of empty reads
    sc.reset()
    sc.submit_fastq(b"")
    sc.submit_fastq(b"@a\n\n+\n\n@b\n\n+\n\n")
    assert sc.fastq_rejected() == []
    r = sc.finish()
    assert r.reads == 2 and r.lookups == 0
    sc.close()
    idx.close()


@pytest.mark.parametrize("name", ["global_single", "local_paired", "refs_with_n"])
@pytest.mark.parametrize("variant", ["wide", "slot_counters"])
def test_record_mode_variants(lib, name, variant, monkeypatch):
    """The record-mode layouts a small index never picks by itself, forced through the library's measurement switches:
    `wide` = 64-byte buckets with the per-slot counters inside (tables beyond L2), `slot_counters` = the per-slot side array
    instead of runs over text positions.  Same refinement as the fixture (oracle/_ref) for the k-mer table and the pair table."""
    monkeypatch.setenv("SPEQB_NO_RUNS", "1")
    if variant == "wide":
        monkeypatch.setenv("SPEQB_FORCE_WIDE", "1")
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    for algorithm in (capi.ALGO_KTABLE, capi.ALGO_KPAIR):
        out = driver.run_all(scaff, gs, counts, _reads(seqs, quals), k=run["k"], phred_cutoff=run.get("phred_cutoff", 30),
                             fixed_accuracy=run["fixed_accuracy"], paired=run["paired"], batch_reads=64, algorithm=algorithm)
        _compare(out, exp, len(counts))


@pytest.mark.parametrize("algorithm", [capi.ALGO_ROLLING, capi.ALGO_KTABLE, capi.ALGO_KPAIR])
def test_fusion_sets_many_variants(lib, algorithm):
    """More than 64 variants: the key of an ambiguous unit is (smallest, largest, further variants in between) instead of a bit
    mask.  Checked against the same scan run on every unit alone, lane-per-read and warp-per-read forms."""
    rng = np.random.default_rng(9)
    V, G, k = 70, 2500, 15
    scaff = [np.where((s := rng.integers(0, 4, G)) == 3, 4, s).astype(np.uint8) for _ in range(V)]   # unrelated variants
    idx = engine.Index.from_scaffolds(scaff, list(range(V)), V)
    letters = np.frombuffer(b"ACGNT", np.uint8)
    for L, R in ((150, 180), (2400, 24)):
        parts = []
        for i in range(R):
            a, b, c = rng.choice(V, 3, replace=False)
            p = int(rng.integers(0, G - L + 1))
            r = scaff[int(a)][p:p + L].copy()
            if i % 3 >= 1:
                r[L // 3:2 * L // 3] = scaff[int(b)][p + L // 3:p + 2 * L // 3]
            if i % 3 == 2:
                r[2 * L // 3:] = scaff[int(c)][p + 2 * L // 3:p + L]
            parts.append(letters[r])
        seq = np.concatenate(parts)
        qual = np.full(seq.size, ord("I"), np.uint8)
        offs = np.arange(R + 1, dtype=np.uint64) * L
        sc = engine.Scan(idx, k=k, algorithm=algorithm)
        sc.submit_raw(seq, qual, offs, R)
        r = sc.finish()
        sets = sc.fusion_sets()
        assert sum(sets.values()) == r.ambiguous > 0
        exp = {}
        for u in range(R):
            sc.reset()
            sc.submit_raw(seq, qual, offs[u:u + 2], 1)
            nz = np.nonzero(sc.finish().unique)[0]
            if nz.size > 1:
                t = (int(nz[0]), int(nz[-1])) + (("+",) if nz.size > 2 else ())
                exp[t] = exp.get(t, 0) + 1
        assert sets == exp and any(len(t) == 3 for t in sets)
        sc.close()
    idx.close()


@pytest.mark.parametrize("algorithm", [capi.ALGO_ROLLING, capi.ALGO_KTABLE, capi.ALGO_KPAIR])
@pytest.mark.parametrize("name", ["global_paired", "local_paired"])
def test_submit_fastq_pair_device_parser(lib, name, algorithm):
    """speqg_scan_submit_fastq_pair: a region of each mate file, same number of records, parsed and interleaved on the device.
    Same tallies as the fixture (oracle/_ref) in one and in several region pairs, LF and CRLF; a pair of regions whose record
    counts differ, or with an irregular record in either file, is reported and not scanned."""
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)            # interleaved mates
    s1, q1, s2, q2 = seqs[0::2], quals[0::2], seqs[1::2], quals[1::2]
    local = exp["mode"] == "local"
    idx = engine.Index.from_scaffolds(scaff, gs, len(counts))
    sc = engine.Scan(idx, k=run["k"], phred_cutoff=run.get("phred_cutoff", 30), mode=capi.MODE_LOCAL if local else capi.MODE_GLOBAL,
                     algorithm=algorithm, paired=True, record=True)

    def check(r):
        assert (r.total_kmers, r.ambiguous, r.lookups, r.reads) == (exp["total_kmers"], exp["ambiguous"], exp["lookups"], len(seqs))
        if local:
            assert helpers.rel_close(r.unique_w, exp["unique_w"], TOL)
        else:
            assert [int(x) for x in r.unique] == exp["unique"]

    n = len(s1)
    for nl, trailing, pieces in (("\n", True, 1), ("\r\n", False, 1), ("\n", True, 4)):
        sc.reset()
        cuts = [n * i // pieces for i in range(pieces + 1)]
        for a, b in zip(cuts[:-1], cuts[1:]):
            sc.submit_fastq_pair(_fastq_text(s1[a:b], q1[a:b], nl, trailing), _fastq_text(s2[a:b], q2[a:b], nl, trailing))
        assert sc.fastq_rejected() == []
        check(sc.finish())
    assert {",".join(map(str, key)): c for key, c in sc.fusion_sets().items()} == exp["fusion"]
    # record counts differ / an irregular record in the second file: the pair of regions is handed back
    h = n // 2
    for bad2 in (_fastq_text(s2[h:n - 1], q2[h:n - 1]), _fastq_text(s2[h:], q2[h:]).replace(b"\n+\n", b"\n+\n\n", 1)):
        sc.reset()
        sc.submit_fastq_pair(_fastq_text(s1[:h], q1[:h]), _fastq_text(s2[:h], q2[:h]))
        sc.submit_fastq_pair(_fastq_text(s1[h:], q1[h:]), bad2)
        assert sc.fastq_rejected() == [1]
        seq, q, offs = helpers.flat_reads(seqs[2 * h:], quals[2 * h:])
        sc.submit_raw(seq, q, offs, len(seqs) - 2 * h)
        check(sc.finish())
    with pytest.raises(capi.SpeqError):
        sc.submit_fastq(_fastq_text(s1, q1))                  # a paired scan takes pairs of regions
    sc.close()
    idx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("k", [5, 8, 16, 31])
@pytest.mark.parametrize("paired", [False, True])
def test_walking_scan_matches_rolling(lib, k, paired, monkeypatch):
    """Global mode with the k-mer table runs the walking scan (scanw_kernel): after a hit the read is compared with the packed
    texts and the window answers come from the per-position flags.  Differential test against the rolling FM search and the
    probing table scan (SPEQB_NO_WALK=1) on texts whose lengths sit around multiples of 32 (the walk reads 32-base groups), with
    an N inside a text, and reads built to stress the walk: whole texts (the walk reaches the last base of the last text),
    reads that run over the end of a text into random bases or into another text, variant switches, errors, Ns."""
    rng = np.random.default_rng(1000 + k)
    comp = {"A": "T", "C": "G", "G": "C", "T": "A", "N": "N"}
    anc = "".join("ACGT"[i] for i in rng.integers(0, 4, 260))
    texts, groups = [], []
    for v in range(3):
        for L in (64, 95 + v, 160, 256):
            t = list(anc[:L])
            for i in range(L):
                if rng.random() < 0.03:
                    t[i] = "ACGT"[rng.integers(0, 4)]
            if L == 160 and v == 1:
                t[77] = "N"
            texts.append("".join(t))
            groups.append(v)
    idx = engine.Index.from_scaffolds([engine.ascii_to_ranks(t) for t in texts], groups, 3)
    reads = [texts[-1], texts[0], texts[3] + texts[4][:40]]
    for _ in range(3000):
        kind = rng.integers(0, 4)
        t = texts[rng.integers(0, len(texts))]
        a = int(rng.integers(0, len(t)))
        piece = t[a:a + int(rng.integers(1, 151))]
        if kind == 1:
            piece += "".join("ACGT"[i] for i in rng.integers(0, 4, int(rng.integers(0, 40))))       # over the end of the text
        elif kind == 2:
            t2 = texts[rng.integers(0, len(texts))]
            b = int(rng.integers(0, len(t2)))
            piece += t2[b:b + int(rng.integers(1, 100))]                                           # into another text
        if rng.random() < 0.5:
            piece = "".join(comp[c] for c in reversed(piece))
        r = list(piece)
        for i in range(len(r)):
            u = rng.random()
            if u < 0.01:
                r[i] = "ACGT"[rng.integers(0, 4)]
            elif u < 0.013:
                r[i] = "N"
        reads.append("".join(r))
    if len(reads) % 2:
        reads.append("ACGT")
    seq, q, offs = helpers.flat_reads(reads, ["I" * len(r) for r in reads])

    def run(algorithm):
        sc = engine.Scan(idx, k=k, record=True, paired=paired, algorithm=algorithm)
        sc.submit_raw(seq, q, offs, len(reads))
        r = sc.finish()
        tk = sc.refine(np.full(3, 100.0 / 3), np.full(3, 4, np.int32))
        out = (r, tk, sc.multiplicity_count(), sc.fusion_sets())
        sc.close()
        return out
    ref = run(capi.ALGO_ROLLING)
    monkeypatch.setenv("SPEQB_NO_WALK", "1")
    probing = run(capi.ALGO_KTABLE)
    monkeypatch.delenv("SPEQB_NO_WALK")
    walking = run(capi.ALGO_KTABLE)
    assert walking[0].walked > 0 and walking[0].probes < probing[0].probes
    # the form tables beyond L2 take: positions in a side array, the presence filter in front of the probes
    monkeypatch.setenv("SPEQB_FORCE_SIDE", "1")
    monkeypatch.setenv("SPEQB_FORCE_FILTER", "1")
    walking_big = run(capi.ALGO_KTABLE)
    assert walking_big[0].walked > 0
    assert ref[0].total_kmers > 0 and ref[0].unique.sum() > 0
    for got in (probing, walking, walking_big):
        assert np.array_equal(got[0].unique, ref[0].unique)
        assert got[0].total_kmers == ref[0].total_kmers and got[0].ambiguous == ref[0].ambiguous
        assert got[0].lookups == ref[0].lookups and got[0].reads == ref[0].reads
        assert helpers.rel_close(got[1], ref[1], 1e-11)
        assert got[2] == ref[2]
        assert got[3] == ref[3]
    idx.close()
