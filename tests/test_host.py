"""CPU tier: the C ABI loads and exports every declared symbol; host packer; host index builder vs the oracle's
SA/BWT/C; the kernels' per-lane arithmetic (scan_core.h, compiled for the host by tests/sim) vs the golden fixtures;
compute entry points fail loudly without a CUDA device."""
from __future__ import annotations

import ctypes as C
import re
import subprocess

import numpy as np
import pytest

import helpers
from speq_b200 import capi, engine


def _have_gpu():
    import torch
    return torch.cuda.is_available()


def test_header_symbols_exported(lib):
    hdr = (helpers.ROOT / "include" / "speq_cuda.h").read_text()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(speqg_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    assert declared == set(capi.SIGNATURES), declared ^ set(capi.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.speqg_abi_version() == 2
    # the product library must not contain the oracle
    nm = subprocess.run(["nm", "-D", "--defined-only", str(capi.LIB_PATH)], capture_output=True, text=True).stdout
    assert "oracle" not in nm.lower()


def test_no_cpu_fallback(lib):
    if _have_gpu():
        pytest.skip("a CUDA device is present")
    idx = engine.Index.from_scaffolds([engine.ascii_to_ranks("ACGTACGTTGCA")], [0], 1)
    with pytest.raises(capi.SpeqError) as ei:
        engine.Scan(idx, k=3)
    assert ei.value.code == capi.ECUDA
    with pytest.raises(capi.SpeqError) as ei:
        engine.backcheck(idx, 3)
    assert ei.value.code == capi.ECUDA


def test_index_build_errors(lib):
    with pytest.raises(capi.SpeqError) as ei:
        engine.Index.from_texts([], [], 1)
    assert "cannot be empty" in ei.value.message          # seqan3 fm_index.hpp:240
    with pytest.raises(capi.SpeqError) as ei:
        engine.Index.from_texts([np.zeros(0, np.uint8)] * 2, [0, 0], 1)
    assert "only contains empty texts" in ei.value.message  # fm_index.hpp:257
    with pytest.raises(capi.SpeqError):
        engine.Index.from_texts([np.array([0, 1, 2], np.uint8)], [-1], 1)
    with pytest.raises(capi.SpeqError):
        engine.Index.from_texts([np.array([0, 1, 7], np.uint8)], [0], 1)


def test_pack_reads_matches_numpy(lib):
    rng = np.random.default_rng(3)
    letters = np.frombuffer(b"ACGTUNRYSWKMBDHVacgtunryswkmbdhv", np.uint8)
    lens = [0, 1, 31, 32, 33, 64, 150, 7]
    seqs = [bytes(rng.choice(letters, n)).decode() for n in lens]
    quals = ["".join(chr(int(x)) for x in rng.integers(33, 90, n)) for n in lens]
    seq, q, offs = helpers.flat_reads(seqs, quals)
    cutoff = 30
    pb = engine.pack_reads(seq, q, offs, phred_cutoff=cutoff, with_qual=True, threads=2)
    assert pb.n_reads == len(lens) and pb.n_groups32 == sum((n + 31) // 32 for n in lens)
    code = {"A": 0, "C": 1, "G": 2, "T": 3, "U": 3}
    for i, (s, ql) in enumerate(zip(seqs, quals)):
        g0 = int(pb.grp_off[i])
        assert int(pb.len[i]) == len(s)
        for j, ch in enumerate(s):
            c = (int(pb.bases[g0 + j // 32]) >> (2 * (j % 32))) & 3
            iv = (int(pb.inv[g0 + j // 32]) >> (j % 32)) & 1
            ph = min(max(ord(ql[j]) - 33, 0), 41)
            isn = ch.upper() not in code
            assert c == (0 if isn else code[ch.upper()])
            assert iv == int(isn or not (ph > cutoff))
            assert int(pb.qual[(g0 + j // 32) * 32 + j % 32]) == ph
    with pytest.raises(capi.SpeqError) as ei:
        s2, q2, o2 = helpers.flat_reads(["ACGTXACGT"], ["IIIIIIIII"])
        engine.pack_reads(s2, q2, o2)
    assert "parse_error" in ei.value.message


def test_index_matches_oracle_dump(lib, oracle_bin, tmp_path):
    case = helpers.random_case(21, V=3, I=2, G=300, R=1, n_rate_ref=0.01)
    helpers.write_fasta(tmp_path / "refs.fa", case["scaffolds"])
    subprocess.run([str(oracle_bin), "dump", str(tmp_path / "refs.fa"), str(tmp_path / "dump.bin")], check=True)
    raw = (tmp_path / "dump.bin").read_bytes()
    n = int(np.frombuffer(raw[:8], np.uint64)[0])
    Cc = np.frombuffer(raw[8:72], np.uint64)
    rec = np.frombuffer(raw[72:], np.dtype([("sa", "<u4"), ("bwt", "u1")]))
    idx = engine.Index.from_scaffolds([engine.ascii_to_ranks(s) for s in case["scaffolds"]], case["group_scaffolds"], case["V"], keep_sa=True)
    assert idx.n == n and idx.C == [int(x) for x in Cc]
    ex = idx.export(sa=True)
    assert np.array_equal(ex["sa"], rec["sa"])
    bwt = rec["bwt"].copy()
    bwt[(bwt == 0) | (bwt == 6)] = 0xFF
    assert np.array_equal(ex["bwt"], bwt)
    # lcp against a direct comparison of suffixes of the reversed text
    T = np.zeros(n, np.uint8)
    f = 0
    texts = []
    for s in case["scaffolds"]:
        r = engine.ascii_to_ranks(s)
        texts += [r, engine.revcomp_ranks(r)]
    m = n - 1
    for i, t in enumerate(texts):
        T[m - 1 - f - np.arange(t.size)] = t + 1
        f += t.size
        if i + 1 < len(texts):
            T[m - 1 - f] = 6
            f += 1
    sa = ex["sa"].astype(np.int64)
    for i in range(1, n, 17):
        a, b, h = sa[i - 1], sa[i], 0
        while h < 127 and T[a + h] == T[b + h]:
            h += 1
        assert ex["lcp"][i] == h
    # save / load round trip
    idx.save(str(tmp_path / "x.idx"))
    idx2 = engine.Index.load(str(tmp_path / "x.idx"))
    ex2 = idx2.export(sa=True)
    for key in ex:
        assert np.array_equal(ex[key], ex2[key])
    with pytest.raises(capi.SpeqError) as ei:
        (tmp_path / "bad.idx").write_bytes(b"not an index")
        engine.Index.load(str(tmp_path / "bad.idx"))
    assert ei.value.code == capi.EIO


def _sim_scan(sim, scaff, gs, V, seqs, quals, k, cutoff, paired, algo, backcheck=False, want_mult=False, local=False, nlanes=7,
              seg_len=37, phases=None):
    texts, grp = [], []
    for j, s in enumerate(scaff):
        texts += [np.ascontiguousarray(s), np.ascontiguousarray(engine.revcomp_ranks(s))]
        grp += [int(gs[j])] * 2
    S = len(texts)
    ptrs = (C.c_void_p * S)(*[t.ctypes.data if t.size else None for t in texts])
    lens = np.array([t.size for t in texts], np.uint64)
    grp = np.array(grp, np.int32)
    seq, q, offs = helpers.flat_reads(seqs, quals)
    pb = engine.pack_reads(seq, q, offs, phred_cutoff=cutoff, with_qual=True)
    out = np.zeros(2 * V + 4, np.uint64)
    out_w = np.zeros(V, np.float64)
    n = int(lens.sum()) + S
    mult = np.zeros(n + 1, np.uint64) if want_mult else None
    sim.sim_scan.restype = C.c_int
    rc = sim.sim_scan(ptrs, C.c_void_p(lens.ctypes.data), C.c_uint32(S), C.c_void_p(grp.ctypes.data), C.c_uint32(V),
                      C.c_void_p(pb.bases.ctypes.data), C.c_void_p(pb.inv.ctypes.data), C.c_void_p(pb.qual.ctypes.data),
                      C.c_void_p(pb.grp_off.ctypes.data), C.c_void_p(pb.len.ctypes.data), C.c_uint64(pb.n_reads), C.c_uint32(k),
                      C.c_int(int(paired)), C.c_int(algo), C.c_int(int(backcheck)), C.c_int(int(local)), C.c_uint32(nlanes),
                      C.c_uint32(seg_len), C.c_void_p(out.ctypes.data), C.c_void_p(out_w.ctypes.data),
                      C.c_void_p(mult.ctypes.data if want_mult else None), C.c_void_p(phases.ctypes.data if phases is not None else None))
    assert rc == 0
    return out, out_w, mult


@pytest.mark.parametrize("name", helpers.GOLDEN_NAMES + helpers.GOLDEN_LONG_K)
@pytest.mark.parametrize("algo", [0, 1, 2])
def test_lane_arithmetic_vs_golden(lib, sim, name, algo):
    fx = helpers.load_golden(name)
    exp, run = fx["expected"], fx["run"]
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    V = len(counts)
    local = exp["mode"] == "local"
    out, out_w, _ = _sim_scan(sim, scaff, gs, V, seqs, quals, run["k"], run.get("phred_cutoff", 30), run["paired"], algo,
                              local=(local and algo != 1))
    if not local:
        assert list(out[:V]) == exp["unique"]
    elif algo != 1:
        assert helpers.rel_close(out_w, exp["unique_w"], 1e-12)
    assert int(out[V]) == exp["total_kmers"]
    assert int(out[V + 1]) == exp["ambiguous"]
    assert int(out[V + 2]) == exp["lookups"]
    if algo == 0:
        for nlanes, seg_len in ((1, 5), (7, 37), (64, 100000)):
            bc, _, _ = _sim_scan(sim, scaff, gs, V, seqs, quals, run["k"], 30, False, 0, backcheck=True, nlanes=nlanes, seg_len=seg_len)
            assert list(bc[:V]) == exp["unique_ref"]
            assert list(bc[V:2 * V]) == exp["total_ref"]
    if algo != 1:
        one, _, _ = _sim_scan(sim, scaff, gs, V, seqs, quals, run["k"], run.get("phred_cutoff", 30), run["paired"], algo, nlanes=1, local=local)
        assert np.array_equal(one, out)
        many, _, _ = _sim_scan(sim, scaff, gs, V, seqs, quals, run["k"], run.get("phred_cutoff", 30), run["paired"], algo, nlanes=64, local=local)
        assert np.array_equal(many, out)


@pytest.mark.parametrize("k", [1, 2, 3, 31, 64, 127])
def test_lane_arithmetic_k_extremes(lib, sim, k):
    """rolling == direct for k from 1 to the maximum (long, repetitive texts so that deep intervals stay wide)."""
    rng = np.random.default_rng(k)
    unit = "".join(rng.choice(list("ACGT"), 90))
    scaff_s = [unit * 4 + "ACGTN" + unit[:50], unit[:70] + "T" + unit[70:] + unit * 2, unit[::-1] * 3]
    scaff = [engine.ascii_to_ranks(s) for s in scaff_s]
    seqs = [scaff_s[i % 3][j:j + 200] for i, j in enumerate(rng.integers(0, 100, 40))]
    seqs = ["".join(ch if rng.random() > 0.01 else "ACGT"[int(rng.integers(0, 4))] for ch in s) for s in seqs]
    quals = ["I" * len(s) for s in seqs]
    a, _, ma = _sim_scan(sim, scaff, [0, 1, 2], 3, seqs, quals, k, 30, False, 0, want_mult=True)
    b, _, mb = _sim_scan(sim, scaff, [0, 1, 2], 3, seqs, quals, k, 30, False, 1, want_mult=True)
    c, _, mc = _sim_scan(sim, scaff, [0, 1, 2], 3, seqs, quals, k, 30, False, 2, want_mult=True)
    assert np.array_equal(a, b)
    assert np.array_equal(ma, mb)
    assert np.array_equal(c, b)
    assert np.array_equal(mc, mb)
    assert a[3] > 0


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_lane_arithmetic_low_complexity(lib, sim, seed):
    """Wide SA intervals (poly-A, short tandem repeats, near-identical copies): the e-bit contraction has to leave its two
    blocks and the general lcp scan has to cross many 32-byte chunks.  rolling == direct, tallies and multiplicities."""
    rng = np.random.default_rng(seed)
    parts = ["A" * 300, "AC" * 200, "ACG" * 150, "".join(rng.choice(list("ACGT"), 200))]
    base = "".join(parts[i] for i in rng.permutation(4))
    scaff_s = []
    for v in range(4):
        s = list(base)
        for p in rng.integers(0, len(s), 6):
            s[p] = "ACGTN"[int(rng.integers(0, 5))]
        scaff_s.append("".join(s))
    scaff = [engine.ascii_to_ranks(s) for s in scaff_s]
    seqs = []
    for _ in range(60):
        s = scaff_s[int(rng.integers(0, 4))]
        j = int(rng.integers(0, len(s) - 120))
        r = list(s[j:j + 120])
        for p in rng.integers(0, 120, 2):
            r[p] = "ACGT"[int(rng.integers(0, 4))]
        seqs.append("".join(r))
    quals = ["I" * len(s) for s in seqs]
    for k in (4, 9, 17, 40):
        a, _, ma = _sim_scan(sim, scaff, [0, 1, 2, 3], 4, seqs, quals, k, 30, False, 0, want_mult=True, nlanes=5)
        b, _, mb = _sim_scan(sim, scaff, [0, 1, 2, 3], 4, seqs, quals, k, 30, False, 1, want_mult=True)
        c, _, mc = _sim_scan(sim, scaff, [0, 1, 2, 3], 4, seqs, quals, k, 30, False, 2, want_mult=True, nlanes=5)
        assert np.array_equal(a, b), k
        assert np.array_equal(ma, mb), k
        assert np.array_equal(c, b), k
        assert np.array_equal(mc, mb), k


@pytest.mark.parametrize("fixture,golden", [("ref_global_single.idx", "global_single"), ("ref_refs_with_n.idx", "refs_with_n")])
def test_reference_idx_reader(lib, tmp_path, fixture, golden):
    """speqg_refidx_*: an .idx as the REFERENCE writes it (cereal archive of the metadata + the real sdsl csa_wt, generated by
    tests/golden/make_ref_idx.py through oracle/_ref) -> metadata and the indexed texts, decoded from the wavelet tree.  The
    texts must be scaffold j, reverse complement of scaffold j, ... of the FASTA the index was built from, and the index built
    over them must be the one speqg_index_build makes from the FASTA (same saved bytes)."""
    fx = helpers.load_golden(golden)
    r = engine.ReferenceIndexFile(str(helpers.GOLDEN / fixture))
    names, counts, gs = helpers.parse_groups(fx["groups_txt"])
    assert r.reference_path == "refs.fa"
    assert r.group_names == names and r.group_counts == counts and r.group_scaffolds == [int(g) for g in gs]
    assert r.n == fx["expected"]["n"]
    scaff = [engine.ascii_to_ranks(s) for s in helpers.parse_fasta(fx["refs_fa"])]
    texts = r.texts()
    assert len(texts) == 2 * len(scaff)
    for j, s in enumerate(scaff):
        assert np.array_equal(texts[2 * j], s)
        assert np.array_equal(texts[2 * j + 1], engine.revcomp_ranks(s))
    a = r.build_index()
    b = engine.Index.from_scaffolds(scaff, gs, len(counts))
    assert a.n == b.n == r.n and a.C == b.C
    a.save(str(tmp_path / "a.idx")); b.save(str(tmp_path / "b.idx"))
    assert (tmp_path / "a.idx").read_bytes() == (tmp_path / "b.idx").read_bytes()
    a.close(); b.close(); r.close()


def test_reference_idx_reader_rejects_other_files(lib, tmp_path):
    (tmp_path / "x.idx").write_bytes(b"SPQIDX02" + bytes(200))
    with pytest.raises(capi.SpeqError) as ei:
        engine.ReferenceIndexFile(str(tmp_path / "x.idx"))
    assert ei.value.code == capi.EIO
    (tmp_path / "short.idx").write_bytes(b"\x05\x00\x00\x00\x00\x00\x00\x00abc")
    with pytest.raises(capi.SpeqError):
        engine.ReferenceIndexFile(str(tmp_path / "short.idx"))
