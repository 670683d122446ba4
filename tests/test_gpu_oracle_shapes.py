"""GPU tier (B200): the CUDA path against the CPU oracle ON THE SHAPES BASELINE.json NAMES (C1 in full, C2-C5 with their
full-size indexes and a read count the oracle finishes in seconds), every product algorithm, global and local mode, plus
back-check and refinement.  The oracle is oracle/_ref/speq_ref (the reference's glue over the real vendored sdsl-lite
csa_wt; prebuilt, it travels to the GPU box) when present, else the self-contained restatement.

Bit-exact: unique[V], total_kmers, ambiguous, lookups, unique_ref[V], total_ref[V], the fix-forward fusion map.
<= 1e-12 relative: local-mode unique_w, percent_initial, percent_final; iteration count equal (north_star's tolerance)."""
from __future__ import annotations

import numpy as np
import pytest

import helpers
from speq_b200 import capi, driver, engine, synth
from test_gpu_parity import _compare

pytestmark = pytest.mark.gpu
TOL = 1e-12

# cfg: reads of the scan jobs, reads of the refinement job (the oracle re-searches every read per iteration), whether the
# oracle runs the back-check itself.  C3 / C4: the oracle's back-check locates ~1e3 occurrences per k-mer (C3) or walks
# 2e8 windows (C4) -- minutes to hours -- so there the device back-check's counts are injected into the oracle's solve
# (--unique-ref / --total-ref) and the back-check itself is pinned on reduced shapes (test_backcheck_reduced_shapes).
SHAPES = {
    "C1": dict(R=100_000, R_local=25_000, R_refine=2_000, oracle_backcheck=True),     # C1 in full
    "C2": dict(R=8_000, R_refine=1_000, oracle_backcheck=True),
    "C3": dict(R=3_000, R_refine=300, oracle_backcheck=False),
    "C4": dict(R=30_000, R_refine=2_000, oracle_backcheck=False),
    "C5": dict(R=160, R_refine=16, oracle_backcheck=True),
}


def _oracle(oracle_bin, ref_bin):
    return ref_bin if ref_bin is not None else oracle_bin


def _fusion_dict(sc) -> dict:
    return {",".join(str(g) for g in key): n for key, n in sc.fusion_sets().items()}


@pytest.mark.parametrize("cfg", list(SHAPES))
def test_vs_oracle_config_shapes(lib, oracle_bin, ref_bin, tmp_path, cfg):
    import torch
    spec = SHAPES[cfg]
    sh = synth.CONFIGS[cfg]
    R, Rr, Rl = spec["R"], spec["R_refine"], spec.get("R_local", spec["R"])
    refs = synth.make_refs(sh, synth.SEED0 + int(cfg[1]))                       # CPU generator: same bytes here and on the box
    seq, qual = synth.make_reads(sh, refs, R, synth.SEED0 + 200 + int(cfg[1]), n_rate=0.0005)
    _, qual_l = synth.make_reads(sh, refs, R, synth.SEED0 + 200 + int(cfg[1]), n_rate=0.0005, local_quals=True)
    seq, qual, qual_l = seq.numpy(), qual.numpy(), qual_l.numpy()
    synth.write_fasta(str(tmp_path / "refs.fa"), refs)
    synth.write_groups(str(tmp_path / "groups.txt"), sh)
    helpers.write_fastq_fixed(tmp_path / "g.fq", seq, qual)
    helpers.write_fastq_fixed(tmp_path / "l.fq", seq, qual_l)
    offs = np.arange(R + 1, dtype=np.uint64) * sh.L
    seq_f, qual_f, qual_lf = seq.reshape(-1), qual.reshape(-1), qual_l.reshape(-1)

    idx = engine.Index.from_scaffolds(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), sh.V, device_sa=True)
    uref, tref = engine.backcheck(idx, sh.k)
    inject = {} if spec["oracle_backcheck"] else dict(unique_ref=[int(x) for x in uref], total_ref=[int(x) for x in tref])
    jobs = [dict(reads1=tmp_path / "g.fq", k=sh.k, fixed_accuracy=sh.accuracy, phred_cutoff=30, refine=False, **inject),
            dict(reads1=tmp_path / "l.fq", fixed_accuracy=0.0, phred_cutoff=19, max_reads=Rl),
            dict(reads1=tmp_path / "g.fq", fixed_accuracy=sh.accuracy, phred_cutoff=30, max_reads=Rr, refine=True),
            dict(reads1=tmp_path / "l.fq", fixed_accuracy=0.0, phred_cutoff=19, max_reads=max(1, Rr // 2), refine=True)]
    import os
    exp_g, exp_l, exp_r, exp_rl = helpers.run_oracle_jobs(_oracle(oracle_bin, ref_bin), tmp_path / "refs.fa", tmp_path / "groups.txt", jobs,
                                                          threads=min(32, os.cpu_count() or 4), backcheck=spec["oracle_backcheck"])
    assert exp_g["n"] == idx.n
    if spec["oracle_backcheck"]:
        assert [int(x) for x in uref] == exp_g["unique_ref"] and [int(x) for x in tref] == exp_g["total_ref"]

    algos = [capi.ALGO_AUTO, capi.ALGO_ROLLING, capi.ALGO_KTABLE] + ([capi.ALGO_KPAIR] if sh.V <= 254 else [])
    # ---- global mode: integer tallies bit-exact, every product algorithm
    for alg in algos:
        sc = engine.Scan(idx, k=sh.k, phred_cutoff=30, algorithm=alg)
        sc.submit_raw(seq_f, qual_f, offs, R)
        r = sc.finish()
        assert [int(x) for x in r.unique] == exp_g["unique"], alg
        assert (r.total_kmers, r.ambiguous, r.lookups, r.reads) == (exp_g["total_kmers"], exp_g["ambiguous"], exp_g["lookups"], R), alg
        assert _fusion_dict(sc) == exp_g["fusion"], alg
        sc.close()
    # ---- local mode: phred-weighted tallies within 1e-12
    for alg in algos:
        sc = engine.Scan(idx, k=sh.k, phred_cutoff=19, mode=capi.MODE_LOCAL, algorithm=alg)
        sc.submit_raw(seq_f, qual_lf, offs[:Rl + 1], Rl)
        r = sc.finish()
        assert helpers.rel_close(r.unique_w, exp_l["unique_w"], TOL), alg
        assert (r.total_kmers, r.ambiguous, r.lookups) == (exp_l["total_kmers"], exp_l["ambiguous"], exp_l["lookups"]), alg
        assert _fusion_dict(sc) == exp_l["fusion"], alg
        sc.close()
    # ---- solve + refinement on the first R_refine reads: percent_initial / percent_final / iterations
    for exp, q, n, acc, cutoff in ((exp_r, qual_f, Rr, sh.accuracy, 30), (exp_rl, qual_lf, max(1, Rr // 2), 0.0, 19)):
        for alg in (capi.ALGO_AUTO, capi.ALGO_ROLLING):
            out = driver.run_all(None, None, [sh.I] * sh.V, driver.Reads(seq_f, q, offs[:n + 1]), k=sh.k, phred_cutoff=cutoff,
                                 fixed_accuracy=acc, batch_reads=4096, algorithm=alg, index=idx)
            _compare(out, exp, sh.V)
    idx.close()
    torch.cuda.empty_cache()


@pytest.mark.parametrize("cfg,G,I", [("C3", 1500, 2), ("C4", 200_000, 1)])
def test_backcheck_reduced_shapes(lib, oracle_bin, ref_bin, tmp_path, cfg, G, I):
    """The back-check of the two shapes whose full-size oracle back-check takes minutes to hours, on reduced genomes:
    C3 as 50 variants x 2 isolates x 1.5 kb (k = 25), C4 as 20 strains x 200 kb (k = 31).  unique_ref / total_ref bit-exact."""
    import dataclasses
    sh = dataclasses.replace(synth.CONFIGS[cfg], G=G, I=I)
    refs = synth.make_refs(sh, synth.SEED0 + 30 + int(cfg[1]))
    seq, qual = synth.make_reads(sh, refs, 64, 5)
    synth.write_fasta(str(tmp_path / "refs.fa"), refs)
    synth.write_groups(str(tmp_path / "groups.txt"), sh)
    helpers.write_fastq_fixed(tmp_path / "g.fq", seq.numpy(), qual.numpy())
    import os
    exp = helpers.run_oracle_jobs(_oracle(oracle_bin, ref_bin), tmp_path / "refs.fa", tmp_path / "groups.txt",
                                  [dict(reads1=tmp_path / "g.fq", k=sh.k, fixed_accuracy=sh.accuracy, refine=False)],
                                  threads=min(32, os.cpu_count() or 4))[0]
    idx = engine.Index.from_scaffolds(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), sh.V)
    uref, tref = engine.backcheck(idx, sh.k)
    assert [int(x) for x in uref] == exp["unique_ref"] and [int(x) for x in tref] == exp["total_ref"]
    assert sum(exp["unique_ref"]) > 0
    idx.close()


@pytest.mark.parametrize("k", [32, 33, 64, 70, 127])
@pytest.mark.parametrize("mode", ["global", "local", "global_paired", "local_paired"])
def test_vs_oracle_long_k(lib, oracle_bin, tmp_path, k, mode):
    """k > 31 (scan32_kernel; k = 70 is the reference's default, include/arg_parse.h:20): G = 3000, reads of 230 bases with N,
    IUPAC letters, low-quality bases and ragged lengths, through AUTO and ROLLING, single and paired, global and local."""
    local, paired = mode.startswith("local"), mode.endswith("paired")
    case = helpers.random_case(1000 + k, V=4, I=2, G=3000, R=260, L=230, e=0.003, d_var=0.03, n_rate_read=0.002, lowq_rate=0.003,
                               ragged=True, iupac=True, local_quals="high" if local else False)
    refs, groups, r1, r2 = helpers.case_files(tmp_path, case, paired)
    acc = 0.0 if local else 0.997
    exp = helpers.run_oracle(oracle_bin, refs, groups, r1, reads2=r2, k=k, fixed_accuracy=acc, threads=8)
    scaff = [engine.ascii_to_ranks(s) for s in case["scaffolds"]]
    seq, q, offs = helpers.flat_reads(case["seqs"], case["quals"])
    assert exp["total_kmers"] > 0 and (sum(exp["unique"]) > 0 or sum(exp["unique_w"]) > 0)
    for alg in (capi.ALGO_AUTO, capi.ALGO_ROLLING):
        out = driver.run_all(scaff, case["group_scaffolds"], [case["I"]] * case["V"], driver.Reads(seq, q, offs), k=k,
                             fixed_accuracy=acc, paired=paired, batch_reads=96, algorithm=alg)
        _compare(out, exp, case["V"])
