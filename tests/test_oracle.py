"""CPU tier: pin the oracle.  (1) seqan3's own known-answer vectors, (2) the SPeQ-level golden fixtures produced by the
sdsl-backed build, (3) restatement vs sdsl build on fresh random inputs (when oracle/_ref exists), (4) brute force."""
from __future__ import annotations

import json

import numpy as np
import pytest

import helpers

KATS = json.loads((helpers.GOLDEN / "seqan3_kats.json").read_text())["cases"]


def _check_kats(binary):
    for c in KATS:
        out = helpers.run_oracle_search(binary, c["sigma"], c["texts"], [q["pieces"] for q in c["queries"]], "size" in c)
        if "size" in c:
            assert out[0]["size"] == c["size"], c["source"]
            out = out[1:]
        assert len(out) == len(c["queries"])
        for q, o in zip(c["queries"], out):
            assert o["ok"] == q["ok"], (c["source"], q)
            for key in ("count", "depth", "hits", "locate"):
                if key in q:
                    assert o[key] == q[key], (c["source"], key, q, o)


def test_seqan3_kats_restatement(oracle_bin):
    _check_kats(oracle_bin)


def test_seqan3_kats_sdsl(ref_bin):
    if ref_bin is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    _check_kats(ref_bin)


def _same(a: dict, b: dict):
    for key in ("n", "V", "reads", "unique_ref", "total_ref", "unique", "total_kmers", "ambiguous", "iterations", "lookups",
                "steps_executed", "kmers_hit", "occ", "occ_sectors"):
        assert a[key] == b[key], key
    # same arithmetic, same order (threads=1): bit-identical doubles
    for key in ("unique_w", "unique_totals", "percent_initial", "percent_final"):
        assert a[key] == b[key], key
    assert a["iter_tkpg"] == b["iter_tkpg"]
    assert a["fusion"] == b["fusion"]


@pytest.mark.parametrize("name", helpers.GOLDEN_NAMES + helpers.GOLDEN_LONG_K)
def test_golden_fixtures(oracle_bin, tmp_path, name):
    fx = helpers.load_golden(name)
    refs, groups, r1, r2 = helpers.golden_files(tmp_path, fx)
    run = dict(fx["run"])
    run.pop("paired")
    out = helpers.run_oracle(oracle_bin, refs, groups, r1, reads2=r2, threads=1, **run)
    _same(out, fx["expected"])


@pytest.mark.parametrize("seed", [101, 102])
def test_restatement_vs_sdsl_random(oracle_bin, ref_bin, tmp_path, seed):
    if ref_bin is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    case = helpers.random_case(seed, V=3, I=2, G=700, R=150, L=64, e=0.02, n_rate_ref=0.003, n_rate_read=0.003,
                               lowq_rate=0.01, ragged=True, local_quals=(seed % 2 == 0))
    refs, groups, r1, _ = helpers.case_files(tmp_path, case)
    kw = dict(k=12, fixed_accuracy=0.0 if seed % 2 == 0 else 0.98, phred_cutoff=19, threads=1)
    _same(helpers.run_oracle(oracle_bin, refs, groups, r1, **kw), helpers.run_oracle(ref_bin, refs, groups, r1, **kw))


def test_search_vs_bruteforce(oracle_bin):
    """locate() output == every (text, pos) where the query occurs, by plain string search."""
    rng = np.random.default_rng(5)
    texts = ["".join(rng.choice(list("ACGTN"), p=[.24, .24, .24, .04, .24], size=int(n))) for n in (57, 0, 133, 21, 1)]
    queries = []
    for _ in range(60):
        t = texts[int(rng.choice([0, 2, 3]))]
        L = int(rng.integers(1, 8))
        s = int(rng.integers(0, len(t) - L + 1))
        q = t[s:s + L]
        if rng.random() < 0.3:
            q = "".join(rng.choice(list("ACGT"), size=L))
        queries.append(q)
    out = helpers.run_oracle_search(oracle_bin, 5, texts, [[q] for q in queries])
    for q, o in zip(queries, out):
        want = [[ti, p] for ti, t in enumerate(texts) for p in range(len(t) - len(q) + 1) if t[p:p + len(q)] == q]
        if want:
            assert o["ok"] == [True] and o["hits"] == want, q
        else:
            assert o["ok"] == [False], q


def test_thread_count_does_not_change_integers(oracle_bin, tmp_path):
    case = helpers.random_case(7, R=120)
    refs, groups, r1, _ = helpers.case_files(tmp_path, case)
    a = helpers.run_oracle(oracle_bin, refs, groups, r1, k=11, fixed_accuracy=0.99, threads=1)
    b = helpers.run_oracle(oracle_bin, refs, groups, r1, k=11, fixed_accuracy=0.99, threads=3)
    for key in ("unique", "total_kmers", "ambiguous", "unique_ref", "total_ref", "iterations"):
        assert a[key] == b[key]
    assert helpers.rel_close(a["percent_final"], b["percent_final"])
