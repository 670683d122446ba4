"""The `speq index|scan|all` host binary (speq_b200/host): reference flags, .idx/.dat caching rules, stderr report.
CPU tier: argument handling, groups grammar via `speq index`, staleness rules, and that `scan` refuses to run without a
CUDA device.  GPU tier: `speq all` on the golden fixtures, compared with the sdsl-backed expectations."""
from __future__ import annotations

import json
import os
import subprocess
import time

import numpy as np
import pytest

import helpers


@pytest.fixture(scope="module")
def speq_bin(lib):
    from speq_b200 import build
    b = build.build_cli()
    assert b is not None and b.exists()
    return b


def run(speq_bin, *a, cwd=None):
    return subprocess.run([str(speq_bin), *map(str, a)], capture_output=True, text=True, cwd=cwd)


def test_argument_errors(speq_bin, tmp_path):
    r = run(speq_bin)
    assert r.returncode == 0 and r.stderr.startswith("SPeQ Error: ")
    r = run(speq_bin, "scan", "-t", "1")
    assert "speq scan | argument parsing error:  Validation failed for option -t/--threads" in r.stderr
    r = run(speq_bin, "all", "--fixed-accuracy", "1.5")
    assert "speq all | argument parsing error:  Validation failed for option --fixed-accuracy" in r.stderr
    r = run(speq_bin, "index", "-r", "missing.fa", cwd=tmp_path)
    assert 'The relative-path file "missing.fa" was not found.' in r.stderr
    r = run(speq_bin, "index", "-r", "/nonexistent/dir/x.fa")
    assert 'The full-path file "/nonexistent/dir/x.fa" was not found.' in r.stderr
    (tmp_path / "output.txt").write_text("x")
    r = run(speq_bin, "all", cwd=tmp_path)
    assert "Cowardly refusing to use an existing output file. Use '-f' to overwrite." in r.stderr
    r = run(speq_bin, "index", "--kmer", "5")
    assert "Unknown option --kmer" in r.stderr          # speq index takes no -k (src/arg_parse.cpp:91-126)


def test_index_cache_rules(speq_bin, tmp_path):
    fx = helpers.load_golden("global_single")
    refs, groups, r1, _ = helpers.golden_files(tmp_path, fx)
    idx = tmp_path / "cache.idx"
    r = run(speq_bin, "index", "-r", refs, "-g", groups, "-x", idx, cwd=tmp_path)
    assert r.returncode == 0 and idx.exists(), r.stderr
    m0 = idx.stat().st_mtime_ns
    time.sleep(0.05)
    run(speq_bin, "index", "-r", refs, "-g", groups, "-x", idx, cwd=tmp_path)
    assert idx.stat().st_mtime_ns == m0                  # inputs unchanged: not rebuilt (src/fm_indexer.cpp:68-97)
    run(speq_bin, "index", "-r", refs, "-g", groups, "-x", idx, "-f", cwd=tmp_path)
    m1 = idx.stat().st_mtime_ns
    assert m1 != m0                                      # -f forces a rebuild
    time.sleep(0.05)
    os.utime(groups)                                     # groups file touched: rebuilt
    run(speq_bin, "index", "-r", refs, "-g", groups, "-x", idx, cwd=tmp_path)
    assert idx.stat().st_mtime_ns != m1
    # the payload is loadable by the library at the recorded offset
    from speq_b200 import engine
    raw = idx.read_bytes()
    assert raw[:8] == b"SPQIDX02"


def test_groups_grammar_through_index(speq_bin, tmp_path):
    helpers.write_fasta(tmp_path / "r.fa", ["ACGTACGTAA", "ACGTTCGTAA", "TTGTACGTAA", "ACGAACGTCC"])
    (tmp_path / "g.txt").write_text("# comment line\nalpha (2) : 0 - 1 # trailing comment\nno colon here\nbeta(2):\t3 , x, 2\n")
    r = run(speq_bin, "index", "-r", "r.fa", "-g", "g.txt", "-x", "g.idx", cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert "A non-integer index was detected and ignored at: x" in r.stderr
    raw = (tmp_path / "g.idx").read_bytes()
    assert b"alpha " in raw and b"beta" in raw           # names are kept untrimmed, as the reference does
    # a groups line without (count) makes std::stoi throw, as in the reference (src/file_to_map.cpp:42-43)
    (tmp_path / "bad.txt").write_text("MCV-1p: 0-3\n")
    r = run(speq_bin, "index", "-r", "r.fa", "-g", "bad.txt", "-x", "bad.idx", cwd=tmp_path)
    assert r.returncode != 0 and "stoi" in r.stderr
    # an unassigned scaffold is rejected loudly (reference: undefined behaviour in the refinement loop)
    (tmp_path / "part.txt").write_text("a(1): 0\nb(1): 2-3\n")
    r = run(speq_bin, "index", "-r", "r.fa", "-g", "part.txt", "-x", "part.idx", cwd=tmp_path)
    assert r.returncode != 0 and "every scaffold must be assigned" in r.stderr


def test_scan_needs_gpu(speq_bin, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    fx = helpers.load_golden("global_single")
    refs, groups, r1, _ = helpers.golden_files(tmp_path, fx)
    r = run(speq_bin, "all", "-r", refs, "-g", groups, "-1", r1, "-x", tmp_path / "i.idx", "-k", 13, "--fixed-accuracy", 0.99, cwd=tmp_path)
    assert r.returncode != 0 and ("no usable CUDA device" in r.stderr or "no CUDA device" in r.stderr)


@pytest.mark.gpu
@pytest.mark.parametrize("name", helpers.GOLDEN_NAMES + ["k70_global_paired", "k127_local_single"])
def test_speq_all_matches_golden(speq_bin, tmp_path, name):
    fx = helpers.load_golden(name)
    exp, runp = fx["expected"], fx["run"]
    refs, groups, r1, r2 = helpers.golden_files(tmp_path, fx)
    cmd = ["all", "-r", refs, "-g", groups, "-1", r1, "-x", tmp_path / "i.idx", "-k", runp["k"], "--fixed-accuracy", runp["fixed_accuracy"],
           "--phred-cutoff", runp.get("phred_cutoff", 30), "--json", tmp_path / "out.json", "--batch-reads", 64]
    if r2 is not None:
        cmd += ["-2", r2]
    r = run(speq_bin, *cmd, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    out = json.loads((tmp_path / "out.json").read_text())
    assert out["unique_ref"] == exp["unique_ref"] and out["total_ref"] == exp["total_ref"]
    assert out["total_kmers"] == exp["total_kmers"] and out["ambiguous"] == exp["ambiguous"] and out["lookups"] == exp["lookups"]
    if exp["mode"] == "global":
        assert out["unique"] == exp["unique"]
        assert out["percent_initial"] == exp["percent_initial"]
    else:
        assert helpers.rel_close(out["unique_w"], exp["unique_w"])
    assert out["iterations"] == exp["iterations"]
    assert helpers.rel_close(out["percent_final"], exp["percent_final"], 1e-12)
    # stderr shape: first the back-check vectors (the .dat was built), then percent / totals lines, then the iterations
    lines = r.stderr.strip().split("\n")
    assert lines[0] == "[" + ",".join(map(str, exp["unique_ref"])) + "]"
    assert lines[1] == "[" + ",".join(map(str, exp["total_ref"])) + "]"
    assert f"{exp['total_kmers']}\t{exp['ambiguous']}" in lines
    dat = tmp_path / f"i_{runp['k']}mer.dat"
    assert dat.exists()
    # -o is parsed and, like in the reference, not written unless --write-output asks for it
    assert not (tmp_path / "output.txt").exists()
    # second run: .dat is fresh, so the back-check vectors are not printed again (src/fm_scanner.cpp:101-119)
    cmd[0] = "scan"
    cmd += ["-o", tmp_path / "table.tsv", "--write-output"]
    cmd2 = [c for i, c in enumerate(cmd) if not (i > 0 and str(cmd[i - 1]) in ("-r", "-g")) and str(c) not in ("-r", "-g")]
    r2nd = run(speq_bin, *cmd2, "-f", cwd=tmp_path)
    assert r2nd.returncode == 0, r2nd.stderr
    assert r2nd.stderr.strip().split("\n")[0] != lines[0] or exp["unique_ref"] == exp["total_ref"]
    # sharded over 2 "GPUs" requested on a 1-GPU box clamps to the devices present; result unchanged
    out2 = json.loads((tmp_path / "out.json").read_text())
    assert out2["total_kmers"] == exp["total_kmers"]
    # the per-variant table of --write-output carries the same numbers at full precision
    rows = [ln.split("\t") for ln in (tmp_path / "table.tsv").read_text().strip().split("\n")]
    assert rows[0] == ["# total_kmers", str(exp["total_kmers"])] and rows[1] == ["# ambiguous_reads", str(exp["ambiguous"])]
    assert rows[2] == ["group", "isolates", "unique_kmers", "percent_initial", "percent_final"]
    nv = len(out2["percent_final"])
    assert [float(r[4]) for r in rows[3:3 + nv]] == out2["percent_final"]
    assert [float(r[3]) for r in rows[3:3 + nv]] == out2["percent_initial"]
    # then the fusion report: ambiguous reads / pairs by the variants that met in them; the counts sum to ambiguous_reads
    assert rows[3 + nv][0].startswith("# ambiguous") and rows[4 + nv] == ["fusion", "groups", "count"]
    assert sum(int(r[2]) for r in rows[5 + nv:]) == exp["ambiguous"]
    # the sets are the oracle's fix-forward map: variant names V0, V1, ... joined by commas
    names, _, _ = helpers.parse_groups(fx["groups_txt"])
    got = {",".join(str(names.index(x)) for x in r[1].split(",")): int(r[2]) for r in rows[5 + nv:]}
    assert got == exp["fusion"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["k70_local_single", "k70_local_paired"])
def test_speq_all_default_invocation(speq_bin, tmp_path, name):
    """`speq all` with NO -k, NO --fixed-accuracy and NO --phred-cutoff: the reference's defaults kmer{70}, fixed_accuracy{0.0}
    (= local-error mode, src/fm_scanner.cpp:8-17) and phred_cutoff{30} (include/arg_parse.h:20-23).  The fixtures were generated
    with exactly those values by oracle/_ref."""
    fx = helpers.load_golden(name)
    exp, runp = fx["expected"], fx["run"]
    assert runp["k"] == 70 and runp["fixed_accuracy"] == 0.0 and runp.get("phred_cutoff", 30) == 30
    refs, groups, r1, r2 = helpers.golden_files(tmp_path, fx)
    cmd = ["all", "-r", refs, "-g", groups, "-1", r1, "-x", tmp_path / "i.idx", "--json", tmp_path / "out.json"]
    if r2 is not None:
        cmd += ["-2", r2]
    r = run(speq_bin, *cmd, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    out = json.loads((tmp_path / "out.json").read_text())
    assert out["unique_ref"] == exp["unique_ref"] and out["total_ref"] == exp["total_ref"]
    assert out["total_kmers"] == exp["total_kmers"] and out["ambiguous"] == exp["ambiguous"] and out["lookups"] == exp["lookups"]
    assert helpers.rel_close(out["unique_w"], exp["unique_w"], 1e-12)
    assert out["iterations"] == exp["iterations"]
    assert helpers.rel_close(out["percent_final"], exp["percent_final"], 1e-12)
    assert (tmp_path / "i_70mer.dat").exists()            # <stem>_<k>mer.dat with the default k (src/fm_scanner.cpp:81-82)


@pytest.mark.gpu
def test_speq_scan_on_reference_written_index(speq_bin, tmp_path):
    """`speq scan -x` on an .idx the REFERENCE wrote (tests/golden/ref_global_single.idx: cereal + sdsl bytes, make_ref_idx.py):
    the texts are decoded from the archived wavelet tree, and the scan reports the fixture's numbers.  The <stem>_<k>mer.dat it
    writes has the reference's layout (src/fm_scanner.cpp:1561-1571): int64 index time, then two u64-length-prefixed u64 vectors."""
    import shutil
    import struct
    fx = helpers.load_golden("global_single")
    exp, runp = fx["expected"], fx["run"]
    refs, groups, r1, _ = helpers.golden_files(tmp_path, fx)            # refs.fa must exist: the archive names it (src/fm_scanner.cpp:60-72)
    shutil.copy(helpers.GOLDEN / "ref_global_single.idx", tmp_path / "fromref.idx")
    r = run(speq_bin, "scan", "-1", r1, "-x", tmp_path / "fromref.idx", "-k", runp["k"], "--fixed-accuracy", runp["fixed_accuracy"],
            "--json", tmp_path / "out.json", cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    out = json.loads((tmp_path / "out.json").read_text())
    assert out["unique_ref"] == exp["unique_ref"] and out["total_ref"] == exp["total_ref"]
    assert out["unique"] == exp["unique"] and out["total_kmers"] == exp["total_kmers"] and out["ambiguous"] == exp["ambiguous"]
    assert out["iterations"] == exp["iterations"] and helpers.rel_close(out["percent_final"], exp["percent_final"], 1e-12)
    raw = (tmp_path / f"fromref_{runp['k']}mer.dat").read_bytes()
    V = len(exp["unique_ref"])
    assert len(raw) == 8 + 2 * (8 + 8 * V)
    t, n1 = struct.unpack_from("<qQ", raw, 0)
    assert n1 == V and list(struct.unpack_from(f"<{V}Q", raw, 16)) == exp["unique_ref"]
    n2 = struct.unpack_from("<Q", raw, 16 + 8 * V)[0]
    assert n2 == V and list(struct.unpack_from(f"<{V}Q", raw, 24 + 8 * V)) == exp["total_ref"]
    assert t == os.stat(tmp_path / "fromref.idx").st_mtime_ns or t != 0


@pytest.mark.gpu
@pytest.mark.parametrize("damage", ["none", "multiline", "crlf_blank"])
def test_speq_scan_device_fastq_path(speq_bin, tmp_path, damage):
    """Single-end plain FASTQ takes the device parser (speqg_scan_submit_fastq) in regions of one staging buffer; with
    --batch-reads 64 the file spans several regions.  A multi-line record / blank lines make their region irregular: it is handed
    back and read with the host grammar.  Same numbers as with SPEQ_HOST_FASTQ=1 (the host indexer) and as the fixture."""
    fx = helpers.load_golden("global_single")
    exp, runp = fx["expected"], fx["run"]
    refs, groups, r1, _ = helpers.golden_files(tmp_path, fx)
    seqs, quals = helpers.parse_fastq(fx["reads1_fq"])
    j = next(i for i, s in enumerate(seqs) if len(s) >= 10 and i > 100)
    lines = []
    for i, (sq, ql) in enumerate(zip(seqs, quals)):
        if damage == "multiline" and i == j:
            lines += [f"@read_{i}", sq[:5], sq[5:], "+", ql[:5], ql[5:]]
        else:
            lines += [f"@read_{i}", sq, "+", ql]
        if damage == "crlf_blank" and i == j:
            lines.append("")
    nl = "\r\n" if damage == "crlf_blank" else "\n"
    (tmp_path / "x.fq").write_bytes((nl.join(lines) + nl).encode("latin-1"))
    outs = []
    for env_host in (False, True):
        env = dict(os.environ, SPEQ_HOST_FASTQ="1") if env_host else dict(os.environ)
        r = subprocess.run([str(speq_bin), "all", "-r", str(refs), "-g", str(groups), "-1", str(tmp_path / "x.fq"), "-x", str(tmp_path / "i.idx"),
                            "-k", str(runp["k"]), "--fixed-accuracy", str(runp["fixed_accuracy"]), "--json", str(tmp_path / "o.json"),
                            "--batch-reads", "64", "-f"], capture_output=True, text=True, cwd=tmp_path, env=env)
        assert r.returncode == 0, r.stderr
        outs.append(json.loads((tmp_path / "o.json").read_text()))
    for out in outs:
        assert out["unique"] == exp["unique"] and out["total_kmers"] == exp["total_kmers"] and out["ambiguous"] == exp["ambiguous"]
        assert out["lookups"] == exp["lookups"] and out["iterations"] == exp["iterations"]
        assert helpers.rel_close(out["percent_final"], exp["percent_final"], 1e-12)


def _write_fq(path, recs, newline="\n", trailing=True):
    txt = newline.join(newline.join(["@" + h, s, "+", q]) for h, s, q in recs)
    path.write_text(txt + (newline if trailing else ""), newline="")


@pytest.mark.parametrize("threads,region", [(1, 1 << 20), (4, 4096), (7, 700), (3, 64)])
def test_parallel_fastq_indexer_matches_serial_reader(speq_bin, tmp_path, threads, region):
    """Host-only: the mapped, multi-threaded 4-line FASTQ indexer hands out exactly the records (order, bytes) the serial
    seqan3-grammar reader does -- quality lines that start with '@' or '+', CRLF, no trailing newline, ragged lengths,
    empty reads -- and hands over to the serial reader in front of a multi-line record."""
    import random
    rnd = random.Random(threads * 1000 + region)
    qchars = "@+!#IJ5?"

    def rec(i, L):
        s = "".join(rnd.choice("ACGTN") for _ in range(L))
        q = "".join(rnd.choice(qchars) for _ in range(L))
        return (f"r{i} desc @+ x", s, q)
    recs = [rec(i, rnd.choice([0, 1, 5, 36, 150, 151, 400])) for i in range(600)]
    for name, nl, trailing in (("a.fq", "\n", True), ("b.fastq", "\r\n", True), ("c.fq", "\n", False)):
        p = tmp_path / name
        _write_fq(p, recs, nl, trailing)
        r = run(speq_bin, "ingest-check", p, threads, region)
        assert r.returncode == 0, r.stdout + r.stderr
        assert r.stdout.split("\n")[0].split()[1] == "600" and "handover 0" in r.stdout
    # a multi-line record in the middle: regular records before it go through the indexer, the rest through the serial reader
    p = tmp_path / "d.fq"
    head = "".join(f"@{h}\n{s}\n+\n{q}\n" for h, s, q in recs[:300])
    multi = "@m\nACGT\nACGT\n+\nIIII\nIIII\n"
    tail = "".join(f"@{h}\n{s}\n+\n{q}\n" for h, s, q in recs[300:])
    p.write_text(head + multi + tail)
    r = run(speq_bin, "ingest-check", p, threads, region)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.split("\n")[0].split()[1] == "601"
    assert int(r.stdout.split("handover")[1]) > 0


@pytest.mark.gpu
def test_speq_scan_two_gpus_same_result(speq_bin, tmp_path):
    """`--gpus 2`: the batches / FASTQ regions go round-robin to two devices with the index replicated, the per-device tallies and
    the per-iteration refinement estimates are summed on the host (the reference sums its futures, src/fm_scanner.cpp:224-233).
    Same numbers as on one device.  Needs two visible GPUs (skipped on the one-GPU box)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    fx = helpers.load_golden("global_single")
    exp, runp = fx["expected"], fx["run"]
    refs, groups, r1, _ = helpers.golden_files(tmp_path, fx)
    for env_host in (False, True):
        env = dict(os.environ, SPEQ_HOST_FASTQ="1") if env_host else dict(os.environ)
        r = subprocess.run([str(speq_bin), "all", "-r", str(refs), "-g", str(groups), "-1", str(r1), "-x", str(tmp_path / "i.idx"), "-k", str(runp["k"]),
                            "--fixed-accuracy", str(runp["fixed_accuracy"]), "--json", str(tmp_path / "o.json"), "--batch-reads", "32", "--gpus", "2",
                            "-f", "--write-output", "-o", str(tmp_path / "t.tsv")], capture_output=True, text=True, cwd=tmp_path, env=env)
        assert r.returncode == 0, r.stderr
        out = json.loads((tmp_path / "o.json").read_text())
        assert out["unique"] == exp["unique"] and out["total_kmers"] == exp["total_kmers"] and out["ambiguous"] == exp["ambiguous"]
        assert out["iterations"] == exp["iterations"] and helpers.rel_close(out["percent_final"], exp["percent_final"], 1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("damage", ["none", "multiline_in_mate2", "two_multiline_in_mate2", "crlf"])
def test_speq_scan_device_fastq_paired(speq_bin, tmp_path, damage):
    """Paired plain FASTQ: regions of both files with equal record counts go to the device (speqg_scan_submit_fastq_pair); a
    multi-line record in the second file ends the device path in front of its region (the cut is no record start) or gets its
    region pair handed back, and the host readers take over.  Two 6-line records next to each other keep the LINE count a
    multiple of four while the region holds one record less than its mate region: the files lose step there, the device tallies
    are dropped (speqg_scan_reset) and both files are read again by the host readers.  Same numbers as SPEQ_HOST_FASTQ=1 and
    as the fixture."""
    fx = helpers.load_golden("global_paired")
    exp, runp = fx["expected"], fx["run"]
    refs, groups, _, _ = helpers.golden_files(tmp_path, fx)
    nl = "\r\n" if damage == "crlf" else "\n"
    for name, text, dmg in (("a_1.fq", fx["reads1_fq"], False), ("a_2.fq", fx["reads2_fq"], "multiline_in_mate2" in damage)):
        seqs, quals = helpers.parse_fastq(text)
        j = next(i for i, s in enumerate(seqs[:-1]) if len(s) >= 10 and len(seqs[i + 1]) >= 10 and i > 40)
        hit = (j, j + 1) if damage.startswith("two_") else (j,)
        lines = []
        for i, (sq, ql) in enumerate(zip(seqs, quals)):
            lines += [f"@read_{i}", sq[:5], sq[5:], "+", ql[:5], ql[5:]] if dmg and i in hit else [f"@read_{i}", sq, "+", ql]
        (tmp_path / name).write_bytes((nl.join(lines) + nl).encode("latin-1"))
    for env_host in (False, True):
        env = dict(os.environ, SPEQ_HOST_FASTQ="1") if env_host else dict(os.environ)
        r = subprocess.run([str(speq_bin), "all", "-r", str(refs), "-g", str(groups), "-1", str(tmp_path / "a_1.fq"), "-2", str(tmp_path / "a_2.fq"),
                            "-x", str(tmp_path / "i.idx"), "-k", str(runp["k"]), "--fixed-accuracy", str(runp["fixed_accuracy"]),
                            "--json", str(tmp_path / "o.json"), "--batch-reads", "32", "-f"], capture_output=True, text=True, cwd=tmp_path, env=env)
        assert r.returncode == 0, r.stderr
        assert ("lose step" in r.stderr) == (damage.startswith("two_") and not env_host), r.stderr
        out = json.loads((tmp_path / "o.json").read_text())
        assert out["unique"] == exp["unique"] and out["total_kmers"] == exp["total_kmers"] and out["ambiguous"] == exp["ambiguous"]
        assert out["lookups"] == exp["lookups"] and out["iterations"] == exp["iterations"]
        assert helpers.rel_close(out["percent_final"], exp["percent_final"], 1e-12)
