"""Shared test helpers: build/run the CPU oracle (TEST INFRASTRUCTURE), tiny FASTA/FASTQ writers, random inputs."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
GOLDEN = ROOT / "tests" / "golden"
REFERENCE = Path("/root/reference")


def _stale(target: Path, src: Path) -> bool:
    return (not target.exists()) or target.stat().st_mtime < src.stat().st_mtime


def oracle_binary() -> Path:
    b = ORACLE_DIR / "_build" / "speq_oracle"
    if _stale(b, ORACLE_DIR / "speq_oracle.cpp"):
        subprocess.run(["make", "-C", str(ORACLE_DIR), "_build/speq_oracle"], check=True, capture_output=True)
    return b


def ref_binary() -> Path | None:
    b = ORACLE_DIR / "_ref" / "speq_ref"
    if REFERENCE.exists() and _stale(b, ORACLE_DIR / "speq_oracle.cpp"):
        subprocess.run(["make", "-C", str(ORACLE_DIR), "ref"], check=True, capture_output=True)
    return b if b.exists() else None


def sim_library() -> C.CDLL:
    """tests/sim/roll_sim.cpp: the per-lane arithmetic of the CUDA kernels (scan_core.h) compiled for the host, as a
    white-box check of index layout + rolling window in the CPU tier.  Not a product path."""
    src = ROOT / "tests" / "sim" / "roll_sim.cpp"
    so = ROOT / "tests" / "sim" / "roll_sim.so"
    csrc = ROOT / "speq_b200" / "csrc"
    deps = [src, csrc / "scan_core.h", csrc / "scan32_core.h", csrc / "fm_layout.h", csrc / "index_build.cpp", csrc / "sais.hpp", csrc / "speq_internal.h"]
    if not so.exists() or any(d.stat().st_mtime > so.stat().st_mtime for d in deps):
        subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-pthread", "-o", str(so), str(src),
                        str(csrc / "index_build.cpp")], check=True, capture_output=True)
    return C.CDLL(os.fspath(so))


def run_oracle(binary: Path, refs: Path, groups: Path, reads1: Path, *, k: int, reads2: Path | None = None,
               phred_cutoff: int = 30, fixed_accuracy: float = 0.0, precision: float = 1e-6, threads: int = 4,
               extra: list[str] | None = None) -> dict:
    cmd = [str(binary), "run", "--refs", str(refs), "--groups", str(groups), "--reads1", str(reads1), "-k", str(k),
           "--phred-cutoff", str(phred_cutoff), "--fixed-accuracy", repr(float(fixed_accuracy)), "--precision", repr(precision),
           "--threads", str(threads)]
    if reads2 is not None:
        cmd += ["--reads2", str(reads2)]
    if extra:
        cmd += extra
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"oracle failed: {r.stderr}")
    return json.loads(r.stdout)


def run_oracle_jobs(binary: Path, refs: Path, groups: Path, jobs: list[dict], *, threads: int = 4, backcheck: bool = True) -> list[dict]:
    """Several scan jobs against ONE index build of the oracle (`--next` between them; a job inherits the settings of the one
    before it).  Job keys: reads1, reads2, k, phred_cutoff, fixed_accuracy, precision, max_reads, refine (bool), max_iters,
    unique_ref / total_ref (lists: injected back-check counts)."""
    cmd = [str(binary), "run", "--refs", str(refs), "--groups", str(groups), "--threads", str(threads)]
    if not backcheck:
        cmd.append("--no-backcheck")
    for i, j in enumerate(jobs):
        if i:
            cmd.append("--next")
        if "reads1" in j:
            cmd += ["--reads1", str(j["reads1"])]
        if j.get("reads2") is not None:
            cmd += ["--reads2", str(j["reads2"])]
        for key, flag in (("k", "-k"), ("phred_cutoff", "--phred-cutoff"), ("max_reads", "--max-reads"), ("max_iters", "--max-iters")):
            if key in j:
                cmd += [flag, str(int(j[key]))]
        for key, flag in (("fixed_accuracy", "--fixed-accuracy"), ("precision", "--precision")):
            if key in j:
                cmd += [flag, repr(float(j[key]))]
        if "refine" in j:
            cmd.append("--refine" if j["refine"] else "--no-refine")
        for key, flag in (("unique_ref", "--unique-ref"), ("total_ref", "--total-ref")):
            if key in j:
                cmd += [flag, ",".join(str(int(x)) for x in j[key])]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"oracle failed: {r.stderr}")
    out = json.loads(r.stdout)
    return out if isinstance(out, list) else [out]


def run_oracle_search(binary: Path, sigma: int, texts: list[str], queries: list[list[str]], want_size: bool = False) -> list[dict]:
    s = f"sigma {sigma}\n" + "".join(f"text {t}\n" for t in texts)
    if want_size:
        s += "size\n"
    for q in queries:
        s += "query " + " ".join(p if p else '""' for p in q) + "\n"
    r = subprocess.run([str(binary), "search"], input=s, capture_output=True, text=True, check=True)
    return json.loads(r.stdout)


def write_fasta(path: Path, seqs: list[str], width: int = 60) -> None:
    with open(path, "w") as f:
        for j, s in enumerate(seqs):
            f.write(f">scaffold_{j} synthetic\n")
            for i in range(0, len(s), width):
                f.write(s[i:i + width] + "\n")
            if not s:
                f.write("\n")


def write_fastq(path: Path, seqs: list[str], quals: list[str]) -> None:
    with open(path, "w") as f:
        for i, (s, q) in enumerate(zip(seqs, quals)):
            f.write(f"@read_{i}\n{s}\n+\n{q}\n")


def write_groups(path: Path, names: list[str], counts: list[int], members: list[list[int]]) -> None:
    """`name(count): a-b, c` grammar of src/file_to_map.cpp."""
    with open(path, "w") as f:
        f.write("# groups file\n")
        for n, c, m in zip(names, counts, members):
            toks, i = [], 0
            m = sorted(m)
            while i < len(m):
                j = i
                while j + 1 < len(m) and m[j + 1] == m[j] + 1:
                    j += 1
                toks.append(f"{m[i]}-{m[j]}" if j > i else f"{m[i]}")
                i = j + 1
            f.write(f"{n}({c}): {', '.join(toks)}\n")


def random_case(seed: int, *, V: int = 3, I: int = 2, G: int = 400, R: int = 200, L: int = 50, e: float = 0.01,
                d_var: float = 0.05, d_iso: float = 0.01, n_rate_ref: float = 0.0, n_rate_read: float = 0.0,
                lowq_rate: float = 0.0, ragged: bool = False, local_quals: bool = False, iupac: bool = False):
    """Small random variant collection + reads as strings.  Returns dict(scaffolds, group_scaffolds, V, seqs, quals)."""
    rng = np.random.default_rng(seed)
    alpha = np.array(list("ACGT"))

    def mutate(a, rate):
        a = a.copy()
        hit = rng.random(a.size) < rate
        a[hit] = (a[hit] + rng.integers(1, 4, hit.sum())) & 3
        return a
    anc = rng.integers(0, 4, G)
    codes = []
    for v in range(V):
        var = mutate(anc, d_var)
        for _ in range(I):
            codes.append(mutate(var, d_iso))
    scaffolds = []
    for c in codes:
        s = alpha[c]
        if n_rate_ref > 0:
            s = np.where(rng.random(G) < n_rate_ref, "N", s)
        scaffolds.append("".join(s))
    ab = 1.0 / (np.arange(V) + 1.0)
    ab /= ab.sum()
    seqs, quals = [], []
    for _ in range(R):
        v = rng.choice(V, p=ab)
        i = rng.integers(0, I)
        Lr = int(rng.integers(0, L + 1)) if (ragged and rng.random() < 0.2) else L
        st = int(rng.integers(0, G - Lr + 1))
        c = codes[v * I + i][st:st + Lr]
        if rng.random() < 0.5:
            c = (3 - c)[::-1]
        c = mutate(c, e)
        s = alpha[c] if Lr else np.array([], dtype="<U1")
        if n_rate_read > 0 and Lr:
            sub = np.array(list("NRYKMSWBDHVn"))[rng.integers(0, 12, Lr)] if iupac else np.full(Lr, "N")
            s = np.where(rng.random(Lr) < n_rate_read, sub, s)
        if iupac and Lr:
            low = rng.random(Lr) < 0.1
            s = np.where(low, np.char.lower(s), s)
            s = np.where((s == "T") & (rng.random(Lr) < 0.1), "U", s)
        if local_quals == "high":
            # the reference's default invocation (no --fixed-accuracy, --phred-cutoff 30): qualities above the cutoff that
            # still differ per base, so the phred-weighted tallies are not a constant times the counts
            u = rng.random(Lr)
            q = np.where(u < 0.10, 33, np.where(u < 0.40, 37, 40))
        elif local_quals:
            u = rng.random(Lr)
            q = np.where(u < 0.05, 20, np.where(u < 0.20, 30, 37))
        else:
            q = np.full(Lr, 40)
        if lowq_rate > 0 and Lr:
            q = np.where(rng.random(Lr) < lowq_rate, rng.integers(0, 31, Lr), q)
        seqs.append("".join(s))
        quals.append("".join(chr(33 + int(x)) for x in q))
    return {"scaffolds": scaffolds, "group_scaffolds": np.repeat(np.arange(V), I).astype(np.int32), "V": V, "I": I,
            "seqs": seqs, "quals": quals}


def case_files(tmp: Path, case: dict, paired: bool = False):
    tmp.mkdir(parents=True, exist_ok=True)
    V, I = case["V"], case["I"]
    write_fasta(tmp / "refs.fa", case["scaffolds"])
    write_groups(tmp / "groups.txt", [f"V{v}" for v in range(V)], case.get("group_counts", [I] * V),
                 [[int(j) for j in np.nonzero(case["group_scaffolds"] == v)[0]] for v in range(V)])
    if paired:
        write_fastq(tmp / "r1.fq", case["seqs"][0::2], case["quals"][0::2])
        write_fastq(tmp / "r2.fq", case["seqs"][1::2], case["quals"][1::2])
        return tmp / "refs.fa", tmp / "groups.txt", tmp / "r1.fq", tmp / "r2.fq"
    write_fastq(tmp / "r1.fq", case["seqs"], case["quals"])
    return tmp / "refs.fa", tmp / "groups.txt", tmp / "r1.fq", None


def write_fastq_fixed(path: Path, seq: np.ndarray, qual: np.ndarray) -> None:
    """4-line FASTQ of equal-length reads ([n, L] ASCII arrays), written as one fixed-width byte matrix."""
    n, L = seq.shape
    rec = np.empty((n, 3 + L + 3 + L + 1), np.uint8)
    rec[:, 0:3] = np.frombuffer(b"@r\n", np.uint8)
    rec[:, 3:3 + L] = seq
    rec[:, 3 + L:6 + L] = np.frombuffer(b"\n+\n", np.uint8)
    rec[:, 6 + L:6 + 2 * L] = qual
    rec[:, 6 + 2 * L] = 10
    rec.tofile(str(path))


def flat_reads(seqs: list[str], quals: list[str] | None):
    """-> (seq uint8 flat, qual uint8 flat | None, offsets uint64)"""
    offs = np.zeros(len(seqs) + 1, np.uint64)
    offs[1:] = np.cumsum([len(s) for s in seqs])
    seq = np.frombuffer("".join(seqs).encode(), np.uint8).copy()
    q = np.frombuffer("".join(quals).encode("latin-1"), np.uint8).copy() if quals is not None else None
    return seq, q, offs


def parse_fasta(text: str) -> list[str]:
    out, cur = [], None
    for ln in text.splitlines():
        if ln.startswith(">"):
            if cur is not None:
                out.append("".join(cur))
            cur = []
        elif cur is not None:
            cur.append(ln.strip())
    if cur is not None:
        out.append("".join(cur))
    return out


def parse_fastq(text: str):
    lines = text.split("\n")
    seqs, quals = [], []
    for i in range(0, len(lines) - 3, 4):
        assert lines[i].startswith("@") and lines[i + 2].startswith("+")
        seqs.append(lines[i + 1])
        quals.append(lines[i + 3])
    return seqs, quals


def parse_groups(text: str):
    """Minimal reader of the `name(count): a-b, c` grammar for test fixtures written by write_groups."""
    names, counts, gs = [], [], {}
    for ln in text.splitlines():
        ln = ln.split("#")[0]
        if ":" not in ln:
            continue
        head, tail = ln.split(":", 1)
        names.append(head[:head.index("(")])
        counts.append(int(head[head.index("(") + 1:head.index(")")]))
        for tok in tail.split(","):
            tok = tok.replace(" ", "").replace("\t", "")
            if "-" in tok:
                a, b = tok.split("-")
                for j in range(int(a), int(b) + 1):
                    gs[j] = len(names) - 1
            elif tok:
                gs[int(tok)] = len(names) - 1
    arr = np.full(max(gs) + 1, -1, np.int32)
    for j, g in gs.items():
        arr[j] = g
    return names, counts, arr


def load_golden(name: str) -> dict:
    return json.loads((GOLDEN / f"speq_{name}.json").read_text())


GOLDEN_NAMES = ["global_single", "local_single", "global_paired", "local_paired", "refs_with_n", "short_k"]
# k > 31: the rolling FM search is the product path (the k-mer table covers k <= 31).  k = 70 in local-error mode with the
# default --phred-cutoff is the reference's default invocation (include/arg_parse.h:20-23)
GOLDEN_LONG_K = ["k70_global_single", "k70_local_single", "k70_global_paired", "k70_local_paired", "k127_local_single", "k33_global_paired"]


def golden_files(tmp: Path, fx: dict):
    tmp.mkdir(parents=True, exist_ok=True)
    (tmp / "refs.fa").write_text(fx["refs_fa"])
    (tmp / "groups.txt").write_text(fx["groups_txt"])
    (tmp / "r1.fq").write_text(fx["reads1_fq"])
    r2 = None
    if fx["reads2_fq"] is not None:
        (tmp / "r2.fq").write_text(fx["reads2_fq"])
        r2 = tmp / "r2.fq"
    return tmp / "refs.fa", tmp / "groups.txt", tmp / "r1.fq", r2


def golden_inputs(fx: dict):
    """-> (scaffold rank arrays, group_scaffolds, group_counts, seqs, quals [interleaved when paired])"""
    from speq_b200 import engine
    scaff = [engine.ascii_to_ranks(s) for s in parse_fasta(fx["refs_fa"])]
    _, counts, gs = parse_groups(fx["groups_txt"])
    s1, q1 = parse_fastq(fx["reads1_fq"])
    if fx["reads2_fq"] is not None:
        s2, q2 = parse_fastq(fx["reads2_fq"])
        n = min(len(s1), len(s2))
        seqs = [x for p in zip(s1[:n], s2[:n]) for x in p]
        quals = [x for p in zip(q1[:n], q2[:n]) for x in p]
    else:
        seqs, quals = s1, q1
    return scaff, gs, counts, seqs, quals


def rel_close(a, b, tol=1e-12) -> bool:
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return bool(np.all(np.abs(a - b) <= tol * np.maximum(np.abs(a), np.abs(b))))
