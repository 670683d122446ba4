"""Wall-clock of the host CLI (`speq all`, then `speq scan` on the cached index) on a synthetic C2-shaped FASTQ file:
index build, back-check, parallel FASTQ ingest, scan and refinement through the C ABI.  Usage: python profiles/cli_e2e.py [reads] [paired]   (paired: the reads are written as two mate files)"""
import json, os, subprocess, sys, time, tempfile
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from speq_b200 import synth, build

R = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
PAIRED = len(sys.argv) > 2 and sys.argv[2] == "paired"
sh = synth.CONFIGS["C2"]
exe = build.CLI_BIN
with tempfile.TemporaryDirectory(dir="/dev/shm") as d:
    d = Path(d)
    refs = synth.make_refs(sh, synth.SEED0 + 2)
    synth.write_fasta(str(d / "refs.fa"), refs)
    synth.write_groups(str(d / "groups.txt"), sh)
    seq, qual = synth.make_reads(sh, refs, R, 11)
    t0 = time.time()
    if PAIRED:
        synth.write_fastq(str(d / "reads.fq"), seq.numpy()[0::2], qual.numpy()[0::2])
        synth.write_fastq(str(d / "reads_2.fq"), seq.numpy()[1::2], qual.numpy()[1::2])
    else:
        synth.write_fastq(str(d / "reads.fq"), seq.numpy(), qual.numpy())
    tw = time.time() - t0
    size = (d / "reads.fq").stat().st_size + ((d / "reads_2.fq").stat().st_size if PAIRED else 0)
    mates = ["-2", "reads_2.fq"] if PAIRED else []
    out = {"reads": R, "paired": PAIRED, "fastq_bytes": size, "write_s": tw}
    for name, cmd in (("all", ["all", "-r", "refs.fa", "-g", "groups.txt"]), ("scan", ["scan"])):
        for t in (16,) if name == "all" else (2, 16):
            c = [str(exe), *cmd, "-1", "reads.fq", *mates, "-x", "i.idx", "-k", str(sh.k), "--fixed-accuracy", "0.995", "-t", str(t), "--json", "o.json"]
            t0 = time.time(); r = subprocess.run(c, cwd=d, capture_output=True, text=True, env=dict(os.environ, SPEQ_TIMING="1")); dt = time.time() - t0
            assert r.returncode == 0, r.stderr[-2000:]
            o = json.loads((d / "o.json").read_text())
            out[f"{name}_t{t}"] = {"wall_s": dt, "reads_per_s_wall": R / dt, "scan_s": o.get("t_scan_s"), "backcheck_s": o.get("t_backcheck_s"), "refine_s": o.get("t_refine_s"), "total_kmers": o["total_kmers"],
                                   "iterations": o["iterations"],
                                   "phases": [ln.split("]", 1)[1].split() for ln in r.stderr.splitlines() if ln.startswith("[timing]")]}
print(json.dumps(out))
