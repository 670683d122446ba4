#!/usr/bin/env python
"""bench.py -- throughput of the SPeQ scan hot path on B200 (k-mer lookups/s; reads/s alongside).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--config C3] [--reads R] [--mode global|local] [-k K]

Default workload = the configuration BASELINE.json's target is quoted on: C3 (50 variants x 10 isolates x 30 kb, 200 M x
150 bp reads at 0.5 % error, k = 25), STRONG scaling: the same 200 M reads at every N, sharded over the N GPUs (one rank per
GPU under torchrun, index replicated, no data-path collective; the only collective is the all-reduce of the small tally
vector at the end of a step).  A step = one scan pass over all the reads.  `--config C1|C2|C4|C5` selects the other shapes.

  value     lookups/s of the scan AS `speq scan` RUNS IT (record_multiplicity = 1: the refinement's per-k-mer counts are kept),
            packed reads resident in HBM, CUDA events on the scan's stream, max over ranks.  `tallies_only` = the same without
            the recording.
  e2e       the same through the C ABI from pinned HOST buffers (ASCII reads + quality chars) in 1 M-read submits: H2D copy,
            device packer, scan, D2H of the tallies inside the timed region; on the first `--e2e-reads` reads of every rank
  roofline  the dominant kernel against the roofline its access pattern has: independent random 32-byte sectors (one per
            k-mer table probe / two per rank step).  achieved = the kernel's own sectors x 32 B / its CUDA-event time;
            peak = the random-sector rate of THIS device for a table of the probed structure's size, measured in the same
            process (speqg_probe_random_sectors) x 32 B.  bound says whether that structure is L2- or HBM-resident.
            `reference_work_rate` keeps SURVEY.md 8d's figure: the algorithmic bytes of the reference's own search
            (64 B per executed backward step, 32 B per 32 located occurrences, 0.5 B per base) over the same time.
  parity    the oracle's tallies on the CPU-timed sample == the GPU's tallies on the same reads (default algorithm)
  cpu_baseline  oracle/_ref (the reference's glue over the real vendored sdsl-lite objects) when it was built, else the
            port; all host threads; a bounded sample of the same reads
`--impl reference` times only that CPU arm and prints the same line with "impl": "reference".
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "kmer_lookups_per_sec"
UNIT = "lookups/s"
CHUNK_BASES = 150_000_000      # synthetic reads are generated in chunks of about this many bases (C3: 1 M reads)


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ---------------------------------------------------------------------------------------------------------------
# workload: a fixed global read set, generated chunk by chunk (chunk c has its own seed), so that every N scans the same reads
# ---------------------------------------------------------------------------------------------------------------
def chunk_reads(sh) -> int:
    return max(1, min(1_000_000, CHUNK_BASES // sh.L))


def shard_chunks(n_chunks: int, rank: int, world: int) -> tuple[int, int]:
    return n_chunks * rank // world, n_chunks * (rank + 1) // world


def gen_chunk(sh, cfg_id: int, refs, c: int, n: int, local: bool):
    from speq_b200 import synth
    return synth.make_reads(sh, refs, n, synth.SEED0 * 1000 + cfg_id * 100_000 + c, local_quals=local)


def write_fastq_fixed(path: Path, seq: np.ndarray, qual: np.ndarray) -> None:
    """4-line FASTQ of equal-length reads, written as one fixed-width byte matrix."""
    n, L = seq.shape
    rec = np.empty((n, 3 + L + 3 + L + 1), np.uint8)
    rec[:, 0:3] = np.frombuffer(b"@r\n", np.uint8)
    rec[:, 3:3 + L] = seq
    rec[:, 3 + L:6 + L] = np.frombuffer(b"\n+\n", np.uint8)
    rec[:, 6 + L:6 + 2 * L] = qual
    rec[:, 6 + 2 * L] = 10
    rec.tofile(str(path))


def oracle_choice():
    """(binary, kind): oracle/_ref (real sdsl objects, 'reference') when it was built, else the restatement ('port')."""
    sys.path.insert(0, str(ROOT / "tests"))
    import helpers
    ref = helpers.ref_binary()
    if ref is not None:
        return ref, "reference"
    return helpers.oracle_binary(), "port"


def bind_near_gpu(index: int):
    """One process per GPU: run on (and so first-touch / pin host memory on) the CPUs of the GPU's NUMA node."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(index)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        base = Path("/sys/bus/pci/devices") / bdf
        cpus = set()
        for part in (base / "local_cpulist").read_text().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"bdf": bdf, "numa_node": int((base / "numa_node").read_text()), "cpus": len(cpus)}
    except Exception as ex:   # affinity is an optimisation only
        return {"error": str(ex)[:80]}


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def run_cpu(binary, files, k: int, accuracy: float, cutoff: int, threads: int, n_reads: int) -> dict:
    """One scan pass of the oracle over the first n_reads reads (no back-check, no refinement): the reference's scan."""
    refs, groups, fq = files
    cmd = [str(binary), "run", "--refs", str(refs), "--groups", str(groups), "--reads1", str(fq), "-k", str(k),
           "--fixed-accuracy", repr(accuracy), "--phred-cutoff", str(cutoff), "--threads", str(threads), "--max-reads", str(n_reads),
           "--no-refine", "--no-backcheck"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle failed: " + r.stderr)
    return json.loads(r.stdout)


def cpu_arm(sh, refs_cpu, seq: np.ndarray, qual: np.ndarray, *, k: int, accuracy: float, cutoff: int, steps: int, warmup: int,
            target_s: float):
    """Times the CPU scan on a bounded sample (the first n reads of seq).  -> (lookups/s, sample text, kind, cores, s/step, out, n)"""
    from speq_b200 import synth
    binary, kind = oracle_choice()
    T = host_threads()
    wpr = max(1, sh.L - k + 1)
    with tempfile.TemporaryDirectory() as d:
        d = Path(d)
        synth.write_fasta(str(d / "refs.fa"), refs_cpu)
        synth.write_groups(str(d / "groups.txt"), sh)
        n_cal = min(seq.shape[0], 256 * T)
        write_fastq_fixed(d / "sample.fq", seq[:n_cal], qual[:n_cal])
        files = (d / "refs.fa", d / "groups.txt", d / "sample.fq")
        cal = run_cpu(binary, files, k, accuracy, cutoff, T, n_cal)
        rate = cal["lookups"] / max(cal["t_scan_s"], 1e-6)
        n = int(min(seq.shape[0], max(n_cal, rate * target_s / wpr)))
        write_fastq_fixed(d / "sample.fq", seq[:n], qual[:n])
        times, out = [], None
        for i in range(warmup + steps):
            out = run_cpu(binary, files, k, accuracy, cutoff, T, n)
            if i >= warmup:
                times.append(out["t_scan_s"])
    t = sum(times) / len(times)
    sample = (f"first {n} reads of rank 0's shard ({out['lookups']} lookups) per step, scan pass only (index build, back-check and "
              f"refinement excluded); sample sized for about {target_s:g} s on this box's {T} threads")
    return out["lookups"] / t, sample, kind, T, t, out, n


# ---------------------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons during the timed region: NVML polled every 5 ms from a thread (a `nvidia-smi -lms` child
    needs longer to print its first row than a short timed region lasts)."""
    NAMES = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, gpu_index: int):
        self.rows = []
        self.h = None
        self.stop_flag = False
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            phys = int(ids[gpu_index]) if ids and all(v.isdigit() for v in ids) and gpu_index < len(ids) else gpu_index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
        except Exception as e:
            self.h = None
            self.err = f"NVML unavailable: {e}"

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    why = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    why = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((time.perf_counter(), mhz, why))
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self, t0: float, t1: float) -> dict:
        if self.h is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [self.err], "samples": 0}
        self.stop_flag = True
        self.th.join(timeout=1.0)
        rows = [r for r in self.rows if t0 <= r[0] <= t1] or self.rows[-3:]
        sm = [r[1] for r in rows]
        reasons = sorted(nm for nm, bit in self.NAMES.items() if any(r[2] & bit for r in rows))
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(sm)}


def measured_peak() -> tuple[float, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------------------------
def emit(line: dict) -> None:
    """The ONE JSON line goes to the process's real stdout; everything else a library prints (NCCL's version banner ...)
    was sent to stderr by main()."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main() -> None:
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--config", default="C3", choices=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--reads", type=int, default=None, help="reads of the whole job per step (default: the config's read count)")
    ap.add_argument("--mode", default="global", choices=["global", "local"],
                    help="global: --fixed-accuracy 1-e, integer tallies; local: phred-weighted fp64 tallies (no --fixed-accuracy: the reference's default)")
    ap.add_argument("-k", "--kmer", type=int, default=None, help="k-mer length (default: the config's; the reference's own default is 70)")
    ap.add_argument("--batch-reads", type=int, default=1 << 20, help="reads per C-ABI submit on the e2e leg")
    ap.add_argument("--e2e-reads", type=int, default=10_000_000, help="reads of every rank's shard the host-fed legs stream per step")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the cross-checks (measurement builds whose results are meaningless)")
    ap.add_argument("--no-probe", action="store_true", help="skip the in-run random-sector probe (profiling runs)")
    ap.add_argument("--algorithm", default="auto", choices=["auto", "kpair", "ktable", "rolling", "direct"],
                    help="auto = the library's default: pair / k-mer table scan for k <= 31, else the rolling FM search")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from speq_b200 import synth
    sh = synth.CONFIGS[args.config]
    cfg_id = int(args.config[1])
    k = args.kmer or sh.k
    local = args.mode == "local"
    accuracy = 0.0 if local else sh.accuracy
    cutoff = 19 if local else 30          # local-mode reads carry Q20/Q30/Q37, scanned with --phred-cutoff 19
    R_total = args.reads or sh.R
    CH = chunk_reads(sh)
    n_chunks = (R_total + CH - 1) // CH
    config = {"workload": sh.name, "config": args.config, "reads_per_step": R_total, "read_len": sh.L, "k": k,
              "variants": sh.V, "isolates": sh.I, "genome_len": sh.G, "error_rate": sh.e,
              "mode": ("local (phred-weighted, --phred-cutoff 19)" if local else "global (--fixed-accuracy %g)" % sh.accuracy),
              "parallelism": f"reads sharded over {args.gpus} GPU(s), index replicated", "algorithm": args.algorithm,
              "l2": "inputs larger than L2 (packed reads of one step: 72 B per 150 bp read)"}
    wpr = max(0, sh.L - k + 1)

    # ------------------------------------------------------------------------------------------------- reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        import torch
        refs = synth.make_refs(sh, synth.SEED0 + cfg_id, device="cpu")
        n0 = min(R_total, CH, 400_000)
        a, b = gen_chunk(sh, cfg_id, refs, 0, n0, local)     # the head of chunk 0: the same generator and seed as the native arm
        # every step scans the same bounded sample; the sample is sized so that the whole run stays around 100 s of CPU work
        n_runs = max(1, args.steps) + min(args.warmup, 1)
        v, sample, kind, T, t, out, n = cpu_arm(sh, refs, a.numpy(), b.numpy(), k=k, accuracy=accuracy, cutoff=cutoff,
                                                steps=max(1, args.steps), warmup=min(args.warmup, 1),
                                                target_s=min(args.cpu_seconds, 100.0 / n_runs))
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64" if local else "u64",
                "data": "synthetic", "config": config, "reads_per_sec": out["reads"] / t,
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": T, "kind": kind, "sample": sample},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        emit(line)
        return

    # ---------------------------------------------------------------------------------------------------- native arm
    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py --impl native needs a CUDA device: the scan path has no CPU fallback"
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    numa = bind_near_gpu(local_rank) if world > 1 else None   # pinned staging buffers on the GPU's own NUMA node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    from speq_b200 import capi, engine

    # ---- index (deterministic synthetic references; suffix sort etc. on the GPU) ----
    t0 = time.perf_counter()
    refs = synth.make_refs(sh, synth.SEED0 + cfg_id, device=dev)
    cache = os.environ.get("SPEQ_BENCH_INDEX_CACHE")   # optional: reuse the (deterministic) synthetic index between runs
    cpath = Path(cache) / f"{args.config}.speqidx" if cache else None
    if cpath is not None and cpath.exists():
        idx = engine.Index.load(str(cpath))
    else:
        idx = engine.Index.from_scaffolds(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), sh.V, device_sa=True)
        if cpath is not None and rank == 0:
            cpath.parent.mkdir(parents=True, exist_ok=True)
            idx.save(str(cpath))
    t_index = time.perf_counter() - t0
    idx.upload(local_rank)

    # ---- this rank's shard: chunks [c0, c1) of the global read set, packed on the device into one resident batch; the first
    #      e2e_reads reads are also kept as ASCII in pinned host memory for the host-fed legs and the CPU sample ----
    t0 = time.perf_counter()
    c0, c1 = shard_chunks(n_chunks, rank, world)
    starts = [c * CH for c in range(c0, c1 + 1)]
    starts[-1] = min(starts[-1], R_total)
    R = starts[-1] - starts[0]                                   # reads of this rank
    E = 0 if args.no_e2e and args.no_cpu else min(R, args.e2e_reads)
    L = sh.L
    db = engine.ResidentBatch(local_rank, torch.full((R,), L, dtype=torch.int64, device=dev), with_qual=local)
    seq_h = torch.empty((E, L), dtype=torch.uint8, pin_memory=True) if E else None
    qual_h = torch.empty((E, L), dtype=torch.uint8, pin_memory=True) if E else None
    # the same reads as 4-line FASTQ text ("@r\n" seq "\n+\n" qual "\n"): what the file of `speq scan -1 reads.fq` holds
    FQ = 2 * L + 7
    fq_h = torch.empty((E, FQ), dtype=torch.uint8, pin_memory=True) if E and not args.no_e2e else None
    for c in range(c0, c1):
        n = min(R_total, (c + 1) * CH) - c * CH
        a, b = gen_chunk(sh, cfg_id, refs, c, n, local)
        r0 = c * CH - starts[0]
        offs = torch.arange(n + 1, dtype=torch.int64, device=dev) * L
        db.pack(r0, a.view(-1), b.view(-1), offs, phred_cutoff=cutoff)
        if r0 < E:
            m = min(n, E - r0)
            seq_h[r0:r0 + m].copy_(a[:m])
            qual_h[r0:r0 + m].copy_(b[:m])
            if fq_h is not None:
                rec = torch.empty((m, FQ), dtype=torch.uint8, device=dev)
                rec[:, 0:3] = torch.tensor(list(b"@r\n"), dtype=torch.uint8, device=dev)
                rec[:, 3:3 + L] = a[:m]
                rec[:, 3 + L:6 + L] = torch.tensor(list(b"\n+\n"), dtype=torch.uint8, device=dev)
                rec[:, 6 + L:6 + 2 * L] = b[:m]
                rec[:, 6 + 2 * L] = 10
                fq_h[r0:r0 + m].copy_(rec)
                del rec
        del a, b, offs
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    t_gen = time.perf_counter() - t0
    packed_bytes = db.nbytes()
    log(f"[rank {rank}] index n={idx.n} in {t_index:.1f}s ({idx.device_bytes/1e6:.0f} MB on device); {R} reads generated + packed on "
        f"the device in {t_gen:.1f}s ({packed_bytes/1e9:.2f} GB resident), {E} of them pinned on the host")

    stream = torch.cuda.Stream(device=dev)   # a real (non-default) stream: the scan handle launches on it and the events time it
    torch.cuda.set_stream(stream)
    algo_req = {"rolling": capi.ALGO_ROLLING, "direct": capi.ALGO_DIRECT, "ktable": capi.ALGO_KTABLE, "kpair": capi.ALGO_KPAIR,
                "auto": capi.ALGO_AUTO}[args.algorithm]
    mode = capi.MODE_LOCAL if local else capi.MODE_GLOBAL
    t0 = time.perf_counter()
    # what `speq scan` opens: multiplicities recorded for the refinement (speq_b200/host/speq.cpp)
    scan = engine.Scan(idx, k=k, phred_cutoff=cutoff, mode=mode, record=algo_req != capi.ALGO_DIRECT, algorithm=algo_req, device=local_rank,
                       stream=stream.cuda_stream)
    torch.cuda.synchronize()
    t_scan_open = time.perf_counter() - t0     # per-(index, k) specialisation: rank blocks or the k-mer table
    info = scan.info()
    algo = info["algorithm"]                   # what AUTO resolved to
    algo_name = {capi.ALGO_ROLLING: "rolling", capi.ALGO_DIRECT: "direct", capi.ALGO_KTABLE: "ktable", capi.ALGO_KPAIR: "kpair"}[algo]
    table = algo in (capi.ALGO_KTABLE, capi.ALGO_KPAIR)
    V = sh.V

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def tall_of(r):
        return np.concatenate([r.unique, np.array([r.total_kmers, r.ambiguous, r.lookups, r.reads], np.uint64)]).astype(np.int64)

    def make_step(sc):
        # N > 1: the counters are summed across ranks where they live -- one NCCL all-reduce in place on the scan's stream
        # behind the last kernel -- so a step has a single host sync (finish), which then returns the global tallies
        ci, cf = sc.device_counters() if world > 1 else (None, None)

        def finish_global():
            if world > 1:
                sc.flush()                 # the last FASTQ region's kernels are queued before the reduction, not by finish()
                dist.all_reduce(ci)
                if local:
                    dist.all_reduce(cf)
            r = sc.finish()
            return tall_of(r), r

        def step_resident():
            sc.reset()
            sc.submit_device(db)
            return finish_global()

        def step_e2e():
            sc.reset()
            for s in range(0, E, args.batch_reads):
                e = min(E, s + args.batch_reads)
                sc.submit_raw(seq_np, qual_np, offs_np[s:e + 1], e - s)
            return finish_global()

        def step_fastq():
            # raw FASTQ bytes, region by region; the record grammar runs on the device (speqg_scan_submit_fastq)
            sc.reset()
            for s in range(0, E, args.batch_reads):
                e = min(E, s + args.batch_reads)
                sc.submit_fastq(fq_np[s * FQ:e * FQ])
            return finish_global()
        return step_resident, step_e2e, step_fastq

    if E:
        seq_np, qual_np = seq_h.view(-1).numpy(), qual_h.view(-1).numpy()
        offs_np = (torch.arange(E + 1, dtype=torch.int64) * L).pin_memory().numpy().view(np.uint64)
        fq_np = fq_h.view(-1).numpy() if fq_h is not None else None
    step_resident, step_e2e, step_fastq = make_step(scan)

    # ---- value: resident inputs, CUDA events ----
    for _ in range(args.warmup):
        tall, res = step_resident()
    barrier()
    clocks = ClockSampler(local_rank)
    tc0 = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    launches = 0
    for _ in range(args.steps):
        tall, res = step_resident()
        launches += 1
    ev1.record(stream)
    barrier()
    tc1 = time.perf_counter()
    ms = ev0.elapsed_time(ev1)
    clk = clocks.stop(tc0, tc1)
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_total = float(tms.item())
    lookups_step = int(tall[V + 2])          # all ranks
    reads_step = int(tall[V + 3])
    assert reads_step == R_total and lookups_step == R_total * wpr, (reads_step, lookups_step)
    value = lookups_step * args.steps / (ms_total * 1e-3)
    clocks_all = [clk]
    if world > 1:
        clocks_all = [None] * world
        dist.all_gather_object(clocks_all, clk)

    # ---- the all-reduced tallies are the sum of the ranks' own tallies (checked outside the timed region) ----
    allreduce_check = None
    if world > 1:
        scan.reset(); scan.submit_device(db)
        own = tall_of(scan.finish())
        parts = [None] * world
        dist.all_gather_object(parts, own.tolist())
        ok = bool(np.array_equal(np.sum(np.array(parts, dtype=np.int64), axis=0), tall))
        assert ok, "all-reduced tallies differ from the sum of the per-rank tallies"
        allreduce_check = f"all-reduced unique[V], total, ambiguous, lookups, reads == sum over the {world} ranks' own tallies"

    # ---- kernel-only time of the dominant kernel (same stream, no host work between launches) ----
    def kernel_ms(sc):
        sc.reset()
        barrier()
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0.record(stream)
        for _ in range(args.steps):
            sc.submit_device(db)
        k1.record(stream)
        torch.cuda.synchronize()
        r = sc.finish()
        return k0.elapsed_time(k1) / args.steps, r
    kern_ms, rk = kernel_ms(scan)
    probes_rank = rk.probes // args.steps      # table scans count the 32-byte probes they issue
    walked_rank, wsteps_rank = rk.walked // args.steps, rk.walk_steps // args.steps   # walking table scan (global mode)

    # ---- the same resident step WITHOUT the recording (tallies only: enough for the closed-form percentages, not for the refinement)
    tall_only = None
    if algo != capi.ALGO_DIRECT:      # every rank (kernel_ms holds a barrier); rank 0 reports its own shard
        pscan = engine.Scan(idx, k=k, phred_cutoff=cutoff, mode=mode, record=False, algorithm=algo_req, device=local_rank, stream=stream.cuda_stream)
        for _ in range(2):
            pscan.reset(); pscan.submit_device(db)
        torch.cuda.synchronize()
        p_ms, pr_ = kernel_ms(pscan)
        assert args.no_check or (np.array_equal(pr_.unique, rk.unique) and pr_.total_kmers == rk.total_kmers)
        tall_only = {"kernel_ms": p_ms, "value": R * wpr / (p_ms * 1e-3), "unit": UNIT, "record_over_tallies_only": p_ms / kern_ms,
                     "note": "rank 0's shard, resident batch, scan kernel without the per-k-mer multiplicity recording"}
        pscan.close()

    # ---- cross-check outside every timed region: the k-mer table scan must reproduce the rolling FM search bit for bit
    cross = None
    if table and not args.no_check:
        ref_scan = engine.Scan(idx, k=k, phred_cutoff=cutoff, mode=mode, algorithm=capi.ALGO_ROLLING, device=local_rank, stream=stream.cuda_stream)
        sub = min(R, 25_000_000)
        subv = capi.Batch(db.s.bases, db.s.inv, db.s.qual, db.s.grp_off, db.s.len, sub, int(db.grp_off[sub].item()))

        class _Sub:
            s = subv
        ref_scan.submit_device(_Sub)
        rr = ref_scan.finish()
        scan.reset(); scan.submit_device(_Sub); rm = scan.finish()
        same = np.array_equal(rr.unique, rm.unique) if not local else bool(np.all(np.abs(rr.unique_w - rm.unique_w) <= 1e-12 * np.abs(rr.unique_w)))
        assert same and rr.total_kmers == rm.total_kmers and rr.ambiguous == rm.ambiguous, "k-mer table tallies differ from the rolling FM search"
        cross = f"table-scan tallies == rolling FM-search tallies on the first {sub} reads of this rank (unique[V], total, ambiguous)"
        ref_scan.close()

    # ---- e2e: pinned host buffers through the C ABI ----
    e2e = e2e_roof = e2e_fastq = None
    if not args.no_e2e and E:
        for _ in range(min(args.warmup, 3)):
            tall_e, _ = step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            tall_e, _ = step_e2e()
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        nb = (E + args.batch_reads - 1) // args.batch_reads
        h2d = int(seq_np.nbytes + qual_np.nbytes + 8 * (E + nb))
        if E == R:
            assert np.array_equal(tall_e[:V + 2], tall[:V + 2]) or local, "e2e tallies differ from the resident path"
        e2e = {"value": int(tall_e[V + 2]) * args.steps / dt, "unit": UNIT, "h2d_bytes_per_step": h2d * world,
               "d2h_bytes_per_step": (8 * V + 8 * 8 + 8) * world, "ms_per_step": dt * 1e3 / args.steps,
               "reads_per_sec": int(tall_e[V + 3]) * args.steps / dt, "reads_per_step": int(tall_e[V + 3]),
               "input": f"pinned host ASCII reads + phred chars (the first {E} reads of every rank's shard), {args.batch_reads}-read submits; "
                        "device packer; record_multiplicity = 1",
               "launches_per_step": 2 * nb * world}
        # ---- e2e_fastq: the bytes of a FASTQ file from pinned host memory, parsed on the device
        for _ in range(min(args.warmup, 3)):
            tall_f, _ = step_fastq()
        assert scan.fastq_rejected() == []
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            tall_f, _ = step_fastq()
        barrier()
        dtf = time.perf_counter() - t0
        tt = torch.tensor([dtf], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dtf = float(tt.item())
        assert np.array_equal(tall_f[:V + 4], tall_e[:V + 4]) or local, "device-parsed FASTQ tallies differ from the raw-read path"
        e2e_fastq = {"value": int(tall_f[V + 2]) * args.steps / dtf, "unit": UNIT, "h2d_bytes_per_step": int(E) * FQ * world,
                     "d2h_bytes_per_step": (8 * V + 8 * 8 + 8) * world, "ms_per_step": dtf * 1e3 / args.steps,
                     "reads_per_sec": int(tall_f[V + 3]) * args.steps / dtf,
                     "input": f"pinned host FASTQ text ({FQ} B per record, the first {E} reads of every rank's shard), {args.batch_reads}-record "
                              "regions; line starts, record checks and packing on the device",
                     "launches_per_step": 7 * nb * world}
        # the ceiling of that leg: the same bytes in the same pieces as plain cudaMemcpyAsync from the same pinned buffers on all
        # ranks at once (no kernels) -- tells a box / link limit from a staging limit
        dbuf = torch.empty(2 * args.batch_reads * L, dtype=torch.uint8, device=dev)
        cs = torch.cuda.Stream(device=dev)

        def copy_pass():
            with torch.cuda.stream(cs):
                for s in range(0, E, args.batch_reads):
                    e = min(E, s + args.batch_reads)
                    nbts = (e - s) * L
                    dbuf[:nbts].copy_(seq_h.view(-1)[s * L:e * L], non_blocking=True)
                    dbuf[nbts:2 * nbts].copy_(qual_h.view(-1)[s * L:e * L], non_blocking=True)
            cs.synchronize()
        copy_pass()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            copy_pass()
        barrier()
        dtc = time.perf_counter() - t0
        tt = torch.tensor([dtc], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dtc = float(tt.item())
        gbs = 2.0 * E * L * world * args.steps / dtc / 1e9
        e2e_roof = {"h2d_GBps_aggregate": gbs, "h2d_GBps_per_gpu": gbs / world, "value_at_that_rate": E * wpr * world * args.steps / dtc, "unit": UNIT,
                    "frac": e2e["value"] / (E * wpr * world * args.steps / dtc),
                    "how": "plain cudaMemcpyAsync of the same pinned bytes in the same pieces, all ranks at once, no kernels"}
        del dbuf

    # ---- roofline ----
    roof = None
    if rank == 0:
        sb, keys = info["structure_bytes"], info["keys"]
        l2_bytes = torch.cuda.get_device_properties(local_rank).L2_cache_size
        # the kernel's own random sectors per launch.  A walk step reads two adjacent words of the packed texts and two of the
        # per-position flags: 1.25 sectors each on average (a pair of 8-byte words straddles a 32-byte sector one time in four)
        sectors = (probes_rank + (5 * wsteps_rank) // 2) if table else 2 * R * L
        own_bytes = 32 * sectors + packed_bytes
        ach = 32.0 * sectors / (kern_ms * 1e-3) / 1e9
        peak_sec = None
        if not args.no_probe:
            try:
                peak_sec = engine.probe_random_sectors(local_rank, max(sb, 1 << 20))
            except Exception as ex:
                log("probe failed:", ex)
        hbm_peak, hbm_how = measured_peak()
        kname = {capi.ALGO_ROLLING: "scan32_kernel", capi.ALGO_KTABLE: "scanh_kernel", capi.ALGO_KPAIR: "scanp_kernel"}.get(algo, "scan_direct_kernel")
        if wsteps_rank:
            kname = "scanw_kernel"
        kname += "<%s>" % ("|".join(x for x, on in (("M_LOCAL", local), ("M_RECORD", True)) if on) or "0")
        roof = {"bound": "l2_random_sector" if sb <= l2_bytes * 3 // 4 else "hbm_random_sector",
                "achieved": ach, "peak": None if peak_sec is None else 32.0 * peak_sec, "unit": "GB/s",
                "frac": None if peak_sec is None else ach / (32.0 * peak_sec), "traffic": None,
                "kernel": kname, "kernel_ms": kern_ms, "sectors_per_launch": sectors,
                "achieved_Gsectors_per_s": sectors / (kern_ms * 1e-3) / 1e9, "peak_Gsectors_per_s": peak_sec,
                "peak_source": "speqg_probe_random_sectors in this process: independent random 32-byte loads over a table of the probed "
                               f"structure's size ({sb/1e6:.0f} MB), best of 4 launch shapes",
                "structure_bytes": sb, "keys": keys, "l2_bytes": l2_bytes,
                "algorithmic_bytes_per_launch": own_bytes,
                "algorithmic_bytes_note": "32 B per table probe (KTABLE / KPAIR) or 2 x 32 B per read base (rolling) + the packed input stream",
                "own_bytes_GBps": own_bytes / (kern_ms * 1e-3) / 1e9}
        if wsteps_rank:
            # The walking scan answers most windows from the packed texts (32 per step), so it needs a fraction of the sectors
            # the probing scan needs and is no longer bound by them (issue slots: profiles/r2c_ncu.md).  Next to its own sectors:
            # what one sector per valid window -- the probing scan's roofline, rounds 1-2 -- would have to sustain to match it.
            roof.update({"table_probes_per_launch": probes_rank, "walk_steps_per_launch": wsteps_rank,
                         "windows_answered_by_the_walk": walked_rank, "windows_per_sector": R * wpr / max(1, sectors),
                         "algorithmic_bytes_note": "32 B per table probe + 2.5 x 32 B per walk step (<= 32 windows) + the packed input stream",
                         "probing_scan_equivalent": {"Gsectors_per_s": R * wpr / (kern_ms * 1e-3) / 1e9,
                                                     "frac_of_peak": None if peak_sec is None else R * wpr / (kern_ms * 1e-3) / 1e9 / peak_sec,
                                                     "note": "one 32-byte sector per window, as scanh_kernel issues them: > 1 means no probing "
                                                             "scan can reach this rate on this device"}})
        if sb > l2_bytes * 3 // 4:
            roof["hbm_frac"] = own_bytes / (kern_ms * 1e-3) / 1e9 / hbm_peak
            roof["hbm_peak"] = hbm_peak
            roof["hbm_peak_source"] = hbm_how
        # dram__bytes_read + write of this kernel from the committed `ncu --set full` capture, scaled from the captured launch to
        # this launch by the number of reads (per-read traffic does not depend on the launch size: input stream + table misses)
        try:
            for name in ("r2_traffic.json", "r1_traffic.json"):
                tp = ROOT / "profiles" / name
                if not tp.exists():
                    continue
                tj = json.loads(tp.read_text()).get(f"{args.config}:{kname}" + ("" if k == sh.k else f"@k{k}"))
                if tj:
                    roof["traffic"] = tj["dram_bytes_per_read"] * R
                    roof["traffic_source"] = tj["source"]
                    if "issue" in tj:      # what the same ncu capture says about instruction issue (the walking scan's real bound)
                        roof["issue_from_ncu"] = tj["issue"]
                    break
        except Exception:
            pass
        # SURVEY.md 8d's figure, kept under its own name: the reference's per-window search (k steps from the root + locate)
        # counted with reference semantics on a slice by scan_direct_kernel<1>, scaled to the launch
        try:
            sub = min(R, 2_000_000)
            subv = capi.Batch(db.s.bases, db.s.inv, db.s.qual, db.s.grp_off, db.s.len, sub, int(db.grp_off[sub].item()))

            class _Sub2:
                s = subv
            st = scan.stats(_Sub2)
            alg = (64 * (st["steps_executed"] - st["lookups"]) + 32 * st["occ_sectors"] + sub * L // 2) * (R / sub)
            roof["reference_work_rate"] = {"bytes_per_launch": alg, "GB/s": alg / (kern_ms * 1e-3) / 1e9,
                                           "note": "SURVEY.md 8d: 64 B per executed backward step after the first, 32 B per 32 occurrences "
                                                   "located, 0.5 B per base -- the work the reference does for the same windows; the table "
                                                   "kernel memoises it, so this is a rate of reference work retired, not a roofline fraction",
                                           "steps_per_lookup": st["steps_executed"] / max(1, st["lookups"]), "occ_per_hit": st["occ"] / max(1, st["kmers_hit"]),
                                           "counted_on_reads": sub}
        except Exception as ex:
            log("stats failed:", ex)

    # ---- cpu baseline + parity against it (rank 0, N = 1 only) ----
    cpu = parity = None
    if rank == 0 and world == 1 and not args.no_cpu and E:
        try:
            v, sample, kind, T, t, out, n = cpu_arm(sh, refs.cpu(), seq_h.numpy(), qual_h.numpy(), k=k, accuracy=accuracy, cutoff=cutoff,
                                                    steps=1, warmup=0, target_s=args.cpu_seconds)
            cpu = {"value": v, "unit": UNIT, "cores": T, "kind": kind, "sample": sample}
            # the oracle's tallies on its sample == the GPU's on the same n reads, default algorithm, through the raw-read C ABI
            pscan = engine.Scan(idx, k=k, phred_cutoff=cutoff, mode=mode, record=True, algorithm=algo_req, device=local_rank, stream=stream.cuda_stream)
            pscan.submit_raw(seq_np, qual_np, offs_np[:n + 1], n)
            g = pscan.finish()
            pscan.close()
            if local:
                okp = bool(np.all(np.abs(g.unique_w - np.array(out["unique_w"])) <= 1e-12 * np.abs(g.unique_w)))
            else:
                okp = [int(x) for x in g.unique] == out["unique"]
            okp = okp and (g.total_kmers, g.ambiguous, g.lookups) == (out["total_kmers"], out["ambiguous"], out["lookups"])
            parity = {"vs": "oracle/_ref" if kind == "reference" else "oracle (port)", "reads": n, "ok": bool(okp),
                      "compared": ("unique_w[V] (<= 1e-12 rel)" if local else "unique[V]") + ", total_kmers, ambiguous, lookups"}
            assert okp, ("GPU tallies differ from the oracle's on the CPU-timed sample", [int(x) for x in g.unique], out["unique"])
        except AssertionError:
            raise
        except Exception as ex:   # the checker must never take the bench line down
            cpu = {"value": None, "unit": UNIT, "cores": host_threads(), "kind": "port", "sample": f"failed: {ex}"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64" if local else "u64", "data": "synthetic", "config": config,
                "reads_per_sec": reads_step * args.steps / (ms_total * 1e-3),
                "value_is": "scan with record_multiplicity = 1 (what `speq scan` runs), resident packed reads",
                "resolved_algorithm": algo_name, "clocks": clk, "clocks_per_rank": clocks_all if world > 1 else None,
                "e2e": e2e, "e2e_fastq": e2e_fastq, "e2e_roofline": e2e_roof, "gpu_launches": launches * world, "roofline": roof, "cpu_baseline": cpu, "parity": parity,
                "tallies_only": tall_only,
                "tallies": {"unique": [int(x) for x in tall[:V]], "total_kmers": int(tall[V]), "ambiguous": int(tall[V + 1])},
                "setup": {"index_build_s": t_index, "index_n": idx.n, "index_device_mb": idx.device_bytes / 1e6, "generate_and_pack_s": t_gen,
                          "scan_open_s": t_scan_open, "reads_this_rank": R, "packed_bytes_this_rank": packed_bytes, "numa": numa},
                "cross_check": cross, "allreduce_check": allreduce_check}
        emit(line)
    scan.close()
    db.close()
    idx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
