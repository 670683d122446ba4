#!/usr/bin/env python
"""bench.py -- throughput of the SPeQ scan hot path on B200 (k-mer lookups/s; reads/s alongside).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--config C2] [--reads R]

A step = one scan pass over one batch of synthetic reads of the named BASELINE.json config (default C2: 10 variants x 20
isolates x 10 kb, 10 M x 150 bp reads, 0.5 % error, k = 21) per GPU.  N > 1: one rank per GPU (torchrun), every rank
scans its own 10 M-read shard against its replica of the index (weak scaling, no data-path collective); the only
collective is the all-reduce of the per-variant tallies at the end of the step.
  value     lookups/s with the packed reads already resident in HBM (CUDA events, max over ranks)
  e2e       lookups/s through the C ABI from pinned HOST buffers (ASCII reads + qualities): H2D copy, device packer,
            scan, D2H of the tallies all inside the timed region
  roofline  dominant kernel (scan_roll_kernel): ALGORITHMIC bytes of the reference search (SURVEY.md section 8d:
            64 B per executed backward step after the first, 32 B per 32 occurrences of a hit, 0.5 B per base of input)
            / measured kernel time, against MEASURED_PEAKS.json's HBM copy bandwidth
  cpu_baseline  the oracle (oracle/_ref = real sdsl-lite objects when present, else the port) on a bounded sample of the
            same reads, all host threads
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


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ---------------------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------------------
def workload(cfg: str, reads: int | None, rank: int, device):
    """-> (shape, refs codes [V*I, G] on device, seq ASCII [R, L] uint8 host-pinned, qual likewise)"""
    import torch
    from speq_b200 import synth
    sh = synth.CONFIGS[cfg]
    R = reads or sh.R
    refs = synth.make_refs(sh, synth.SEED0 + int(cfg[1]), device=device)
    pin = device != "cpu"
    seq = torch.empty((R, sh.L), dtype=torch.uint8, pin_memory=pin)
    qual = torch.empty((R, sh.L), dtype=torch.uint8, pin_memory=pin)
    chunk = max(1, min(R, (1 << 25) // sh.L))
    for i, s in enumerate(range(0, R, chunk)):
        e = min(R, s + chunk)
        a, b = synth.make_reads(sh, refs, e - s, synth.SEED0 * 1000 + int(cfg[1]) * 100000 + rank * 1000 + i)
        seq[s:e].copy_(a)
        qual[s:e].copy_(b)
        del a, b
    return sh, refs, seq, qual


def cpu_sample(sh, refs, seq, qual, n_reads: int, tmp: Path):
    from speq_b200 import synth
    synth.write_fasta(str(tmp / "refs.fa"), refs)
    synth.write_groups(str(tmp / "groups.txt"), sh)
    write_fastq_fixed(tmp / "sample.fq", seq[:n_reads].numpy(), qual[:n_reads].numpy())
    return tmp / "refs.fa", tmp / "groups.txt", tmp / "sample.fq"


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


def run_cpu(binary, files, sh, threads: int, n_reads: int) -> dict:
    """One scan pass of the oracle over the first n_reads reads (no back-check, no refinement): the reference's scan."""
    refs, groups, fq = files
    cmd = [str(binary), "run", "--refs", str(refs), "--groups", str(groups), "--reads1", str(fq), "-k", str(sh.k),
           "--fixed-accuracy", repr(sh.accuracy), "--threads", str(threads), "--max-reads", str(n_reads), "--no-refine", "--no-backcheck"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle failed: " + r.stderr)
    return json.loads(r.stdout)


def cpu_arm(sh, refs, seq, qual, *, steps: int, warmup: int, target_s: float, max_sample: int):
    """Times the CPU scan on a bounded sample.  Returns (lookups/s, sample description, kind, cores, per-step seconds, out)."""
    binary, kind = oracle_choice()
    T = host_threads()
    with tempfile.TemporaryDirectory() as d:
        n_cal = min(seq.shape[0], 256 * T)
        files = cpu_sample(sh, refs, seq, qual, n_cal, Path(d))
        cal = run_cpu(binary, files, sh, T, n_cal)
        rate = cal["lookups"] / max(cal["t_scan_s"], 1e-6)
        n = int(min(seq.shape[0], max_sample, max(n_cal, rate * target_s / max(1, sh.windows_per_read))))
        files = cpu_sample(sh, refs, seq, qual, n, Path(d))
        times, out = [], None
        for i in range(warmup + steps):
            out = run_cpu(binary, files, sh, T, n)
            if i >= warmup:
                times.append(out["t_scan_s"])
    t = sum(times) / len(times)
    return out["lookups"] / t, f"first {n} reads of the workload ({out['lookups']} lookups) per step, scan pass only", kind, T, t, out


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


# independent random 32-byte sector loads, G sectors/s, by table size (profiles/r1_gather_probe.txt, measured on this pool)
RANDOM_SECTOR_PEAKS = [
    (0.12e9, 235.8, "L2-resident table (0.1 GB probe)"),
    (0.4e9, 72.4, "0.3 GB table probe"),
    (0.8e9, 48.7, "0.5 GB table probe"),
    (1.6e9, 42.7, "1.1 GB table probe"),
    (3.2e9, 39.6, "2.1 GB table probe"),
    (float("inf"), 37.9, "4.3 GB table probe"),
]


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
    ap.add_argument("--config", default="C2", choices=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--reads", type=int, default=None, help="reads per GPU per step (default: the config's read count; C3: /8)")
    ap.add_argument("--batch-reads", type=int, default=1 << 20, help="reads per C-ABI submit on the e2e leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--algorithm", default="auto", choices=["auto", "kpair", "ktable", "rolling", "direct"],
                    help="auto = the library's default: pair / k-mer table scan for k <= 31, else the rolling FM search")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from speq_b200 import synth
    sh0 = synth.CONFIGS[args.config]
    reads = args.reads or (sh0.R // 8 if args.config == "C3" else sh0.R)
    config = {"workload": sh0.name, "config": args.config, "reads_per_gpu_per_step": reads, "read_len": sh0.L, "k": sh0.k,
              "variants": sh0.V, "isolates": sh0.I, "genome_len": sh0.G, "error_rate": sh0.e, "mode": "global (--fixed-accuracy %g)" % sh0.accuracy,
              "parallelism": f"reads sharded over {world} GPU(s), index replicated", "algorithm": args.algorithm}

    # ------------------------------------------------------------------------------------------------- reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        import torch
        sh, refs, seq, qual = workload(args.config, min(reads, 2_000_000), 0, "cpu")
        v, sample, kind, T, t, out = cpu_arm(sh, refs, seq, qual, steps=max(1, args.steps), warmup=min(args.warmup, 1),
                                             target_s=args.cpu_seconds, max_sample=seq.shape[0])
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
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

    t_gen = time.perf_counter()
    sh, refs, seq, qual = workload(args.config, reads, rank, dev)
    R, L = seq.shape
    t_gen = time.perf_counter() - t_gen
    t0 = time.perf_counter()
    cache = os.environ.get("SPEQ_BENCH_INDEX_CACHE")   # optional: reuse the (deterministic) synthetic index between runs
    cpath = Path(cache) / f"{args.config}.speqidx" if cache else None
    if cpath is not None and cpath.exists():
        idx = engine.Index.load(str(cpath))
    else:
        idx = engine.Index.from_scaffolds(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), sh.V, device_sa=True)   # suffix sort etc. on the GPU
        if cpath is not None and rank == 0:
            cpath.parent.mkdir(parents=True, exist_ok=True)
            idx.save(str(cpath))
    t_index = time.perf_counter() - t0
    idx.upload(local_rank)
    log(f"[rank {rank}] workload {R} reads generated in {t_gen:.1f}s; index n={idx.n} built in {t_index:.1f}s, {idx.device_bytes/1e6:.1f} MB on device")

    seq_f, qual_f = seq.view(-1), qual.view(-1)
    offs = (torch.arange(R + 1, dtype=torch.int64) * L).pin_memory()
    offs_np = offs.numpy().view(np.uint64)
    seq_np, qual_np = seq_f.numpy(), qual_f.numpy()

    # resident packed batch (host packer once, outside every timed region)
    t0 = time.perf_counter()
    pb = engine.pack_reads(seq_np, qual_np, offs_np, phred_cutoff=30, threads=min(host_threads(), 64))
    db = engine.DeviceBatch(local_rank, pb)
    packed_bytes = pb.nbytes(False)
    t_pack = time.perf_counter() - t0
    del pb

    stream = torch.cuda.Stream(device=dev)   # a real (non-default) stream: the scan handle launches on it and the events time it
    torch.cuda.set_stream(stream)
    algo_req = {"rolling": capi.ALGO_ROLLING, "direct": capi.ALGO_DIRECT, "ktable": capi.ALGO_KTABLE, "kpair": capi.ALGO_KPAIR,
                "auto": capi.ALGO_AUTO}[args.algorithm]
    t0 = time.perf_counter()
    scan = engine.Scan(idx, k=sh.k, phred_cutoff=30, mode=capi.MODE_GLOBAL, algorithm=algo_req, device=local_rank, stream=stream.cuda_stream)
    algo = scan.info()["algorithm"]           # what AUTO resolved to
    config["algorithm"] = {capi.ALGO_ROLLING: "rolling", capi.ALGO_DIRECT: "direct", capi.ALGO_KTABLE: "ktable",
                           capi.ALGO_KPAIR: "kpair"}[algo] + (" (auto)" if args.algorithm == "auto" else "")
    torch.cuda.synchronize()
    t_scan_open = time.perf_counter() - t0     # per-(index, k) specialisation: rank blocks or the k-mer table
    V = sh.V

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: the counters are summed across ranks where they live -- one NCCL all-reduce in place on the scan's stream behind
    # the last kernel -- so a step has a single host sync (finish), which then returns the global tallies
    cnt_i64 = scan.device_counters()[0] if world > 1 else None

    def finish_global():
        if world > 1:
            dist.all_reduce(cnt_i64)
        r = scan.finish()
        return np.concatenate([r.unique, np.array([r.total_kmers, r.ambiguous, r.lookups, r.reads], np.uint64)]).astype(np.int64)

    def step_resident():
        scan.reset()
        scan.submit_device(db)
        return finish_global()

    def step_e2e():
        scan.reset()
        for s in range(0, R, args.batch_reads):
            e = min(R, s + args.batch_reads)
            scan.submit_raw(seq_np, qual_np, offs_np[s:e + 1], e - s)
        return finish_global()

    # ---- value: resident inputs, CUDA events ----
    for _ in range(args.warmup):
        tall = step_resident()
    barrier()
    clocks = ClockSampler(local_rank)
    tc0 = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    launches = 0
    for _ in range(args.steps):
        tall = step_resident()
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
    value = lookups_step * args.steps / (ms_total * 1e-3)

    # ---- kernel-only time of the dominant kernel (same stream, no host work between launches) ----
    scan.reset()
    barrier()
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(stream)
    for _ in range(args.steps):
        scan.submit_device(db)
    k1.record(stream)
    torch.cuda.synchronize()
    kern_ms = k0.elapsed_time(k1) / args.steps
    scan.finish()

    # ---- the same resident step with record_multiplicity = 1 (what `speq scan` runs: per-k-mer counts for the refinement) ----
    rec = None
    if rank == 0 and algo != capi.ALGO_DIRECT:
        rscan = engine.Scan(idx, k=sh.k, phred_cutoff=30, mode=capi.MODE_GLOBAL, record=True, algorithm=algo_req, device=local_rank,
                            stream=stream.cuda_stream)
        for _ in range(2):
            rscan.reset(); rscan.submit_device(db)
        torch.cuda.synchronize()
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        rscan.reset()
        r0.record(stream)
        for _ in range(args.steps):
            rscan.submit_device(db)
        r1.record(stream)
        torch.cuda.synchronize()
        rr_ = rscan.finish()
        rec_ms = r0.elapsed_time(r1) / args.steps
        rec = {"kernel_ms": rec_ms, "value": rr_.lookups / args.steps / (rec_ms * 1e-3), "unit": UNIT,
               "note": "one GPU, resident batch, scan kernel with one 8-byte atomic per hit window on top"}
        rscan.close()

    # ---- cross-check outside every timed region: the k-mer table scan must reproduce the rolling FM search bit for bit
    cross = None
    if algo in (capi.ALGO_KTABLE, capi.ALGO_KPAIR):
        ref_scan = engine.Scan(idx, k=sh.k, phred_cutoff=30, mode=capi.MODE_GLOBAL, algorithm=capi.ALGO_ROLLING, device=local_rank,
                               stream=stream.cuda_stream)
        ref_scan.submit_device(db)
        rr = ref_scan.finish()
        scan.reset(); scan.submit_device(db); rm = scan.finish()
        assert np.array_equal(rr.unique, rm.unique) and rr.total_kmers == rm.total_kmers and rr.ambiguous == rm.ambiguous, \
            "k-mer table tallies differ from the rolling FM search"
        cross = "table-scan tallies == rolling FM-search tallies on this rank's reads (unique[V], total, ambiguous)"
        probes_rank = rm.probes
        ref_scan.close()

    # ---- e2e: pinned host buffers through the C ABI ----
    e2e = None
    if not args.no_e2e:
        for _ in range(min(args.warmup, 3)):
            tall_e = step_e2e()
        barrier()
        t0 = time.perf_counter()
        l0 = 0
        for _ in range(args.steps):
            tall_e = step_e2e()
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        assert np.array_equal(tall_e[:V + 2], tall[:V + 2]), "e2e tallies differ from the resident path"
        nb = (R + args.batch_reads - 1) // args.batch_reads
        h2d = int(seq_np.nbytes + qual_np.nbytes + 8 * (R + nb))
        e2e = {"value": int(tall_e[V + 2]) * args.steps / dt, "unit": UNIT, "h2d_bytes_per_step": h2d * world,
               "d2h_bytes_per_step": (8 * V + 8 * 8 + 8) * world, "ms_per_step": dt * 1e3 / args.steps,
               "reads_per_sec": int(tall_e[V + 3]) * args.steps / dt, "input": "pinned host ASCII reads + phred chars; device packer",
               "launches_per_step": 2 * nb * world}

    # ---- e2e_packed: what the streaming ingest hands over (2-bit packed batches in pinned host buffers, speqg_scan_submit_packed);
    #      the host packer runs outside the timed region, the H2D copy of the packed batch and the D2H of the tallies inside
    e2e_packed = None
    if not args.no_e2e:
        nbp = (R + args.batch_reads - 1) // args.batch_reads
        pbs = []
        for s0 in range(0, R, args.batch_reads):
            e0 = min(R, s0 + args.batch_reads)
            o = offs_np[s0:e0 + 1] - offs_np[s0]
            pbs.append(engine.pack_reads(seq_np[s0 * L:e0 * L], qual_np[s0 * L:e0 * L], o, phred_cutoff=30,
                                         threads=min(host_threads(), 64), pinned=True))

        def step_packed():
            scan.reset()
            for pb_ in pbs:
                scan.submit_packed(pb_)
            return finish_global()
        for _ in range(min(args.warmup, 3)):
            tall_p = step_packed()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            tall_p = step_packed()
        barrier()
        dtp = time.perf_counter() - t0
        tt = torch.tensor([dtp], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dtp = float(tt.item())
        assert np.array_equal(tall_p[:V + 2], tall[:V + 2]), "packed-host tallies differ from the resident path"
        e2e_packed = {"value": int(tall_p[V + 2]) * args.steps / dtp, "unit": UNIT, "h2d_bytes_per_step": sum(b_.nbytes(False) for b_ in pbs) * world,
                      "d2h_bytes_per_step": (8 * V + 8 * 8 + 8) * world, "ms_per_step": dtp * 1e3 / args.steps,
                      "input": "pinned host 2-bit packed batches (host packer outside the timed region)", "launches_per_step": nbp * world}
        del pbs

    # ---- roofline: reference-semantics work counters (direct kernel, outside every timed region) ----
    roof = None
    if rank == 0:
        st = scan.stats(db)
        input_bytes = R * L // 2   # 2 bit base + 1 bit N + 1 bit q-pass per base
        alg = 64 * (st["steps_executed"] - st["lookups"]) + 32 * st["occ_sectors"] + input_bytes
        peak, how = measured_peak()
        ach = alg / (kern_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                "kernel": {capi.ALGO_ROLLING: "scan32_kernel<0>", capi.ALGO_KTABLE: "scanh_kernel<0>",
                           capi.ALGO_KPAIR: "scanp_kernel<0>"}.get(algo, "scan_direct_kernel<0>"),
                "kernel_ms": kern_ms,
                "algorithmic_bytes_per_launch": alg, "peak_source": how,
                "work": {"steps_executed": st["steps_executed"], "lookups": st["lookups"], "kmers_hit": st["kmers_hit"],
                         "occ": st["occ"], "occ_sectors": st["occ_sectors"], "input_bytes": input_bytes},
                "note": "algorithmic bytes follow the reference's per-window search (k steps from the root); the rolling kernel "
                        "shares search state between overlapping windows and the k-mer table kernel memoises the search, so their "
                        f"real traffic is lower (own_sectors); index footprint {idx.device_bytes/1e6:.0f} MB"}
        # dram__bytes_read + write of this kernel from the committed `ncu --set full` capture, scaled from the captured launch to
        # this launch by the number of reads (per-read traffic does not depend on the launch size: input stream + table misses)
        try:
            tj = json.loads((ROOT / "profiles" / "r1_traffic.json").read_text()).get(f"{args.config}:{roof['kernel']}")
            if tj:
                roof["traffic"] = tj["dram_bytes_per_read"] * R
                roof["traffic_source"] = tj["source"]
        except Exception:
            pass
        # what the kernel itself must move: one 32-byte sector per valid window (ktable) / two per read base (rolling), plus
        # the packed input stream; against the HBM peak when the structure misses L2, else informative only
        table = algo in (capi.ALGO_KTABLE, capi.ALGO_KPAIR)
        own = (32 * probes_rank if table else 64 * R * L) + packed_bytes
        roof["own_sectors"] = {"bytes_per_launch": own, "GB/s": own / (kern_ms * 1e-3) / 1e9, "frac_of_hbm_peak": own / (kern_ms * 1e-3) / 1e9 / peak}
        # the roofline this access pattern really has: independent random 32-byte sectors.  Peaks measured on this pool's B200
        # by profiles/gather_probe.cu (profiles/r1_gather_probe.txt), chosen by the size of the structure the kernel probes.
        info = scan.info()
        sb = info["structure_bytes"]
        ra_peak, ra_src = next((p_, s_) for lim, p_, s_ in RANDOM_SECTOR_PEAKS if sb <= lim)
        sectors = probes_rank if table else 2 * R * L     # table scans count the probes they issue
        roof["random_access"] = {"sectors_per_launch": sectors, "achieved_Gsectors_per_s": sectors / (kern_ms * 1e-3) / 1e9,
                                 "peak_Gsectors_per_s": ra_peak, "frac": sectors / (kern_ms * 1e-3) / 1e9 / ra_peak,
                                 "structure_bytes": sb, "keys": info["keys"], "peak_source": ra_src}

    # ---- cpu baseline (rank 0, N = 1 only) ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            v, sample, kind, T, t, _ = cpu_arm(sh, refs.cpu(), seq, qual, steps=1, warmup=0, target_s=args.cpu_seconds,
                                               max_sample=min(R, 2_000_000))
            cpu = {"value": v, "unit": UNIT, "cores": T, "kind": kind, "sample": sample}
        except Exception as ex:   # the checker must never take the bench line down
            cpu = {"value": None, "unit": UNIT, "cores": host_threads(), "kind": "port", "sample": f"failed: {ex}"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u64", "data": "synthetic", "config": dict(config, l2="inputs larger than L2: %d MB of packed reads per step" % (packed_bytes >> 20)),
                "reads_per_sec": reads_step * args.steps / (ms_total * 1e-3),
                "clocks": clk, "e2e": e2e, "e2e_packed": e2e_packed, "gpu_launches": launches * world, "roofline": roof, "cpu_baseline": cpu, "record_mode": rec,
                "tallies": {"unique": [int(x) for x in tall[:V]], "total_kmers": int(tall[V]), "ambiguous": int(tall[V + 1])},
                "setup": {"index_build_s": t_index, "index_n": idx.n, "index_device_mb": idx.device_bytes / 1e6, "host_pack_s": t_pack, "scan_open_s": t_scan_open, "numa": numa},
                "cross_check": cross}
        emit(line)
    scan.close()
    db.close()
    idx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
