"""`speq all` on in-memory inputs: index -> back-check -> scan -> closed-form solve -> refinement loop.

Python mirror of the reference's driver functions for the hot path (paths relative to the reference checkout):
  speq::fm::index                               src/fm_indexer.cpp:55-110
  speq::scan::async_one / async_two             src/fm_scanner.cpp:5-33   (mode selection, percent_perfect = pow(acc, k))
  _async_{one,two}_with_{global,local}_error_rate  src/fm_scanner.cpp:35-279, 310-545, 548-792, 795-1065
Every number is produced by libspeq_cuda.so (CUDA); this file sequences the calls, shards read batches over ranks
and all-reduces the small tallies with torch.distributed (NCCL on the GPU box, gloo in the CPU tests).
The C++ `speq` binary (speq_b200/host) does the same over the same C ABI.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from . import capi, engine


@dataclass
class Reads:
    """Flat read set: read i = seq[offsets[i]:offsets[i+1]] (ASCII letters / phred+33 chars).  Paired input is
    interleaved: reads 2i and 2i+1 are the mates of pair i (what views::zip(fin1, fin2) yields, src/fm_scanner.cpp:651)."""
    seq: np.ndarray
    qual: np.ndarray | None
    offsets: np.ndarray

    @property
    def n(self) -> int:
        return self.offsets.size - 1


def shard_bounds(n_units: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous shard [lo, hi) of n_units (reads or pairs) for `rank` of `world`."""
    return n_units * rank // world, n_units * (rank + 1) // world


def batch_ranges(lo: int, hi: int, batch: int):
    s = lo
    while s < hi:
        e = min(hi, s + batch)
        yield s, e
        s = e


class Reducer:
    """Sum-all-reduce of the small tally vectors across ranks.  world == 1: identity."""

    def __init__(self, group=None, device=None):
        self.group = group
        self.device = device
        self.world = 1
        if group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(group)

    def sum_u64(self, a: np.ndarray) -> np.ndarray:
        if self.world == 1:
            return a
        import torch
        import torch.distributed as dist
        t = torch.from_numpy(a.astype(np.int64)).to(self.device or "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t.cpu().numpy().astype(np.uint64)

    def sum_f64(self, a: np.ndarray) -> np.ndarray:
        if self.world == 1:
            return a
        import torch
        import torch.distributed as dist
        t = torch.from_numpy(np.array(a, dtype=np.float64, copy=True)).to(self.device or "cpu")   # never alias the caller's array
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t.cpu().numpy()


def percent_perfect(fixed_accuracy: float, k: int) -> float:
    """pow(fixed_accuracy, k) with the C library's pow, as src/fm_scanner.cpp:15 computes it."""
    import ctypes
    import ctypes.util
    libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.pow.restype = ctypes.c_double
    libm.pow.argtypes = [ctypes.c_double, ctypes.c_double]
    return float(libm.pow(float(fixed_accuracy), float(k)))


def scan_reads(scan: engine.Scan, reads: Reads, lo: int, hi: int, *, paired: bool, batch_reads: int = 1 << 20,
               raw: bool = True, phred_cutoff: int = 30, local: bool = False) -> None:
    """Submit reads [lo, hi) (pairs when paired) in batches.  raw: ship ASCII bytes and pack on the device;
    else pack on the host (speqg_pack_reads) and ship the packed batch."""
    mul = 2 if paired else 1
    for s, e in batch_ranges(lo, hi, max(1, batch_reads // mul)):
        a, b = s * mul, e * mul
        offs = reads.offsets[a:b + 1]
        if raw:
            scan.submit_raw(reads.seq, reads.qual, np.ascontiguousarray(offs), b - a)
        else:
            base = int(offs[0])
            pb = engine.pack_reads(reads.seq[base:int(offs[-1])], None if reads.qual is None else reads.qual[base:int(offs[-1])],
                                   offs - offs[0], phred_cutoff=phred_cutoff, with_qual=local)
            scan.submit_packed(pb)
            scan.sync()   # pb's arrays are about to go out of scope


def run_all(scaffolds: list[np.ndarray], group_scaffolds, group_counts, reads: Reads, *, k: int, phred_cutoff: int = 30,
            fixed_accuracy: float = 0.0, precision: float = 1e-6, paired: bool = False, device: int = 0,
            group=None, reduce_device=None, algorithm: int = capi.ALGO_AUTO, raw: bool = True,
            batch_reads: int = 1 << 20, refine: bool = True, max_iters: int = 100000, index: engine.Index | None = None) -> dict:
    """scaffolds: dna5-rank arrays in FASTA order.  Returns the same keys the oracle's `run` mode prints."""
    V = len(group_counts)
    red = Reducer(group, reduce_device)
    rank = 0
    if group is not None:
        import torch.distributed as dist
        rank = dist.get_rank(group)
    idx = index or engine.Index.from_scaffolds(scaffolds, group_scaffolds, V)
    # back-check (what the reference caches as <stem>_<k>mer.dat) -- every rank computes it: it is tiny next to the scan
    unique_ref, total_ref = engine.backcheck(idx, k, device)
    local = (fixed_accuracy == 0.0)   # src/fm_scanner.cpp:8-17
    scan = engine.Scan(idx, k=k, phred_cutoff=phred_cutoff, mode=capi.MODE_LOCAL if local else capi.MODE_GLOBAL,
                       paired=paired, record=refine, algorithm=algorithm, device=device)
    try:
        units = reads.n // 2 if paired else reads.n
        lo, hi = shard_bounds(units, rank, red.world)
        scan_reads(scan, reads, lo, hi, paired=paired, batch_reads=batch_reads, raw=raw, phred_cutoff=phred_cutoff, local=local)
        r = scan.finish()
        ints = red.sum_u64(np.concatenate([r.unique, np.array([r.total_kmers, r.ambiguous, r.lookups, r.reads], np.uint64)]))
        unique_int, (total_kmers, ambiguous, lookups, n_reads) = ints[:V], (int(x) for x in ints[V:])
        unique_w = red.sum_f64(r.unique_w)
        if local:
            unique_totals = unique_w.copy()
        else:
            pp = percent_perfect(fixed_accuracy, k)
            unique_totals = unique_int.astype(np.float64) / pp          # src/fm_scanner.cpp:234-236
        gc = np.ascontiguousarray(group_counts, np.int32)

        def estimate(percent):
            return red.sum_f64(scan.refine(percent, gc))

        if refine:
            initial, final, iters, hist = engine.solve(unique_totals, total_kmers, unique_ref, total_ref, gc, estimate, precision, max_iters)
        else:
            initial = engine.unique_to_percent(unique_totals, total_kmers, unique_ref.astype(np.float64), total_ref.astype(np.float64))
            final, iters, hist = initial, 0, []
        return {"n": idx.n, "V": V, "k": k, "mode": "local" if local else "global", "paired": paired, "reads": n_reads,
                "unique_ref": unique_ref, "total_ref": total_ref, "unique": unique_int, "unique_w": unique_w,
                "unique_totals": unique_totals, "total_kmers": total_kmers, "ambiguous": ambiguous, "lookups": lookups,
                "percent_initial": initial, "percent_final": final, "iterations": iters,
                "iter_percent": [h[0] for h in hist], "iter_tkpg": [h[1] for h in hist],
                "kernel_launches": r.kernel_launches}
    finally:
        scan.close()
        if index is None:
            idx.close()
