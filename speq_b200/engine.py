"""Thin object layer over the C ABI (capi.py) used by the tests, bench.py and the Python front end.

Mirrors the reference's call structure for the hot path (paths relative to the reference checkout):
  Index.from_scaffolds   <- speq::fm::generate_fm_index            src/fm_indexer.cpp:8-52
  backcheck              <- _async_count_unique_kmers_per_group    src/fm_scanner.cpp:1476-1576
  Scan.submit_* / finish <- worker bodies + future reduction       src/fm_scanner.cpp:145-236
  solve                  <- unique_to_percent + refinement loop    src/fm_scanner.cpp:238-279
All arithmetic happens in libspeq_cuda.so; this file only moves pointers.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

from . import capi

_COMPLEMENT = np.array([4, 2, 1, 3, 0], dtype=np.uint8)   # dna5 ranks A0 C1 G2 N3 T4


def revcomp_ranks(r: np.ndarray) -> np.ndarray:
    """complement(reverse(seq)) on dna5 ranks (src/fm_indexer.cpp:30-33)."""
    return _COMPLEMENT[r[::-1]]


_ASCII_RANK = np.full(256, 3, dtype=np.uint8)
for _ch, _r in (("A", 0), ("C", 1), ("G", 2), ("T", 4), ("U", 4)):
    _ASCII_RANK[ord(_ch)] = _r
    _ASCII_RANK[ord(_ch.lower())] = _r


def ascii_to_ranks(s: bytes | str | np.ndarray) -> np.ndarray:
    if isinstance(s, str):
        s = s.encode()
    a = np.frombuffer(s, dtype=np.uint8) if isinstance(s, (bytes, bytearray)) else s
    return _ASCII_RANK[a]


class Index:
    """FM index over {scaffold_0, revcomp(scaffold_0), scaffold_1, ...} (seqan3 collection layout)."""

    def __init__(self, handle: int):
        self.h = C.c_void_p(handle)
        info = capi.IndexInfo()
        capi.check(capi.lib().speqg_index_get_info(self.h, C.byref(info)))
        self.n = int(info.n)
        self.n_texts = int(info.n_texts)
        self.n_groups = int(info.n_groups)
        self.C = [int(x) for x in info.C]
        self.n_exceptions = int(info.n_exceptions)
        self.device_bytes = int(info.device_bytes)

    @classmethod
    def from_texts(cls, texts: list[np.ndarray], group_of_text, n_groups: int, keep_sa: bool = False, device_sa: bool = False) -> "Index":
        S = len(texts)
        arrs = [np.ascontiguousarray(t, dtype=np.uint8) for t in texts]
        ptrs = (C.c_void_p * max(S, 1))(*[a.ctypes.data if a.size else None for a in arrs])
        lens = np.array([a.size for a in arrs], dtype=np.uint64)
        grp = np.ascontiguousarray(group_of_text, dtype=np.int32)
        assert grp.size == S
        out = C.c_void_p()
        capi.check(capi.lib().speqg_index_build(ptrs, lens.ctypes.data_as(capi.u64p), S, grp.ctypes.data_as(capi.i32p),
                                                n_groups, (capi.BUILD_KEEP_SA if keep_sa else 0) | (capi.BUILD_DEVICE_SA if device_sa else 0),
                                                C.byref(out)))
        return cls(out.value)

    @classmethod
    def from_scaffolds(cls, scaffolds: list[np.ndarray], group_scaffolds, n_groups: int, keep_sa: bool = False,
                       device_sa: bool = False) -> "Index":
        """scaffolds: dna5-rank arrays in FASTA order; group_scaffolds[j]: variant of scaffold j (file_to_map)."""
        texts, grp = [], []
        for j, s in enumerate(scaffolds):
            s = np.ascontiguousarray(s, dtype=np.uint8)
            texts.append(s)
            texts.append(revcomp_ranks(s))
            grp += [int(group_scaffolds[j])] * 2   # double_group_scaffolds, src/fm_scanner.cpp:74-77
        return cls.from_texts(texts, grp, n_groups, keep_sa, device_sa)

    @classmethod
    def load(cls, path: str) -> "Index":
        out = C.c_void_p()
        capi.check(capi.lib().speqg_index_load(os.fsencode(path), C.byref(out)))
        return cls(out.value)

    def save(self, path: str) -> None:
        capi.check(capi.lib().speqg_index_save(self.h, os.fsencode(path)))

    def export(self, sa: bool = False):
        n = self.n
        bwt = np.empty(n, np.uint8)
        grp = np.empty(n, np.uint32)
        lcp = np.empty(n, np.uint8)
        sa_a = np.empty(n, np.uint32) if sa else None
        capi.check(capi.lib().speqg_index_export(self.h, capi.ptr(sa_a), bwt.ctypes.data, grp.ctypes.data, lcp.ctypes.data))
        return {"sa": sa_a, "bwt": bwt, "group_of_sa": grp, "lcp": lcp}

    def upload(self, device: int = 0) -> None:
        capi.check(capi.lib().speqg_index_upload(self.h, device))

    def close(self) -> None:
        if self.h:
            capi.lib().speqg_index_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ReferenceIndexFile:
    """An .idx written by the reference's `speq index` (cereal archive + sdsl csa_wt): metadata and the indexed texts, decoded
    from the wavelet tree (speqg_refidx_*)."""

    def __init__(self, path: str):
        self.h = C.c_void_p()
        capi.check(capi.lib().speqg_refidx_open(os.fsencode(path), C.byref(self.h)))
        l = capi.lib()
        self.reference_path = l.speqg_refidx_reference_path(self.h).decode()
        self.file_times = (int(l.speqg_refidx_file_time(self.h, 0)), int(l.speqg_refidx_file_time(self.h, 1)))
        self.n = int(l.speqg_refidx_size(self.h))
        V = int(l.speqg_refidx_n_groups(self.h))
        self.group_names = [l.speqg_refidx_group_name(self.h, g).decode() for g in range(V)]
        self.group_counts = [int(l.speqg_refidx_group_count(self.h, g)) for g in range(V)]
        self.group_scaffolds = [int(l.speqg_refidx_group_of_scaffold(self.h, j)) for j in range(int(l.speqg_refidx_n_scaffolds(self.h)))]

    def texts(self) -> list[np.ndarray]:
        out = []
        for i in range(int(capi.lib().speqg_refidx_n_texts(self.h))):
            p, n = C.c_void_p(), C.c_uint64()
            capi.check(capi.lib().speqg_refidx_text(self.h, i, C.byref(p), C.byref(n)))
            out.append(np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,)).copy() if n.value else np.zeros(0, np.uint8))
        return out

    def build_index(self, device_sa: bool = False) -> Index:
        out = C.c_void_p()
        capi.check(capi.lib().speqg_refidx_build_index(self.h, capi.BUILD_DEVICE_SA if device_sa else 0, C.byref(out)))
        return Index(out.value)

    def close(self) -> None:
        if self.h:
            capi.lib().speqg_refidx_close(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


@dataclass
class PackedBatch:
    """Host-side packed batch (numpy arrays, or pinned torch tensors via from_arrays)."""
    bases: object
    inv: object
    qual: object
    grp_off: object
    len: object
    n_reads: int
    n_groups32: int
    keep: list = field(default_factory=list)

    def struct(self) -> capi.Batch:
        return capi.Batch(capi.ptr(self.bases), capi.ptr(self.inv), capi.ptr(self.qual), capi.ptr(self.grp_off),
                          capi.ptr(self.len), self.n_reads, self.n_groups32)

    def nbytes(self, with_qual: bool) -> int:
        return self.n_groups32 * 12 + (self.n_reads + 1) * 8 + self.n_reads * 4 + (self.n_groups32 * 32 if with_qual else 0)


def pack_reads(seq: np.ndarray, qual: np.ndarray | None, offsets: np.ndarray, *, encoding: int = capi.ENC_ASCII,
               phred_cutoff: int = 30, n_only: bool = False, with_qual: bool = False, threads: int = 0,
               pinned: bool = False) -> PackedBatch:
    """Host packer (speqg_pack_reads).  seq/qual: flat uint8 arrays; offsets: uint64[n_reads+1]."""
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    R = offsets.size - 1
    G = int(capi.lib().speqg_pack_groups(offsets.ctypes.data, R))
    keep = []
    if pinned:
        import torch

        def alloc(n, dt):
            t = torch.empty(max(n, 1), dtype=dt, pin_memory=True)
            keep.append(t)
            return t.numpy()
        bases = alloc(G, torch.int64).view(np.uint64)
        inv = alloc(G, torch.int32).view(np.uint32)
        grp_off = alloc(R + 1, torch.int64).view(np.uint64)
        ln = alloc(R, torch.int32).view(np.uint32)
        q = alloc(G * 32, torch.uint8) if with_qual else None
    else:
        bases = np.zeros(max(G, 1), np.uint64)
        inv = np.zeros(max(G, 1), np.uint32)
        grp_off = np.zeros(R + 1, np.uint64)
        ln = np.zeros(max(R, 1), np.uint32)
        q = np.zeros(max(G * 32, 1), np.uint8) if with_qual else None
    b = PackedBatch(bases, inv, q, grp_off, ln, R, G, keep)
    st = b.struct()
    if threads <= 0:
        threads = min(os.cpu_count() or 1, 32)
    capi.check(capi.lib().speqg_pack_reads(capi.ptr(np.ascontiguousarray(seq)) if R else None,
                                           capi.ptr(np.ascontiguousarray(qual)) if qual is not None else None,
                                           offsets.ctypes.data, R, encoding, phred_cutoff, int(n_only), threads, C.byref(st)))
    return b


class DeviceBatch:
    def __init__(self, device: int, host: PackedBatch):
        self.device = device
        self.s = capi.Batch()
        hs = host.struct()
        capi.check(capi.lib().speqg_batch_to_device(device, C.byref(hs), C.byref(self.s)))
        self.n_reads = host.n_reads

    def close(self):
        if self.s.bases:
            capi.lib().speqg_batch_free_device(self.device, C.byref(self.s))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ResidentBatch:
    """A packed batch in device memory whose arrays are torch tensors, filled range by range by the device packer
    (speqg_pack_device) from raw reads that already live on the device -- how bench.py builds a resident batch of 2e8 reads
    without staging them through the host.  grp_off (first 32-base group of every read) is the running sum of ceil(len / 32)."""

    def __init__(self, device: int, lens, with_qual: bool = False):
        import torch
        dev = torch.device("cuda", device)
        self.device = device
        lens = torch.as_tensor(lens, device=dev, dtype=torch.int64)
        self.n_reads = int(lens.numel())
        self.grp_off = torch.zeros(self.n_reads + 1, dtype=torch.int64, device=dev)
        torch.cumsum((lens + 31) >> 5, 0, out=self.grp_off[1:])
        self.n_groups32 = int(self.grp_off[-1].item())
        self.bases = torch.zeros(max(1, self.n_groups32), dtype=torch.int64, device=dev)
        self.inv = torch.zeros(max(1, self.n_groups32), dtype=torch.int32, device=dev)
        self.len = torch.zeros(max(1, self.n_reads), dtype=torch.int32, device=dev)
        self.qual = torch.zeros(max(1, self.n_groups32 * 32), dtype=torch.uint8, device=dev) if with_qual else None
        self.s = capi.Batch(self.bases.data_ptr(), self.inv.data_ptr(), self.qual.data_ptr() if with_qual else None,
                            self.grp_off.data_ptr(), self.len.data_ptr(), self.n_reads, self.n_groups32)

    def nbytes(self) -> int:
        return self.n_groups32 * 12 + (self.n_reads + 1) * 8 + self.n_reads * 4 + (self.n_groups32 * 32 if self.qual is not None else 0)

    def pack(self, r0: int, seq, qual, offsets, *, encoding: int = capi.ENC_ASCII, phred_cutoff: int = 30, n_only: bool = False) -> None:
        """reads [r0, r0 + n) from device tensors: seq / qual flat uint8, offsets int64[n + 1] relative to seq."""
        import torch
        n = int(offsets.numel()) - 1
        assert 0 <= r0 and r0 + n <= self.n_reads
        st = capi.Batch(self.bases.data_ptr(), self.inv.data_ptr(), self.qual.data_ptr() if self.qual is not None else None,
                        self.grp_off.data_ptr() + 8 * r0, self.len.data_ptr() + 4 * r0, n, self.n_groups32)
        capi.check(capi.lib().speqg_pack_device(self.device, seq.data_ptr(), qual.data_ptr() if qual is not None else None,
                                                offsets.data_ptr(), n, encoding, phred_cutoff, int(n_only), C.byref(st),
                                                torch.cuda.current_stream(self.device).cuda_stream))

    def close(self) -> None:
        self.bases = self.inv = self.len = self.qual = self.grp_off = None


def probe_random_sectors(device: int, table_bytes: int) -> float:
    """Giga independent random 32-byte sector loads per second from a table of table_bytes (speqg_probe_random_sectors)."""
    out = C.c_double()
    capi.check(capi.lib().speqg_probe_random_sectors(device, int(table_bytes), C.byref(out)))
    return float(out.value)


@dataclass
class ScanResult:
    unique: np.ndarray        # u64[V]   (global mode tallies; back-check: unique_ref)
    unique_w: np.ndarray      # f64[V]   (local mode)
    total_kmers: int
    ambiguous: int
    lookups: int
    reads: int
    kernel_launches: int
    probes: int = 0           # table algorithms: 32-byte table probes issued
    walked: int = 0           # walking table scan: windows answered from the packed texts, without a probe
    walk_steps: int = 0       # ... and the steps (<= 32 windows each) that answered them


class Scan:
    def __init__(self, index: Index, *, k: int, phred_cutoff: int = 30, mode: int = capi.MODE_GLOBAL, paired: bool = False,
                 record: bool = False, algorithm: int = capi.ALGO_AUTO, device: int = 0, stream: int | None = None):
        self.index = index
        self.V = index.n_groups
        self.mode = mode
        self.device = device
        self.params = capi.Params(k, phred_cutoff, mode, int(paired), int(record), algorithm, stream)
        self.h = C.c_void_p()
        capi.check(capi.lib().speqg_scan_open(index.h, device, C.byref(self.params), C.byref(self.h)))

    def submit_packed(self, b: PackedBatch) -> None:
        st = b.struct()
        capi.check(capi.lib().speqg_scan_submit_packed(self.h, C.byref(st)))

    def submit_raw(self, seq, qual, offsets, n_reads: int, encoding: int = capi.ENC_ASCII) -> None:
        capi.check(capi.lib().speqg_scan_submit_raw(self.h, capi.ptr(seq), capi.ptr(qual), capi.ptr(offsets), n_reads, encoding))

    def submit_fastq(self, raw, n_bytes: int | None = None) -> tuple[int, int]:
        """raw: bytes / numpy uint8 array / pinned torch tensor holding whole 4-line FASTQ records.  -> (region, ticket)"""
        if isinstance(raw, (bytes, bytearray)):
            raw = np.frombuffer(raw, np.uint8)
        n = int(n_bytes if n_bytes is not None else raw.size if isinstance(raw, np.ndarray) else raw.numel())
        reg, tick = C.c_uint64(), C.c_uint64()
        capi.check(capi.lib().speqg_scan_submit_fastq(self.h, capi.ptr(raw) if n else None, n, C.byref(reg), C.byref(tick)))
        self._keep = raw          # pageable host memory: the copy is staged before the call returns; pinned: caller waits the ticket
        return int(reg.value), int(tick.value)

    def submit_fastq_pair(self, raw1, raw2) -> tuple[int, int]:
        """Two raw FASTQ regions with the same number of records: first mates, second mates.  -> (region, ticket)"""
        a = np.frombuffer(raw1, np.uint8) if isinstance(raw1, (bytes, bytearray)) else raw1
        b = np.frombuffer(raw2, np.uint8) if isinstance(raw2, (bytes, bytearray)) else raw2
        reg, tick = C.c_uint64(), C.c_uint64()
        capi.check(capi.lib().speqg_scan_submit_fastq_pair(self.h, capi.ptr(a) if a.size else None, a.size, capi.ptr(b) if b.size else None, b.size,
                                                           C.byref(reg), C.byref(tick)))
        self._keep = (a, b)
        return int(reg.value), int(tick.value)

    def flush(self) -> None:
        """Everything submitted so far is queued on the scan's stream (needed before an in-place all-reduce of device_counters)."""
        capi.check(capi.lib().speqg_scan_flush(self.h))

    def wait_ticket(self, ticket: int) -> None:
        capi.check(capi.lib().speqg_scan_wait_ticket(self.h, ticket))

    def fastq_rejected(self) -> list[int]:
        n = C.c_uint64()
        capi.check(capi.lib().speqg_scan_fastq_rejected(self.h, None, 0, C.byref(n)))
        out = np.zeros(max(1, n.value), np.uint64)
        capi.check(capi.lib().speqg_scan_fastq_rejected(self.h, out.ctypes.data, n.value, C.byref(n)))
        return [int(x) for x in out[:n.value]]

    def submit_device(self, b: DeviceBatch) -> None:
        capi.check(capi.lib().speqg_scan_submit_device(self.h, C.byref(b.s)))

    def sync(self) -> None:
        capi.check(capi.lib().speqg_scan_sync(self.h))

    def reset(self) -> None:
        capi.check(capi.lib().speqg_scan_reset(self.h))

    def finish(self) -> ScanResult:
        u = np.zeros(self.V, np.uint64)
        w = np.zeros(self.V, np.float64)
        c = capi.Counts()
        capi.check(capi.lib().speqg_scan_finish(self.h, u.ctypes.data, w.ctypes.data, C.byref(c)))
        return ScanResult(u, w, int(c.total_kmers), int(c.ambiguous_reads), int(c.lookups), int(c.reads), int(c.kernel_launches), int(c.steps_executed),
                          int(c.kmers_hit), int(c.occ_sectors))

    def stats(self, b: DeviceBatch) -> dict:
        c = capi.Counts()
        capi.check(capi.lib().speqg_scan_stats(self.h, C.byref(b.s), C.byref(c)))
        return {"lookups": int(c.lookups), "reads": int(c.reads), "steps_executed": int(c.steps_executed),
                "kmers_hit": int(c.kmers_hit), "occ": int(c.occurrences), "occ_sectors": int(c.occ_sectors)}

    def info(self) -> dict:
        """Resolved algorithm (AUTO -> KTABLE / ROLLING) and the size of the per-(index, k) structure it scans."""
        a, nb, nk = C.c_uint32(), C.c_uint64(), C.c_uint64()
        capi.check(capi.lib().speqg_scan_info(self.h, C.byref(a), C.byref(nb), C.byref(nk)))
        return {"algorithm": int(a.value), "structure_bytes": int(nb.value), "keys": int(nk.value)}

    def fusion_counts(self) -> np.ndarray:
        """u64[V, V]: ambiguous reads / pairs by (smallest, largest) unique variant; sums to ScanResult.ambiguous."""
        out = np.zeros((self.V, self.V), np.uint64)
        capi.check(capi.lib().speqg_scan_fusion_counts(self.h, out.ctypes.data))
        return out

    def fusion_sets(self) -> dict:
        """{tuple of variants: ambiguous reads / pairs}: the fix-forward of the reference's fusion_totals map
        (src/fm_scanner.cpp:915,947-952,983-995).  At most 64 variants: the full set of variants with a unique k-mer in the unit;
        above: (smallest, largest) with a trailing "+" entry when further variants lie between them."""
        n, lost = C.c_uint64(), C.c_uint64()
        capi.check(capi.lib().speqg_scan_fusion_sets(self.h, None, None, 0, C.byref(n), C.byref(lost)))
        keys, cnts = np.zeros(max(1, n.value), np.uint64), np.zeros(max(1, n.value), np.uint64)
        capi.check(capi.lib().speqg_scan_fusion_sets(self.h, keys.ctypes.data, cnts.ctypes.data, n.value, C.byref(n), C.byref(lost)))
        if lost.value:
            raise capi.SpeqError(capi.ESTATE, f"fusion map overflowed: {lost.value} ambiguous units without a set")
        out = {}
        for key, c in zip(keys[:n.value].tolist(), cnts[:n.value].tolist()):
            if self.V <= 64:
                t = tuple(g for g in range(64) if (key >> g) & 1)
            else:
                t = (key & 0xFFFF, (key >> 16) & 0xFFFF) + (("+",) if (key >> 32) & 1 else ())
            out[t] = out.get(t, 0) + int(c)
        return out

    def device_counters(self):
        """The handle's counters as torch tensors over its device memory (int64 view of the u64 counters, float64 local-mode
        tallies): a multi-GPU driver all-reduces them in place on the scan's stream before finish()."""
        import torch
        pu, pf, nu, nf = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint64()
        capi.check(capi.lib().speqg_scan_device_counters(self.h, C.byref(pu), C.byref(nu), C.byref(pf), C.byref(nf)))

        class _View:
            def __init__(self, ptr, n, typestr):
                self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}

        dev = torch.device("cuda", self.device)
        return (torch.as_tensor(_View(pu.value, nu.value, "<i8"), device=dev), torch.as_tensor(_View(pf.value, nf.value, "<f8"), device=dev))

    def refine(self, percent: np.ndarray, group_counts: np.ndarray) -> np.ndarray:
        p = np.ascontiguousarray(percent, np.float64)
        g = np.ascontiguousarray(group_counts, np.int32)
        out = np.zeros(self.V, np.float64)
        capi.check(capi.lib().speqg_refine(self.h, p.ctypes.data, g.ctypes.data, out.ctypes.data))
        return out

    def multiplicity_count(self) -> int:
        n = C.c_uint64()
        capi.check(capi.lib().speqg_scan_multiplicity_count(self.h, C.byref(n)))
        return int(n.value)

    def close(self) -> None:
        if self.h:
            capi.lib().speqg_scan_close(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def backcheck(index: Index, k: int, device: int = 0) -> tuple[np.ndarray, np.ndarray]:
    """unique_ref[V], total_ref[V] (what the reference caches in <stem>_<k>mer.dat)."""
    u = np.zeros(index.n_groups, np.uint64)
    t = np.zeros(index.n_groups, np.uint64)
    capi.check(capi.lib().speqg_backcheck(index.h, device, k, u.ctypes.data, t.ctypes.data))
    return u, t


def unique_to_percent(unique_in_reads, total_in_reads: int, unique_in_refs, total_in_refs) -> np.ndarray:
    a = np.ascontiguousarray(unique_in_reads, np.float64)
    b = np.ascontiguousarray(unique_in_refs, np.float64)
    c = np.ascontiguousarray(total_in_refs, np.float64)
    out = np.zeros(a.size, np.float64)
    capi.check(capi.lib().speqg_unique_to_percent(a.ctypes.data, int(total_in_reads), b.ctypes.data, c.ctypes.data, a.size,
                                                  out.ctypes.data))
    return out


def solve(unique_totals: np.ndarray, total_kmers: int, unique_ref, total_ref, group_counts, estimate, precision: float = 1e-6,
          max_iters: int = 100000):
    """The loop of src/fm_scanner.cpp:238-279.  estimate(percent) -> tkpg[V] (Scan.refine, all-reduced by the caller when
    the reads are sharded).  Returns (percent_initial, percent_final, iterations, history)."""
    V = len(unique_totals)
    percent = unique_to_percent(unique_totals, total_kmers, np.asarray(unique_ref, np.float64), np.asarray(total_ref, np.float64))
    initial = percent.copy()
    diff = np.ones(V)
    iters = 0
    hist = []
    while V > 0 and diff.max() > precision and iters < max_iters:
        tkpg = estimate(percent)
        nxt = unique_to_percent(unique_totals, total_kmers, unique_totals, tkpg)
        diff = np.abs(nxt - percent)
        percent = nxt
        hist.append((percent.copy(), tkpg.copy()))
        iters += 1
    return initial, percent, iters, hist
