"""Deterministic synthetic inputs of the shapes BASELINE.json names.

refs : ancestor of length G uniform over ACGT; variant v = ancestor with i.i.d. substitutions at rate d_var; isolate =
       variant with substitutions at rate d_iso; FASTA order variant-major; groups file `V{v}({I}): {vI}-{vI+I-1}`.
reads: variant ~ abundance (prop. to 1/(v+1)), isolate uniform, strand uniform, start uniform in [0, G-L], substitution
       errors i.i.d. at rate e; qualities constant 'I' (Q40), or per base from {Q20 5%, Q30 15%, Q37 80%} (local mode).
Generation runs on torch (CPU for the tests, CUDA for the bench) -- data plumbing only.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

ASCII = np.frombuffer(b"ACGT", dtype=np.uint8)


@dataclass(frozen=True)
class Shape:
    name: str
    G: int          # genome length
    V: int          # variants
    I: int          # isolates per variant
    R: int          # reads
    L: int          # read length
    k: int
    e: float        # substitution error rate of the reads
    d_var: float = 0.02
    d_iso: float = 0.002
    fixed_accuracy: float | None = None   # --fixed-accuracy for global mode (default 1 - e)

    @property
    def accuracy(self) -> float:
        return self.fixed_accuracy if self.fixed_accuracy is not None else 1.0 - self.e

    @property
    def windows_per_read(self) -> int:
        return max(0, self.L - self.k + 1)


# the five configs of BASELINE.json (ids 1..5) -- SURVEY.md §8 shorthand C1..C5
CONFIGS = {
    "C1": Shape("C1: 2 var x 3 iso x 30 kb, 100k x 150 bp, k=21", 30_000, 2, 3, 100_000, 150, 21, 0.005),
    "C2": Shape("C2: 10 var x 20 iso x 10 kb, 10M x 150 bp, k=21", 10_000, 10, 20, 10_000_000, 150, 21, 0.005),
    "C3": Shape("C3: 50 var x 10 iso x 30 kb, 200M x 150 bp, k=25", 30_000, 50, 10, 200_000_000, 150, 25, 0.005),
    "C4": Shape("C4: 20 strains x 5 Mb, 100M x 150 bp, k=31", 5_000_000, 20, 1, 100_000_000, 150, 31, 0.005, d_var=0.01),
    "C5": Shape("C5: 20 var x 30 kb, 1M x 10 kb, 5% error, k=15", 30_000, 20, 1, 1_000_000, 10_000, 15, 0.05, fixed_accuracy=0.95),
}
SEED0 = 0x5BE0


def _gen(seed: int, device) -> torch.Generator:
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def _mutate(codes: torch.Tensor, rate: float, g: torch.Generator) -> torch.Tensor:
    """i.i.d. substitutions: a hit base becomes one of the three other bases."""
    if rate <= 0:
        return codes.clone()
    hit = torch.rand(codes.shape, generator=g, device=codes.device) < rate
    shift = torch.randint(1, 4, codes.shape, generator=g, device=codes.device, dtype=torch.uint8)
    return torch.where(hit, (codes + shift) & 3, codes)


def make_refs(sh: Shape, seed: int, device="cpu") -> torch.Tensor:
    """[V*I, G] uint8 codes (A0 C1 G2 T3), variant-major."""
    g = _gen(seed, device)
    anc = torch.randint(0, 4, (sh.G,), generator=g, device=device, dtype=torch.uint8)
    rows = []
    for _ in range(sh.V):
        var = _mutate(anc, sh.d_var, g)
        for _ in range(sh.I):
            rows.append(_mutate(var, sh.d_iso, g))
    return torch.stack(rows)


def abundance(V: int) -> np.ndarray:
    w = 1.0 / (np.arange(V) + 1.0)
    return w / w.sum()


def make_reads(sh: Shape, refs: torch.Tensor, n_reads: int, seed: int, *, local_quals: bool = False, n_rate: float = 0.0):
    """-> (seq ASCII uint8 [n, L], qual ASCII uint8 [n, L]) on refs.device."""
    dev = refs.device
    g = _gen(seed, dev)
    ab = torch.tensor(abundance(sh.V), dtype=torch.float64, device=dev)
    var = torch.multinomial(ab, n_reads, replacement=True, generator=g)
    iso = torch.randint(0, sh.I, (n_reads,), generator=g, device=dev)
    strand = torch.randint(0, 2, (n_reads,), generator=g, device=dev, dtype=torch.uint8)
    start = torch.randint(0, sh.G - sh.L + 1, (n_reads,), generator=g, device=dev)
    row = var * sh.I + iso
    flat = refs.reshape(-1)
    idx = (row * sh.G + start)[:, None] + torch.arange(sh.L, device=dev)[None, :]
    codes = flat[idx]
    del idx
    rc = (3 - codes).flip(1)
    codes = torch.where(strand[:, None] == 1, rc, codes)
    del rc
    codes = _mutate(codes, sh.e, g)
    lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device=dev)
    seq = lut[codes.long()]
    if n_rate > 0:
        seq = torch.where(torch.rand(seq.shape, generator=g, device=dev) < n_rate, torch.tensor(ord("N"), dtype=torch.uint8, device=dev), seq)
    if local_quals:
        u = torch.rand(seq.shape, generator=g, device=dev)
        q = torch.where(u < 0.05, 20, torch.where(u < 0.20, 30, 37)).to(torch.uint8) + 33
    else:
        q = torch.full(seq.shape, ord("I"), dtype=torch.uint8, device=dev)
    return seq, q


def refs_to_ranks(refs: torch.Tensor) -> list[np.ndarray]:
    """codes (A0 C1 G2 T3) -> dna5 ranks (A0 C1 G2 T4) per scaffold."""
    a = refs.cpu().numpy()
    lut = np.array([0, 1, 2, 4], dtype=np.uint8)
    return [lut[r] for r in a]


def group_scaffolds(sh: Shape) -> np.ndarray:
    return np.repeat(np.arange(sh.V, dtype=np.int32), sh.I)


def write_fasta(path: str, refs: torch.Tensor, width: int = 70) -> None:
    a = ASCII[refs.cpu().numpy()]
    with open(path, "wb") as f:
        for j, r in enumerate(a):
            f.write(b">scaffold_%d\n" % j)
            b = r.tobytes()
            for i in range(0, len(b), width):
                f.write(b[i:i + width])
                f.write(b"\n")


def write_groups(path: str, sh: Shape) -> None:
    with open(path, "w") as f:
        f.write("# synthetic groups file (format of src/file_to_map.cpp)\n")
        for v in range(sh.V):
            lo, hi = v * sh.I, v * sh.I + sh.I - 1
            f.write(f"V{v}({sh.I}): {lo}-{hi}\n" if sh.I > 1 else f"V{v}({sh.I}): {lo}\n")


def write_fastq(path: str, seq: np.ndarray, qual: np.ndarray) -> None:
    """seq/qual: [n, L] ASCII arrays (or lists of 1-D arrays for ragged reads)."""
    with open(path, "wb") as f:
        for i in range(len(seq)):
            f.write(b"@r%d\n" % i)
            f.write(np.asarray(seq[i], np.uint8).tobytes())
            f.write(b"\n+\n")
            f.write(np.asarray(qual[i], np.uint8).tobytes())
            f.write(b"\n")
