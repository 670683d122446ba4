"""speqg_refine_noalloc: the refinement estimate for callers that cannot call cudaMalloc -- caller's device arrays in and out,
the per-block partial sums in a file-scope __device__ array of fixed size (speq_b200/csrc/noalloc.cu; no counterpart in the
reference).  Checked through the C ABI against speqg_refine on the same handle.
This file is collected last on purpose."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

import helpers
from speq_b200 import capi, engine

pytestmark = pytest.mark.gpu
TOL = 1e-12          # the tolerance of the fp64 parity tests (tests/test_gpu_parity.py)


def _noalloc(sc, percent: np.ndarray, counts: np.ndarray) -> np.ndarray:
    import torch
    dev = torch.device("cuda", sc.device)
    p = torch.as_tensor(np.ascontiguousarray(percent, np.float64)).to(dev)
    g = torch.as_tensor(np.ascontiguousarray(counts, np.int32)).to(dev)
    out = torch.full((sc.V,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    capi.check(capi.lib().speqg_refine_noalloc(sc.h, C.c_void_p(p.data_ptr()), C.c_void_p(g.data_ptr()), C.c_void_p(out.data_ptr())))
    sc.sync()                                      # the kernels run on the scan's stream
    return out.cpu().numpy()


@pytest.mark.parametrize("algorithm", [capi.ALGO_AUTO, capi.ALGO_ROLLING])
def test_refine_noalloc_matches_refine(lib, algorithm):
    fx = helpers.load_golden("global_single")
    scaff, gs, counts, seqs, quals = helpers.golden_inputs(fx)
    V = len(counts)
    idx = engine.Index.from_scaffolds(scaff, gs, V)
    sc = engine.Scan(idx, k=fx["run"]["k"], record=True, algorithm=algorithm)
    seq, q, offs = helpers.flat_reads(seqs, quals)
    sc.submit_raw(seq, q, offs, len(seqs))
    r = sc.finish()
    assert r.total_kmers > 0
    gc = np.asarray(counts, np.int32)
    p1 = np.full(V, 100.0 / V)
    # work is still due on the handle (the recorded multiplicities are not folded yet): refused, nothing launched
    if algorithm == capi.ALGO_AUTO:
        with pytest.raises(capi.SpeqError) as ei:
            _noalloc(sc, p1, gc)
        assert ei.value.code == capi.ESTATE and "pending" in ei.value.message
    want1 = sc.refine(p1, gc)                      # the existing entry point (folds the multiplicities)
    assert want1.sum() > 0
    # two calls in a row, with different inputs: the second must not see anything of the first in the shared scratch
    rng = np.random.default_rng(5)
    p2 = rng.random(V) * 100.0
    p2[0] = 0.0                                    # a variant at zero drops out of the estimate (src/fm_scanner.cpp:1099-1119)
    got1 = _noalloc(sc, p1, gc)
    got2 = _noalloc(sc, p2, gc)
    want2 = sc.refine(p2, gc)
    assert helpers.rel_close(got1, want1, TOL), (got1, want1)
    assert helpers.rel_close(got2, want2, TOL), (got2, want2)
    assert not np.array_equal(want1, want2)
    # and once more in the first form, after the second
    assert helpers.rel_close(_noalloc(sc, p1, gc), want1, TOL)
    sc.close()
    idx.close()


def test_refine_noalloc_rejects_sizes_beyond_the_scratch(lib):
    """variants x refinement grid (4 blocks per multiprocessor) must fit SPEQG_NOALLOC_MAX_DOUBLES = 131072 doubles: 300
    variants do not on any device with more than 109 multiprocessors."""
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    V = max(300, 131072 // (4 * sms) + 1)
    rng = np.random.default_rng(9)
    letters = "ACGT"
    scaff = [engine.ascii_to_ranks("".join(letters[i] for i in rng.integers(0, 4, 48))) for _ in range(V)]
    idx = engine.Index.from_scaffolds(scaff, list(range(V)), V)
    sc = engine.Scan(idx, k=12, record=True)
    with pytest.raises(capi.SpeqError) as ei:
        _noalloc(sc, np.full(V, 100.0 / V), np.ones(V, np.int32))
    assert "exceeds the fixed scratch" in str(ei.value)
    # a scan without recorded multiplicities: wrong state, as for speqg_refine
    s2 = engine.Scan(idx, k=12)
    with pytest.raises(capi.SpeqError):
        _noalloc(s2, np.full(V, 100.0 / V), np.ones(V, np.int32))
    s2.close()
    sc.close()
    idx.close()
