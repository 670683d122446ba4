"""Index build wall-clock, host SA-IS vs suffix array on the device (SPEQG_BUILD_DEVICE_SA), with the per-phase lines of
SPEQB_TIMING.  Usage: SPEQB_TIMING=1 python profiles/index_build_time.py C3 C4"""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from speq_b200 import synth, engine
for cfg in sys.argv[1:] or ["C3"]:
    sh = synth.CONFIGS[cfg]
    refs = synth.make_refs(sh, synth.SEED0 + int(cfg[1]), device="cuda")
    sc, gs = synth.refs_to_ranks(refs), synth.group_scaffolds(sh)
    out = {}
    for dev in (True, False):
        t = time.time(); idx = engine.Index.from_scaffolds(sc, gs, sh.V, device_sa=dev); out[dev] = time.time() - t
        print(cfg, "device_sa" if dev else "host sa-is", round(out[dev], 2), "s  n =", idx.n, flush=True)
        if dev: ex = idx.export()
        else:
            ex2 = idx.export()
            print(cfg, "identical index arrays:", all(np.array_equal(ex[k], ex2[k]) for k in ("bwt", "group_of_sa", "lcp")), flush=True)
        idx.close()
    del refs; torch.cuda.empty_cache()
