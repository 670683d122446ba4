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
        import hashlib, tempfile, os
        with tempfile.TemporaryDirectory(dir="/dev/shm") as d:
            idx.save(os.path.join(d, "x.idx"))
            hsh = hashlib.sha256()
            with open(os.path.join(d, "x.idx"), "rb") as f:
                while chunk := f.read(1 << 24):
                    hsh.update(chunk)
        if dev: ex = hsh.hexdigest()
        else: print(cfg, "identical .idx payload (sha256):", ex == hsh.hexdigest(), flush=True)
        idx.close()
    del refs; torch.cuda.empty_cache()
