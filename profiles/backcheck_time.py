import time, sys, torch
sys.path.insert(0, '.')
from speq_b200 import synth, engine
for cfg in ("C2", "C3", "C4"):
    sh = synth.CONFIGS[cfg]
    refs = synth.make_refs(sh, synth.SEED0 + int(cfg[1]), device="cuda")
    t = time.time(); idx = engine.Index.from_scaffolds(synth.refs_to_ranks(refs), synth.group_scaffolds(sh), sh.V); tb = time.time() - t
    for i in range(3):
        torch.cuda.synchronize(); t = time.time(); u, tt = engine.backcheck(idx, sh.k); torch.cuda.synchronize(); dt = time.time() - t
        print(cfg, "index build", round(tb, 2), "s; backcheck run", i, round(dt, 4), "s", "windows", int(tt.sum()), flush=True)
    idx.close(); del refs; torch.cuda.empty_cache()
