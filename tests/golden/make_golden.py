"""Generate the SPeQ-level golden fixtures (tests/golden/speq_*.json).

Run in the build container, where /root/reference exists:   python tests/golden/make_golden.py
Each fixture = the input files (as strings) + the output of oracle/_ref/speq_ref, i.e. the reference's glue restated
over the REAL vendored sdsl-lite csa_wt that seqan3 instantiates (compiled from /root/reference by oracle/Makefile).
SPeQ itself ships no golden outputs, so these vectors pin the oracle restatement
(oracle/_build/speq_oracle) and the CUDA path to the sdsl arithmetic.  The fixtures travel; /root/reference does not.
"""
from __future__ import annotations

import json
import sys
import tempfile
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import helpers  # noqa: E402

CASES = {
    # name: (random_case kwargs, run kwargs)
    "global_single": (dict(seed=11, V=3, I=2, G=500, R=240, L=60, e=0.01, n_rate_read=0.004, lowq_rate=0.01, ragged=True, iupac=True),
                      dict(k=13, fixed_accuracy=0.99, paired=False)),
    "local_single": (dict(seed=12, V=3, I=2, G=500, R=200, L=60, e=0.01, local_quals=True, n_rate_read=0.003),
                     dict(k=13, fixed_accuracy=0.0, phred_cutoff=19, paired=False)),
    "global_paired": (dict(seed=13, V=4, I=2, G=400, R=240, L=50, e=0.02, ragged=True, lowq_rate=0.01),
                      dict(k=11, fixed_accuracy=0.98, paired=True)),
    "local_paired": (dict(seed=14, V=2, I=3, G=600, R=200, L=70, e=0.01, local_quals=True),
                     dict(k=15, fixed_accuracy=0.0, phred_cutoff=19, paired=True)),
    "refs_with_n": (dict(seed=15, V=3, I=2, G=400, R=200, L=50, e=0.01, n_rate_ref=0.01, n_rate_read=0.01),
                    dict(k=9, fixed_accuracy=0.99, paired=False)),
    "short_k": (dict(seed=16, V=5, I=1, G=300, R=150, L=40, e=0.03, d_var=0.15),
                dict(k=5, fixed_accuracy=0.97, paired=False)),
    # k > 31 (round 2): reads long enough to hold windows, N / IUPAC letters and low-quality bases so that restarts and parent
    # jumps at depths up to k-1 trigger; ragged lengths give reads below k.  k = 70 + local mode + --phred-cutoff 30 is the
    # reference's default invocation (include/arg_parse.h:20-23)
    "k70_global_single": (dict(seed=21, V=3, I=2, G=2500, R=120, L=180, e=0.004, d_var=0.03, n_rate_read=0.002, lowq_rate=0.003, ragged=True, iupac=True),
                          dict(k=70, fixed_accuracy=0.996, paired=False)),
    "k70_local_single": (dict(seed=22, V=3, I=2, G=2500, R=120, L=180, e=0.004, d_var=0.03, n_rate_read=0.002, lowq_rate=0.003, local_quals="high"),
                         dict(k=70, fixed_accuracy=0.0, paired=False)),
    "k70_global_paired": (dict(seed=23, V=4, I=2, G=2000, R=120, L=160, e=0.004, d_var=0.03, n_rate_read=0.002, lowq_rate=0.003, ragged=True),
                          dict(k=70, fixed_accuracy=0.996, paired=True)),
    "k70_local_paired": (dict(seed=24, V=3, I=2, G=2500, R=120, L=180, e=0.004, d_var=0.03, n_rate_read=0.002, lowq_rate=0.003, local_quals="high"),
                         dict(k=70, fixed_accuracy=0.0, paired=True)),
    "k127_local_single": (dict(seed=25, V=3, I=2, G=2500, R=100, L=240, e=0.002, d_var=0.03, n_rate_read=0.001, lowq_rate=0.002, local_quals="high"),
                          dict(k=127, fixed_accuracy=0.0, paired=False)),
    "k33_global_paired": (dict(seed=26, V=4, I=2, G=1500, R=120, L=100, e=0.01, d_var=0.04, n_rate_read=0.003, lowq_rate=0.005, ragged=True),
                          dict(k=33, fixed_accuracy=0.99, paired=True)),
}


def main(only: list[str] | None = None) -> None:
    ref = helpers.ref_binary()
    assert ref is not None, "needs /root/reference to build oracle/_ref/speq_ref"
    for name, (ckw, rkw) in CASES.items():
        if only and name not in only:
            continue
        case = helpers.random_case(**ckw)
        rkw = dict(rkw)
        paired = rkw.pop("paired")
        with tempfile.TemporaryDirectory() as d:
            refs, groups, r1, r2 = helpers.case_files(Path(d), case, paired)
            out = helpers.run_oracle(ref, refs, groups, r1, reads2=r2, threads=1, **rkw)
            assert out["backend"] == "sdsl"
            for t in [k for k in out if k.startswith("t_")]:
                del out[t]
            fx = {"generator": "tests/golden/make_golden.py", "backend": "oracle/_ref/speq_ref (vendored sdsl-lite a654b57)",
                  "case": ckw, "run": dict(rkw, paired=paired),
                  "refs_fa": refs.read_text(), "groups_txt": groups.read_text(), "reads1_fq": r1.read_text(),
                  "reads2_fq": r2.read_text() if r2 else None, "expected": out}
        (helpers.GOLDEN / f"speq_{name}.json").write_text(json.dumps(fx, indent=0))
        print(name, out["unique"], out["total_kmers"], out["ambiguous"], out["iterations"], out["percent_final"])


if __name__ == "__main__":
    main(sys.argv[1:])     # names of the fixtures to (re)generate; none = all
