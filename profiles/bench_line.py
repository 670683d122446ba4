"""Print the headline fields of bench.py JSON lines: python profiles/bench_line.py file.json ..."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get("roofline") or {}
        t = d.get("tallies_only") or {}
        ra = r.get("random_access")          # round-1 lines
        if ra:
            print(f, round(d["value"] / 1e9, 1), "G/s", round(d["ms_per_step"], 3), "ms", "frac", round(ra["frac"], 3), ra["structure_bytes"] >> 20, "MB")
            continue
        e = lambda k: round(d[k]["value"] / 1e9, 2) if d.get(k) else None
        print(f, d["config"]["config"], "N=%d" % d["n_gpus"], d.get("resolved_algorithm"), "value", round(d["value"] / 1e9, 1), "G lookups/s",
              round(d["ms_per_step"], 3), "ms/step", "| tallies only", round(t["value"] / 1e9, 1) if t else None,
              "| roofline", r.get("bound"), "frac", None if r.get("frac") is None else round(r["frac"], 3),
              "peak", None if not r.get("peak_Gsectors_per_s") else round(r["peak_Gsectors_per_s"], 1), "G sectors/s",
              "| e2e", e("e2e"), "fastq", e("e2e_fastq"), "| cpu", None if not d.get("cpu_baseline") else d["cpu_baseline"]["value"],
              "| parity", (d.get("parity") or {}).get("ok"))
    except Exception as ex:
        print(f, "unreadable:", ex)
