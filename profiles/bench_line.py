"""Print the headline fields of bench.py JSON lines: python profiles/bench_line.py file.json ..."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        ra = d["roofline"]["random_access"]
        print(f, round(d["value"] / 1e9, 1), "G/s", round(d["ms_per_step"], 3), "ms", "sectors/window",
              round(ra["sectors_per_launch"] / max(1, d["tallies"]["total_kmers"]), 4), "frac", round(ra["frac"], 3),
              ra["structure_bytes"] >> 20, "MB", "e2e", d["e2e"]["value"] / 1e9 if d.get("e2e") else None)
    except Exception as e:
        print(f, "unreadable:", e)
