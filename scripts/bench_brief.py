"""Read a bench.py JSON line on stdin and print the headline numbers; usage: python bench.py ... | python scripts/bench_brief.py [tag]"""
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1] if len(sys.argv) > 1 else "", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "lat",
      d.get("latency_us") and round(d["latency_us"]["p50"], 1), {k: round(v["ms_per_step"], 4) for k, v in d.get("kernels", {}).items()})
