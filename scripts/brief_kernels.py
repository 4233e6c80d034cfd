"""one-line digest of a bench.py JSON line on stdin: frames/s, ms/step, the dense kernels' ms/step"""
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
keep = ("clahe_lut", "clahe_apply", "pyr_down", "pyr_pad", "lk", "shi_tomasi")
print(round(d["value"]), round(d["ms_per_step"], 4), {k: round(v["ms_per_step"], 4) for k, v in d.get("kernels", {}).items() if k in keep})
