import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["value"]), round(d["ms_per_step"],4), {k:round(v["ms_per_step"],4) for k,v in d["kernels"].items() if k in ("clahe_lut","clahe_apply","pyr_down","pyr_pad","lk","shi_tomasi")})
