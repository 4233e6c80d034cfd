"""Copy one round's GPU evidence from gpurun_out/ into profiles/ (tracked): bench lines, launch list, ncu summaries, DRAM traffic.
usage: python scripts/store_round.py <tag> [streams]"""
import json, os, re, subprocess, sys
tag = sys.argv[1]
streams = int(sys.argv[2]) if len(sys.argv) > 2 else 512
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go, pr = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")
unit = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}
traffic = {"workload": f"C2 752x480, {streams} streams per launch, steady state (bench.py --steps 3 --warmup 3)",
           "source": f"ncu --set full --clock-control none, profiles/{tag}_ncu_<kernel>.txt: dram__bytes_read.sum + dram__bytes_write.sum per launch",
           "streams": streams, "kernels": {}}
for k, name in [("lk_fast21", "lk"), ("shi_tomasi", "shi_tomasi"), ("clahe_lut", "clahe_lut"), ("clahe_apply", "clahe_apply"), ("pyr_down", "pyr_down"),
                ("pyr_pad_x", "pyr_pad_x"), ("pyr_pad_y", "pyr_pad_y"),
                ("post_lk_ransac_set_mask", "post_lk_ransac_set_mask"), ("select_finish", "select_finish")]:
    rep = os.path.join(go, f"{tag}_{k}_kernel.ncu-rep")
    if not os.path.exists(rep):
        continue
    txt = subprocess.run([sys.executable, os.path.join(root, "scripts", "ncu_summary.py"), rep], capture_output=True, text=True).stdout
    open(os.path.join(pr, f"{tag}_ncu_{k}.txt"), "w").write(txt)
    r = re.search(r"dram__bytes_read.sum \[(\w+)\] = ([\d.]+)", txt)
    w = re.search(r"dram__bytes_write.sum \[(\w+)\] = ([\d.]+)", txt)
    d = re.search(r"gpu__time_duration.sum \[(\w+)\] = ([\d.]+)", txt)
    b = float(r.group(2)) * unit[r.group(1)] + float(w.group(2)) * unit[w.group(1)]
    traffic["kernels"][name] = {"dram_bytes_per_launch": b, "dram_bytes_per_stream": b / streams, "ncu_duration_" + d.group(1): float(d.group(2))}
json.dump(traffic, open(os.path.join(pr, f"{tag}_traffic.json"), "w"), indent=1)
lat = os.path.join(go, f"{tag}_latency.txt")
if os.path.exists(lat):
    open(os.path.join(pr, f"{tag}_latency.txt"), "w").write(open(lat).read())
src = os.path.join(go, f"{tag}_launches.csv")
if os.path.exists(src):
    open(os.path.join(pr, f"{tag}_launches.csv"), "w").write(open(src).read())
for f in ("c2", "c1", "c3", "c4", "ref"):
    p = os.path.join(go, f"{tag}_bench_{f}.json")
    if os.path.exists(p) and os.path.getsize(p):
        d = json.loads(open(p).read().strip().splitlines()[-1])
        json.dump(d, open(os.path.join(pr, f"{tag}_bench_{f}.json"), "w"), indent=1)
        print(f, "value", round(d["value"]), "e2e", d.get("e2e", {}).get("value"), "lat", d.get("latency_us"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
        if "kernels" in d:
            print("   ", {k: round(v["ms_per_step"], 3) for k, v in d["kernels"].items()}, "step", round(d["ms_per_step"], 3))
            print("   ", {k: round(v["frac"], 3) for k, v in d["stages"].items()}, "frame frac", round(d["frame_hbm_frac"], 4), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 4), d["roofline"].get("traffic"))
