python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-cpu --no-c1 --latency-frames 20 2>gpurun_out/q.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d.get('consumers'))"
tail -3 gpurun_out/q.err
