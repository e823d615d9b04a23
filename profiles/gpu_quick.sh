# scratch: last sanity check of the shipped build
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 2
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3
timeout 600 python bench.py --steps 200 --warmup 5 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print(d['value'] / 1e6, d['ms_per_step'], d['e2e']['value'] / 1e6, d['roofline']['frac'], d['roofline']['kernels']['post_physics']['frac'])"
