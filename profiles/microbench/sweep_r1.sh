run() { python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],1), round(d['roofline']['kernel_ms'],1), round(d['roofline']['frac'],3), round(d['e2e']['value'],3))"; }
run reg_default
