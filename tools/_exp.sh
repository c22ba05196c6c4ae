python -m pytest tests/test_project.py tests/test_kernels.py tests/test_golden.py tests/test_baseline_configs.py -x -q -m gpu 2>&1 | tail -2
run() { python bench.py --ncell $1 --steps 5 --warmup 3 --no-e2e --no-cpu --no-workloads $2 > /tmp/o.json 2> /tmp/e.log; if [ -s /tmp/o.json ]; then tail -1 /tmp/o.json | python -c "
import sys,json
r=json.loads(sys.stdin.read()); print('  n=$1 $2', 'ms_solve/vc', round(r['ms_solve_per_vcycle'],4), 'ms/step', round(r['ms_per_step'],3), 'vc', r['vcycles_per_step'], r['rel_res_history'][-1])"; else tail -3 /tmp/e.log; fi; }
for n in 64 256 512; do run $n; done
