python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for w in C3-hopper C2-swimmer C1-pointmass; do python bench.py --workload $w --steps 10 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])"; done
MJRL_B200_CG_LOOP=py python bench.py --workload C2-swimmer --steps 10 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C2 python loop', d['value'], d['ms_per_step'])"
