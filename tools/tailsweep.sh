for t in 0 9 17 33 65 129; do python bench.py --steps 30 --warmup 3 --no-cpu-baseline --coarse-tail $t | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('tail',$t, d['ms_per_step'], d['gpu_launches']/30)"; done
