for r in 3 4 6; do for v in pyrxw desc3; do cp tools/scratch/var/$v.so my_orb_slam2_b200/liborbfe.so; python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-matchers --no-shapes --latency-frames 50 --rigs $r 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v rigs $r', 'value %.0f e2e %.0f left_only %.0f' % (d['value'], d['e2e']['value'], d['e2e']['left_only']['value']), d['host'].get('h2d_gbs'), d['e2e'].get('e2e_frac_of_link'))
"; done; done
