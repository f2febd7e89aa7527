cd $GRAFT_REPO_ROOT
show() { python -c "
import sys,json
for l in sys.stdin:
    n,j=l.split(' ',1); d=json.loads(j); r=d['resident']
    print('%-10s e2e_s %.4f band_ms %.2f step_ms %.2f algorithmic(step) %.0f TOP/s' % (n, d['e2e_s'], r['ms_band'], r['ms_step'], r['step_algorithmic_top_s']))"; }
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -1
timeout 600 python scripts/scale_check.py wide_band cfg5_slice ukb_slice 2>&1 | grep -E "^(ukb|wide|cfg5)" | show
python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('cfg2 ms/step %.3f e2e %.3f band %.3f frac %.3f' % (d['ms_per_step'], d['e2e']['ms_per_step'], r['ms_band'], r['frac']))"
