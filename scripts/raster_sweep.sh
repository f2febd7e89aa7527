cd $GRAFT_REPO_ROOT
for r in 1 4 8 16 32; do
  echo "== NLDSC_RASTER=$r"
  NLDSC_RASTER=$r timeout 300 python scripts/scale_check.py cfg5_slice 2>&1 | tail -1 | python -c "
import sys,json
n,j=sys.stdin.read().split(' ',1); d=json.loads(j); r=d['resident']; print('e2e_s %.4f band_ms %.2f executed %.0f algorithmic %.0f' % (d['e2e_s'], r['ms_band'], r['band_executed_top_s'], r['band_algorithmic_top_s']))"
done
for r in 1 8 32; do
  echo "== cfg2 NLDSC_RASTER=$r"
  NLDSC_RASTER=$r python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('ms/step %.3f e2e %.3f band %.3f frac %.3f' % (d['ms_per_step'], d['e2e']['ms_per_step'], r['ms_band'], r['frac']))"
done
