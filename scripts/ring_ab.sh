cd $GRAFT_REPO_ROOT
show() { python -c "
import sys,json
for l in sys.stdin:
    n,j=l.split(' ',1); d=json.loads(j); r=d['resident']
    print('%-10s e2e_s %.4f band_ms %.2f step_ms %.2f chunks %d algorithmic(step) %.0f TOP/s' % (n, d['e2e_s'], r['ms_band'], r['ms_step'], r['chunks'], r['step_algorithmic_top_s']))"; }
for pb in 12e9 20e9 32e9; do
  for cr in 4096 8192; do
  echo "== PLANE_BYTES=$pb CHUNK_ROWS=$cr"
  NLDSC_PLANE_BYTES=$pb NLDSC_CHUNK_ROWS=$cr timeout 300 python scripts/scale_check.py ukb_slice 2>&1 | grep -E "^ukb" | show
  done
done
echo "== cfg5_full wide, PLANE_BYTES=24e9 CHUNK_ROWS=8192"
NLDSC_PLANE_BYTES=24e9 NLDSC_CHUNK_ROWS=8192 timeout 600 python scripts/scale_check.py cfg5_full 2>&1 | grep -E "^cfg5" | show
echo "== cfg5_full wide, default ring"
timeout 600 python scripts/scale_check.py cfg5_full 2>&1 | grep -E "^cfg5" | show
