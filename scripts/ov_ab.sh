cd $GRAFT_REPO_ROOT
show() { python -c "
import sys,json
for l in sys.stdin:
    n,j=l.split(' ',1); d=json.loads(j); r=d['resident']
    print('%-10s e2e_s %.4f band_ms %.2f step_ms %.2f chunks %d algorithmic(step) %.0f TOP/s' % (n, d['e2e_s'], r['ms_band'], r['ms_step'], r['chunks'], r['step_algorithmic_top_s']))"; }
echo "== WIDE_OVERLAP=1"
NLDSC_WIDE_OVERLAP=1 timeout 600 python scripts/scale_check.py ukb_slice wide_band cfg5_slice cfg1 2>&1 | grep -E "^(ukb|wide|cfg)" | show
echo "== turns (default)"
timeout 600 python scripts/scale_check.py cfg1 2>&1 | grep -E "^(ukb|wide|cfg)" | show
echo "== narrow"
NLDSC_WIDE=0 timeout 600 python scripts/scale_check.py cfg1 2>&1 | grep -E "^(ukb|wide|cfg)" | show
