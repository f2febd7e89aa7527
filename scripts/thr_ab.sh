cd $GRAFT_REPO_ROOT
show() { python -c "
import sys,json
for l in sys.stdin:
    n,j=l.split(' ',1); d=json.loads(j); r=d['resident']
    print('%-10s e2e_ms %.3f band_ms %.3f step_ms %.3f pairs/s(step) %.4g' % (n, 1e3*d['e2e_s'], r['ms_band'], r['ms_step'], d['pairs']/(r['ms_step']*1e-3)))"; }
for w in 0 1; do echo "== WIDE=$w"; NLDSC_WIDE=$w timeout 300 python scripts/scale_check.py kg_panel mid_cohort 2>&1 | grep -E "^(kg|mid)" | show; done
