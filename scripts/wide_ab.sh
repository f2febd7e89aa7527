cd $GRAFT_REPO_ROOT
for w in 0 1; do
  echo "== NLDSC_WIDE=$w"
  NLDSC_WIDE=$w timeout 600 python scripts/scale_check.py ukb_slice ukb_ring wide_band cfg5_slice 2>&1 | grep -E "^(ukb|wide|cfg5)" | python -c "
import sys,json
for l in sys.stdin:
    n,j=l.split(' ',1); d=json.loads(j); r=d['resident']
    print('%-10s e2e_s %.4f band_ms %.2f step_ms %.2f algorithmic %.0f TOP/s (%.3f of 3296) oracle %s ring_equal %s' % (n, d['e2e_s'], r['ms_band'], r['ms_step'], r['band_algorithmic_top_s'], r['band_algorithmic_top_s']/3296, [(round(o['l2_rel'],16), o['counts_equal']) for o in d['oracle']], d.get('bitwise_equal_to_ukb_slice')))"
  cp gpurun_out/scale_check_ukb_slice_ukb_ring_wide_band_cfg5_slice.json gpurun_out/scale_wide$w.json
done
