# e2e sweep over chunk-schedule knobs: each line "VAR=val VAR2=val2"
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
while read -r line; do
  [ -z "$line" ] && continue
  echo "== $line"
  env $line timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('ms/step %.3f e2e %.3f ms frac %.3f band %.3f dev %.3f chunks %d' % (d['ms_per_step'], d['e2e']['ms_per_step'], r['frac'], r['ms_band'], r['ms_step_device'], r['chunks']))"
done <<'EOL'
NLDSC_CHUNK_ROWS=6144
NLDSC_CHUNK_ROWS=8192
NLDSC_CHUNK_ROWS=10240
NLDSC_CHUNK_ROWS=12288
NLDSC_CHUNK_ROWS=16384
NLDSC_CHUNK_ROWS=6144
NLDSC_TAPER=0
EOL
