# A/B of one env knob on the bench (arg1 = VAR, arg2.. = values) after the GPU tests.
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_all.log
VAR=$1; shift
for v in "$@"; do
  echo "== $VAR=$v"
  env $VAR=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench_$v.err | tee gpurun_out/bench_$VAR$v.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.4g ms/step %.3f e2e %.3f ms frac %.3f band %.3f unpack %.3f dev %.3f' % (d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], r['frac'], r['ms_band'], r['ms_unpack'], r['ms_step_device']))"
  tail -3 gpurun_out/bench_$v.err
done
