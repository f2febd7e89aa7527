#!/bin/bash
# round 2, call ZG (1 GPU): clocks / power / throttle reasons during the 1.25 s end-to-end configs[4] call, both bodies
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=timestamp,clocks.sm,clocks.max.sm,power.draw,power.limit,temperature.gpu,clocks_throttle_reasons.active,clocks_throttle_reasons.sw_power_cap,clocks_throttle_reasons.hw_slowdown,clocks_throttle_reasons.sw_thermal_slowdown --format=csv -lms 50 > gpurun_out/r02zg_smi.csv 2>&1 &
SMI=$!
AB_OUT=r02zg_cfg4_e2e_ab.json timeout 400 python scripts/cfg4_e2e_ab.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-250
kill $SMI
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02zg_smi.csv')))
hdr=rows[0]; data=[r for r in rows[1:] if len(r)==len(hdr)]
busy=[r for r in data if float(r[3].split()[0])>400]
print(len(data),'samples,',len(busy),'above 400 W')
import statistics
if busy:
    clk=[int(r[1].split()[0]) for r in busy]; pw=[float(r[3].split()[0]) for r in busy]
    print('sm MHz under load: min',min(clk),'median',statistics.median(clk),'max',max(clk),'| max clock',busy[0][2],'| power median',statistics.median(pw),'max',max(pw),'limit',busy[0][4])
    print('sw_power_cap active in', sum('Active' in r[7] and 'Not' not in r[7] for r in busy),'of',len(busy),'; hw_slowdown',sum('Active' in r[8] and 'Not' not in r[8] for r in busy),'; sw_thermal',sum('Active' in r[9] and 'Not' not in r[9] for r in busy))
PY
