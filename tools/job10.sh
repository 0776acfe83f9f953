mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gputest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputest2.log; tail -3 gpurun_out/gputest2.log
python bench.py --config c4 --no-cpu-baseline > gpurun_out/c4_new.json 2> gpurun_out/c4_new.err
python -c "
import json; d=json.loads(open('gpurun_out/c4_new.json').read().strip().splitlines()[-1]); s=d['sustained']
print('c4', round(d['ms_per_step'],3), d['roofline']['frac'], 'sust', s['ms_per_step'], s['roofline_frac'], s['clocks']['sm_mhz'], 'e2e', d['e2e']['ms_per_step'], 'parity', d['parity_vs_oracle'])"
echo "== order 4 (K=34), scalar"; python tools/tune.py 40 4 1 "4,0,1,0" "4,0,2,0" "4,0,0,0" 2>&1 | tail -3
echo "== order 3 scalar"; python tools/tune.py 64 3 1 "4,0,1,0" "4,0,2,0" "4,0,3,0" "4,0,0,0" 2>&1 | tail -4
