mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/verify_gputest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/verify_gputest.log; tail -3 gpurun_out/verify_gputest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/verify_bench.json 2> gpurun_out/verify_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/verify_bench.json').read().strip().splitlines()[-1]); e=d['e2e']; s=d['sustained']; print('dev', round(d['ms_per_step'],3), d['roofline']['frac'], 'sust', s['ms_per_step'], s['roofline_frac'], 'e2e', round(e['ms_per_step'],3), 'div', round(e['div']['ms_per_step'],3), 'div_flux_once', round(e['div_flux_once']['ms_per_step'],3), 'maxdiff', e.get('max_abs_diff_vs_device_path'), 'parity', d.get('parity_vs_oracle'))"
