mkdir -p gpurun_out
for s in 32 16 24 48 64 32; do
WENOEXT_E2E_SLABS=$s python bench.py --no-cpu-baseline --no-parity --sustained-s 0 > gpurun_out/e2e_s$s.json 2> gpurun_out/e2e_s$s.err
python - $s <<'PY'
import json,sys
s=sys.argv[1]
d=json.loads(open(f"gpurun_out/e2e_s{s}.json").read().strip().splitlines()[-1])
print("slabs",s,"dev ms",round(d["ms_per_step"],3),"e2e ms",round(d["e2e"]["ms_per_step"],3),"div",round(d["e2e"]["div"]["ms_per_step"],3),"div_flux_once",round(d["e2e"]["div_flux_once"]["ms_per_step"],3))
PY
done
python bench.py --no-cpu-baseline --no-parity --sustained-s 0 > gpurun_out/e2e_default.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/e2e_default.json').read().strip().splitlines()[-1]); print('default dev', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3))"
