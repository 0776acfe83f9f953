mkdir -p gpurun_out
V=$PWD/build/variants
L=libwenoext_b200.so
for m in 1 2; do
WENOEXT_PFL2=$m WENOEXT_LIB=$V/pfl2/$L timeout 300 python bench.py --n 48 --steps 3 --no-e2e --no-cpu-baseline --sustained-s 0 2>gpurun_out/pf_parity.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('pfl2=$m parity', d['parity_vs_oracle'])" || { tail -5 gpurun_out/pf_parity.err; exit 1; }
done
sh tools/ab_n1.sh off=WENOEXT_LIB=$V/pfl2/$L,WENOEXT_PFL2=0 pf1=WENOEXT_LIB=$V/pfl2/$L,WENOEXT_PFL2=1 pf2=WENOEXT_LIB=$V/pfl2/$L,WENOEXT_PFL2=2 off2=WENOEXT_LIB=$V/pfl2/$L,WENOEXT_PFL2=0 pf1b=WENOEXT_LIB=$V/pfl2/$L,WENOEXT_PFL2=1 pf2b=WENOEXT_LIB=$V/pfl2/$L,WENOEXT_PFL2=2
for t in off pf1 pf2 off2 pf1b pf2b; do python - $t <<'PY'
import json,sys
t=sys.argv[1]
try:
    d=json.loads(open(f"gpurun_out/ab1_{t}.json").read().strip().splitlines()[-1])
    s=d.get("sustained",{})
    print(t, "ms/step", round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],4), "sust", round(s.get("ms_per_step"),4), round(s.get("roofline_frac"),4), s.get("clocks",{}).get("sm_mhz"))
except Exception as e:
    print(t, "failed", e); print(open(f"gpurun_out/ab1_{t}.err").read()[-800:])
PY
done
