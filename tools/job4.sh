mkdir -p gpurun_out
V=$PWD/build/variants
L=libwenoext_b200.so
# parity of each variant at 48^3 first
for v in pf6 ispw ispwpf6; do
WENOEXT_LIB=$V/$v/$L python bench.py --n 48 --steps 3 --no-e2e --no-cpu-baseline --sustained-s 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v parity', d['parity_vs_oracle'])"
done
sh tools/ab_n1.sh base=A=1 pf6=WENOEXT_LIB=$V/pf6/$L pf3=WENOEXT_LIB=$V/pf3/$L ispw=WENOEXT_LIB=$V/ispw/$L ispwpf6=WENOEXT_LIB=$V/ispwpf6/$L base2=A=1 pf6b=WENOEXT_LIB=$V/pf6/$L ispwb=WENOEXT_LIB=$V/ispw/$L
for t in base pf6 pf3 ispw ispwpf6 base2 pf6b ispwb; do python - $t <<'PY'
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
