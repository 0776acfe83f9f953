mkdir -p gpurun_out
V=$PWD/build/variants
L=libwenoext_b200.so
WENOEXT_LIB=$V/vr/$L python bench.py --n 48 --steps 3 --no-e2e --no-cpu-baseline --sustained-s 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('vr parity', d['parity_vs_oracle'])"
sh tools/ab_n1.sh head=WENOEXT_LIB=$V/head/$L cur=A=1 vr=WENOEXT_LIB=$V/vr/$L cur2=A=1 vr2=WENOEXT_LIB=$V/vr/$L
BENCH_ARGS="--config c4" sh tools/ab_n1.sh c4head=WENOEXT_LIB=$V/head/$L c4cur=A=1 c4head2=WENOEXT_LIB=$V/head/$L c4cur2=A=1
for t in head cur vr cur2 vr2 c4head c4cur c4head2 c4cur2; do python - $t <<'PY'
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
