mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gputest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputest.log
tail -4 gpurun_out/gputest.log
V=$PWD/build/variants
L=libwenoext_b200.so
sh tools/ab_n1.sh head=WENOEXT_LIB=$V/head/$L new=A=1 head2=WENOEXT_LIB=$V/head/$L new2=A=1
BENCH_ARGS="--config c4" sh tools/ab_n1.sh c4head=WENOEXT_LIB=$V/head/$L c4new=A=1
BENCH_ARGS="--config c2" sh tools/ab_n1.sh c2head=WENOEXT_LIB=$V/head/$L c2new=A=1
for t in head new head2 new2 c4head c4new c2head c2new; do python - $t <<'PY'
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
