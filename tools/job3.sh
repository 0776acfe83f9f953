mkdir -p gpurun_out
A="--no-cpu-baseline --no-e2e --no-parity --sustained-s 0"
export WENOEXT_VEC=0
python bench.py --n 64 --steps 2 --warmup 1 $A > gpurun_out/old_ncu_plain.json 2> gpurun_out/old_ncu_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:weno_cell_kernel -s 1 -c 1 -f -o gpurun_out/r02_old_full_64cubed \
    python bench.py --n 64 --steps 2 --warmup 1 $A > gpurun_out/old_ncu2.log 2>&1
tail -2 gpurun_out/old_ncu2.log
unset WENOEXT_VEC
BENCH_ARGS="--config c4" sh tools/ab_n1.sh c4old=WENOEXT_VEC=0 c4vec=WENOEXT_VEC=1 c4vec12=WENOEXT_VEC=1,WENOEXT_WARPS=12
BENCH_ARGS="--config c4tets" sh tools/ab_n1.sh tetold=WENOEXT_VEC=0 tetvec=WENOEXT_VEC=1
for t in c4old c4vec c4vec12 tetold tetvec; do python - $t <<'PY'
import json,sys
t=sys.argv[1]
try:
    d=json.loads(open(f"gpurun_out/ab1_{t}.json").read().strip().splitlines()[-1])
    s=d.get("sustained",{})
    print(t, "ms/step", d["ms_per_step"], "frac", d["roofline"]["frac"], "sust", s.get("ms_per_step"), s.get("roofline_frac"), s.get("clocks",{}).get("sm_mhz"))
except Exception as e:
    print(t, "failed", e); print(open(f"gpurun_out/ab1_{t}.err").read()[-800:])
PY
done
