# component-lane kernel (VEC): parity, then same-box A/B against the cell-lane kernel
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_div.py tests/test_gpu_tma2d.py -x -q > gpurun_out/vec_parity.log 2>&1; echo "pytest rc=$?" >> gpurun_out/vec_parity.log
tail -3 gpurun_out/vec_parity.log
sh tools/ab_n1.sh old=WENOEXT_VEC=0 vec=WENOEXT_VEC=1 vec12=WENOEXT_VEC=1,WENOEXT_WARPS=12 vecr3=WENOEXT_VEC=1,WENOEXT_SLOTS=3 vecr5=WENOEXT_VEC=1,WENOEXT_SLOTS=5 old2=WENOEXT_VEC=0 vec2=WENOEXT_VEC=1
for t in old vec vec12 vecr3 vecr5 old2 vec2; do python - $t <<'PY'
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
