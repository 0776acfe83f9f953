mkdir -p gpurun_out
V=$PWD/build/variants
L=libwenoext_b200.so
WENOEXT_LIB=$V/ts/$L timeout 300 python bench.py --n 48 --steps 3 --no-e2e --no-cpu-baseline --sustained-s 0 2>gpurun_out/ts_parity.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ts parity', d['parity_vs_oracle'])" || { tail -5 gpurun_out/ts_parity.err; exit 1; }
WENOEXT_LIB=$V/ts/$L timeout 600 python -m pytest tests/test_gpu_parity.py -k "hex8 or cyclic-0 or cyclic-1 or multirank or pipelined or device_buffers or face_flux" -x -q 2>&1 | tail -3
sh tools/ab_n1.sh cur=A=1 ts=WENOEXT_LIB=$V/ts/$L cur2=A=1 ts2=WENOEXT_LIB=$V/ts/$L tsoff=WENOEXT_LIB=$V/ts/$L,WENOEXT_SUM=0
for t in cur ts cur2 ts2 tsoff; do python - $t <<'PY'
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
