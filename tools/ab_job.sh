V=$PWD/build/variants
sh tools/ab_n1.sh cur=A=1 allcoh=WENOEXT_LIB=$V/allcoh/libwenoext_b200.so cur2=A=1 allcoh2=WENOEXT_LIB=$V/allcoh/libwenoext_b200.so
BENCH_ARGS="--config c4" sh tools/ab_n1.sh c4cur=A=1 c4allcoh=WENOEXT_LIB=$V/allcoh/libwenoext_b200.so
WENOEXT_LIB=$V/allcoh/libwenoext_b200.so python bench.py --n 48 --steps 3 --no-e2e --no-cpu-baseline --sustained-s 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('allcoh parity', d['parity_vs_oracle'])" > gpurun_out/ab1_parity.txt 2>&1
