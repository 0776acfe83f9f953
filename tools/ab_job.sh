V=$PWD/build/variants
sh tools/ab_n1.sh base=A=1 w9=WENOEXT_LIB=$V/w9/libwenoext_b200.so w10=WENOEXT_LIB=$V/w10/libwenoext_b200.so w11=WENOEXT_LIB=$V/w11/libwenoext_b200.so w9r3=WENOEXT_LIB=$V/w9/libwenoext_b200.so,WENOEXT_SLOTS=3 base2=A=1
for v in w9 w10 w11; do WENOEXT_LIB=$V/$v/libwenoext_b200.so python bench.py --n 48 --steps 3 --no-e2e --no-cpu-baseline --sustained-s 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v parity', d['parity_vs_oracle'])"; done > gpurun_out/ab1_parity.txt 2>&1
