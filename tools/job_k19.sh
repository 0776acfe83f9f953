mkdir -p gpurun_out
echo "== main lib"; python tools/tune.py 64 3 3 "4,0,1,0" "4,0,2,0" "4,0,3,0" "4,6,2,0" "4,8,1,0" 2>&1 | tail -7
echo "== k19 variant (no face prefetch)"; WENOEXT_LIB=$PWD/build/variants/k19/libwenoext_b200.so python tools/tune.py 64 3 3 "4,0,1,0" "4,0,2,0" "4,0,3,0" "4,6,2,0" 2>&1 | tail -6
