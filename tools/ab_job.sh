V=$PWD/build/variants
python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_gpu_tma2d.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/ab1_tests.log
sh tools/ab_n1.sh new=A=1 base=WENOEXT_LIB=$V/base/libwenoext_b200.so new2=A=1 base2=WENOEXT_LIB=$V/base/libwenoext_b200.so
