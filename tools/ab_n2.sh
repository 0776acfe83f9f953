run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cpu-baseline --no-e2e --no-parity --sustained-s 0 --steps 30 > gpurun_out/ab2_$tag.json 2> gpurun_out/ab2_$tag.err; }
run default A=1
run nomerge WENOEXT_MERGE=0
run nocoh WENOEXT_LIB=$PWD/build/variants/nocoh/libwenoext_b200.so
run noabort WENOEXT_LIB=$PWD/build/variants/noabort/libwenoext_b200.so
run nop2p WENOEXT_P2P=0
run default2 A=1
