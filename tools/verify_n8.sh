mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
cut -c1-400 gpurun_out/bench_n8.json; tail -2 gpurun_out/bench_n8.err
timeout 900 python -m pytest tests/test_gpu_nccl.py -x -q > gpurun_out/nccl_world8.log 2>&1; echo "rc=$?" >> gpurun_out/nccl_world8.log
tail -4 gpurun_out/nccl_world8.log
