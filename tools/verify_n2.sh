mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nccl.py -x -q > gpurun_out/nccl_world2.log 2>&1; echo "rc=$?" >> gpurun_out/nccl_world2.log
tail -4 gpurun_out/nccl_world2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
cut -c1-900 gpurun_out/bench_n2.json; tail -3 gpurun_out/bench_n2.err
