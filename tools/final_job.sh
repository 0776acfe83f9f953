# HEAD on one GPU: GPU tests, smoke(), both bench arms, other configs, then the ncu evidence (after the plain runs)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/verify_gputest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/verify_gputest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/verify_smoke.log 2>&1
python bench.py > gpurun_out/verify_bench.json 2> gpurun_out/verify_bench.err
python bench.py --impl reference > gpurun_out/verify_bench_ref.json 2> gpurun_out/verify_bench_ref.err
for c in c2 c4 c4tets; do python bench.py --config $c --no-cpu-baseline > gpurun_out/verify_bench_$c.json 2> gpurun_out/verify_bench_$c.err; done
tail -3 gpurun_out/verify_gputest.log; tail -1 gpurun_out/verify_smoke.log; cat gpurun_out/verify_bench.json | cut -c1-600
sh tools/ncu_job.sh
