# round-end style check of HEAD on one GPU: GPU tests, smoke(), both bench arms
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/verify_gputest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/verify_gputest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/verify_smoke.log 2>&1
python bench.py > gpurun_out/verify_bench.json 2> gpurun_out/verify_bench.err
python bench.py --impl reference > gpurun_out/verify_bench_ref.json 2> gpurun_out/verify_bench_ref.err
tail -3 gpurun_out/verify_gputest.log; cat gpurun_out/verify_smoke.log | tail -1; cat gpurun_out/verify_bench.json
