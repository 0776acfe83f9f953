# round-2 ncu evidence (run under gpurun after the plain bench has exited 0)
A="--no-cpu-baseline --no-e2e --no-parity --sustained-s 0"
python bench.py --steps 5 --warmup 3 $A > gpurun_out/r02_ncu_plain.json 2> gpurun_out/r02_ncu_plain.err || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv \
    --log-file gpurun_out/r02e_launches_bench_128cubed.csv python bench.py --steps 5 --warmup 3 $A > gpurun_out/r02_ncu1.log 2>&1
python bench.py --n 64 --steps 2 --warmup 1 $A > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:weno_cell_kernel -s 1 -c 1 -f -o gpurun_out/r02e_full_64cubed \
    python bench.py --n 64 --steps 2 --warmup 1 $A > gpurun_out/r02_ncu2.log 2>&1
