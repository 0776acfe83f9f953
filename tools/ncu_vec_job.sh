# ncu --set full capture of the component-lane kernel at 64^3 vector
A="--no-cpu-baseline --no-e2e --no-parity --sustained-s 0"
python bench.py --n 64 --steps 2 --warmup 1 $A > gpurun_out/vec_ncu_plain.json 2> gpurun_out/vec_ncu_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:weno_cell_kernel -s 1 -c 1 -f -o gpurun_out/r02_vec_full_64cubed \
    python bench.py --n 64 --steps 2 --warmup 1 $A > gpurun_out/vec_ncu2.log 2>&1
tail -2 gpurun_out/vec_ncu2.log
