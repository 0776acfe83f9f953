# same-box A/B of library variants at N=1: tools/ab_n1.sh tag=ENV... ; results in gpurun_out/ab1_<tag>.json
for spec in "$@"; do
  tag=${spec%%=*}; envs=$(echo "${spec#*=}" | tr ',' ' ')
  env $envs python bench.py --no-cpu-baseline --no-e2e --no-parity --sustained-s 3 --steps 20 $BENCH_ARGS > gpurun_out/ab1_$tag.json 2> gpurun_out/ab1_$tag.err
done
