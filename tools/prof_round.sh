# Profile commands of a round (run on the GPU box through gpurun; outputs in gpurun_out/, summarised into profiles/ by
# tools/summarise_profiles.py <tag> <suffix>).  Plain run first, then the ncu launch list, then one full capture of the
# chained kernel.  usage: bash tools/prof_round.sh [suffix = r2b]
S=${1:-r2b}
set -x
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_$S.json 2> gpurun_out/bench_$S.err; tail -c 400 gpurun_out/bench_$S.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$S.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-scaling-base > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:chain_kernel -c 1 -f -o gpurun_out/prof_${S}_chain python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-scaling-base > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -6
