# Round-2 profile commands (run on the GPU box through gpurun; outputs in gpurun_out/, summarised into profiles/ by
# tools/summarise_profiles.py).  Plain run first, then the ncu launch list, then one full capture of the chained kernel.
set -x
python bench.py > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err; tail -c 400 gpurun_out/bench_r2.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r2_ref.json 2>> gpurun_out/bench_r2.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-scaling-base > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:chain_kernel -c 1 -f -o gpurun_out/prof_r2_chain python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-scaling-base > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -6
