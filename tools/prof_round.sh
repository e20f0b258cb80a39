set -x
python bench.py > gpurun_out/bench_r1d.json 2> gpurun_out/bench_r1d.err; tail -c 600 gpurun_out/bench_r1d.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r1d_ref.json 2>> gpurun_out/bench_r1d.err; tail -c 400 gpurun_out/bench_r1d_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1d.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-scaling-base > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:wf_kernel -c 3 -f -o gpurun_out/prof_r1d_wf python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-scaling-base > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -5
