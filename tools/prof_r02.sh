# r02 baseline: GPU tests, the default bench line, the ncu launch list and full captures of the two
# top kernels (each ncu pass only after the plain command exited 0).  Usage: gpurun -- bash tools/prof_r02.sh TAG
TAG=${1:-r02a}
set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/${TAG}_gputests.log 2>&1; tail -5 gpurun_out/${TAG}_gputests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_C2.json 2> gpurun_out/${TAG}_bench_C2.err; tail -c 3000 gpurun_out/${TAG}_bench_C2.json
CMD="python bench.py --steps 2 --warmup 1 --spp 4 --no-cpu-baseline --no-ref-cuda --no-fast-rng --no-also"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_launches_bench_C2_spp4.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_pt_trace -s 6 -c 3 -f -o gpurun_out/${TAG}_k_pt_trace $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_pt_shade -s 6 -c 3 -f -o gpurun_out/${TAG}_k_pt_shade $CMD > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/plain.log; tail -3 gpurun_out/ncu2.log; ls -la gpurun_out
