set -x
CMD="python bench.py --steps 2 --warmup 1 --spp 4 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01h_launches_bench_C2_spp4.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_pt_trace -s 4 -c 3 -f -o gpurun_out/r01h_k_pt_trace $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_pt_shade -s 4 -c 3 -f -o gpurun_out/r01h_k_pt_shade $CMD > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/plain.log; tail -3 gpurun_out/ncu2.log; ls -la gpurun_out
