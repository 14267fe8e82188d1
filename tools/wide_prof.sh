set -x
CMD="python bench.py --steps 2 --warmup 1 --spp 4 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_pt_trace_wide -s 6 -c 2 -f -o gpurun_out/r02e_k_pt_trace_wide $CMD > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
