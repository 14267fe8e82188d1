# Whole-frame physical traffic of the trace kernel (roofline.traffic): tools/trace_traffic.py.
# Usage: gpurun -- bash tools/prof_traffic.sh TAG WORKLOAD SPP
TAG=${1:-r02b}; WL=${2:-C2}; SPP=${3:-4}
set -x
python tools/trace_traffic.py run $WL $SPP gpurun_out/${TAG}_trace_rays_$WL.json > gpurun_out/traffic_plain.log 2>&1 &&
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum \
    --clock-control none -k regex:k_pt_trace --csv --log-file gpurun_out/${TAG}_trace_metrics_$WL.csv \
    python tools/trace_traffic.py run $WL $SPP gpurun_out/${TAG}_trace_rays_$WL.json > gpurun_out/traffic_ncu.log 2>&1
tail -2 gpurun_out/traffic_ncu.log
