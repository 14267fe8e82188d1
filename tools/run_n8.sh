# One 8-GPU box: the default bench line (C2 + fast_rng + also C4), C4 as its own line, C5 at 16 spp and at full size.
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
set -x
$TR --master-port 29531 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02n_bench_C2_N8.json 2> gpurun_out/r02n_bench_C2_N8.err; echo rc=$?
$TR --master-port 29532 bench.py --gpus 8 --workload C4 --steps 8 --warmup 3 > gpurun_out/r02n_bench_C4_N8.json 2> gpurun_out/r02n_bench_C4_N8.err; echo rc=$?
$TR --master-port 29533 bench.py --gpus 8 --workload C5 --spp 16 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02n_bench_C5_N8_spp16.json 2> gpurun_out/r02n_bench_C5_N8_spp16.err; echo rc=$?
$TR --master-port 29534 bench.py --gpus 8 --workload C5 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02n_bench_C5_N8.json 2> gpurun_out/r02n_bench_C5_N8.err; echo rc=$?
for f in gpurun_out/r02n_bench_*.json; do python - $f <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1], "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1),
          "fast", (d["config"].get("fast_rng") or {}).get("value"), "also", (d.get("also") or {}).get("C4", {}).get("pass_ms"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
tail -3 gpurun_out/r02n_bench_*.err
