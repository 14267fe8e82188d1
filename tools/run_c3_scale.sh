# C3 (nested glass + DoF) tile-sharded at 4 and 8 GPUs on one 8-GPU box.
for N in 8 4; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29560 + N)) bench.py --gpus $N --workload C3 --steps 8 --warmup 3 --no-also > gpurun_out/r02L_bench_C3_N$N.json 2> gpurun_out/r02L_bench_C3_N$N.err
  echo "N=$N rc=$?"
  python - gpurun_out/r02L_bench_C3_N$N.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(sys.argv[1], round(d["value"], 1), round(d["ms_per_step"], 2), round(d["e2e"]["value"], 1), (d["config"].get("fast_rng") or {}).get("value"), d.get("parity", {}).get("ok"), d.get("parity", {}).get("frac_within_1e-3"))
PY
done
