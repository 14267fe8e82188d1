# A/B of environment knobs on the C2 frame at 16 spp (value = Mrays/s, frame ms).  Usage: bash tools/quick_sweep.sh "VAR=val VAR2=val" ...
for cfg in "$@"; do
  out=$(env $cfg python bench.py --spp 16 --steps 4 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-fast-rng --no-also 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],2), d['config']['rank0_stage_ms'])")
  echo "$cfg -> $out"
done
