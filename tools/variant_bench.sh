# usage: bash tools/variant_bench.sh  (runs bench.py against every variants/libgpt_*.so)
for f in variants/libgpt_*.so; do
  GPT_LIB=$PWD/$f python bench.py --steps 3 --warmup 2 --spp 8 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$f', round(d['value'],1), round(d['ms_per_step'],2), d['config']['rank0_stage_ms'])"
done
