# usage: bash tools/variant_bench.sh  (runs bench.py against the normal build and every variants/libgpt_*.so)
for f in gpu_pathtracer_b200/csrc/libgpt.so variants/libgpt_*.so; do
  GPT_LIB=$PWD/$f python bench.py --steps 3 --warmup 3 --spp 8 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read()); c=d['config']; print('$f', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms', {k: round(v,2) for k,v in c['rank0_stage_ms'].items()}, 'V_int', round(c['avg_internal_visits_per_ray'],2), 'V_leaf', round(c['avg_leaf_tests_per_ray'],2))"
done
