# usage: bash tools/quick_ab.sh [spp]  -- bench.py on the normal build (wide walk, then binary walk) and every variants/libgpt_*.so
SPP=${1:-8}
one() {
  python bench.py --steps 3 --warmup 3 --spp $SPP --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']
print('$1', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms', {k: round(v,2) for k,v in c['rank0_stage_ms'].items()}, 'V_int', round(c['avg_internal_visits_per_ray'],2), 'V_leaf', round(c['avg_leaf_tests_per_ray'],2), 'build', round(c['lbvh_build_ms'],3), 'all', round(c['build_all_ms'],2))"
}
one wide
GPT_TRACE_WIDE=0 one binary
for f in variants/libgpt_*.so; do [ -f $f ] && GPT_LIB=$PWD/$f one $f; done
true
