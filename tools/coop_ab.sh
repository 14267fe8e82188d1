# A/B of the cooperative-tail variants: full frame and 1/8-frame shard of C2 at 16 spp
export PYTHONPATH=.
python -m pytest tests/test_gpu_render.py -m gpu -x -q 2>&1 | tail -2 | tee gpurun_out/coop_ab.log
for f in "" $(ls variants/libgpt_*.so 2>/dev/null); do
  for c in 0 1; do
    if [ -z "$f" ]; then lib=$PWD/gpu_pathtracer_b200/csrc/libgpt.so; else lib=$PWD/$f; [ $c = 0 ] && continue; fi
    echo "== $lib COOP=$c" | tee -a gpurun_out/coop_ab.log
    GPT_LIB=$lib GPT_TRACE_COOP=$c python tools/shard_time.py --nranks 8 --spp 16 --only-rank0 | grep "stage_timing\": false" | cut -c1-90 | tee -a gpurun_out/coop_ab.log
  done
done
