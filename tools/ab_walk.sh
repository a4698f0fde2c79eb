set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_t1_tests.log
tail -3 gpurun_out/r2_t1_tests.log
for cfg in "" "RM_RUN_DEPTH=2" "RM_RUN_DEPTH=4" "RM_RUN_MINB=16" "RM_RUN_MINB=16 RM_RUN_DEPTH=2" "RM_RUN_STAGE=0"; do
  echo "== $cfg" >> gpurun_out/r2_t1_ab.log
  env $cfg python bench.py --steps 20 --warmup 5 --stages --no-cpu 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('ms/step', d['ms_per_step'], d.get('stages_ms_per_step'), 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
    else: print(l.rstrip()[-300:])
" >> gpurun_out/r2_t1_ab.log 2>&1
done
for wl in target c5r; do
  echo "== $wl" >> gpurun_out/r2_t1_ab.log
  python bench.py --steps 10 --warmup 3 --stages --no-cpu --workload $wl 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('ms/step', d['ms_per_step'], d.get('stages_ms_per_step'))
    else: print(l.rstrip()[-300:])
" >> gpurun_out/r2_t1_ab.log 2>&1
done
cat gpurun_out/r2_t1_ab.log
