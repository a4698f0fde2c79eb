# A/B of the walk / split options + ncu captures of K1 and K3 (run under gpurun; outputs in gpurun_out/)
out=gpurun_out/r2_t4
summ() { python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('ms/step', round(d['ms_per_step'],5), d.get('stages_ms_per_step'), 'e2e', round(d['e2e']['value']/1e6,2), 'launches', d['gpu_launches'])
    else: print(l.rstrip()[-300:])
"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > ${out}_tests.log
cat ${out}_tests.log
for cfg in "RM_RUN_STAGE=0" "RM_RUN_STAGE=0 RM_SPLIT_ROLES=0" "RM_RUN_STAGE=1"; do
  echo "== $cfg" >> ${out}_ab.log
  env $cfg python bench.py --steps 20 --warmup 5 --stages --no-cpu 2>/dev/null | summ >> ${out}_ab.log 2>&1
done
for wl in target c5r; do
  echo "== $wl STAGE=0" >> ${out}_ab.log
  RM_RUN_STAGE=0 python bench.py --steps 10 --warmup 3 --stages --no-cpu --workload $wl 2>/dev/null | summ >> ${out}_ab.log 2>&1
done
cat ${out}_ab.log
export RM_RUN_STAGE=0
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:rm_match_rung_kernel<\(int\)32' -s 4 -c 1 -o ${out}_k1 python bench.py --steps 3 --warmup 3 --no-cpu > ${out}_ncu_k1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:rm_split_kernel -s 4 -c 1 -o ${out}_k3 python bench.py --steps 3 --warmup 3 --no-cpu > ${out}_ncu_k3.log 2>&1
tail -3 ${out}_ncu_k1.log
ls -la gpurun_out/ | tail -8
