out=gpurun_out/r2_final
summ() { python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('ms/step', round(d['ms_per_step'],5), d.get('stages_ms_per_step'))
"; }
for cfg in "RM_EXP_FLAGS=0" "RM_EXP_FLAGS=8" "RM_RUN_DEPTH=2" "RM_EXP_FLAGS=0"; do
  echo "== $cfg" >> ${out}_ab.log
  env $cfg python bench.py --steps 20 --warmup 5 --stages --no-cpu --no-blocks 2>/dev/null | summ >> ${out}_ab.log 2>&1
done
for cfg in "RM_EXP_FLAGS=0" "RM_EXP_FLAGS=8" "RM_RUN_DEPTH=2"; do
  echo "== target $cfg" >> ${out}_ab.log
  env $cfg python bench.py --steps 10 --warmup 3 --stages --no-cpu --no-blocks --workload target 2>/dev/null | summ >> ${out}_ab.log 2>&1
done
cat ${out}_ab.log
B="python bench.py --steps 3 --warmup 3"
N="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
$N -k 'regex:rm_match_rung_kernel<\(int\)32, \(bool\)1' -s 4 -c 1 -o ${out}_k1_c2 $B --no-blocks --no-cpu > ${out}_ncu1.log 2>&1
$N -k 'regex:rm_match_rung_kernel<\(int\)32, \(bool\)1' -s 16 -c 1 -o ${out}_k1_target $B > ${out}_ncu3.log 2>&1
$N -k 'regex:rm_insert_kernel<\(int\)32, \(int\)0' -s 200 -c 1 -o ${out}_k4 $B > ${out}_ncu4.log 2>&1
ls -la gpurun_out | grep "r2_final_k"
