# Round-2 profiling pass (run under gpurun, one GPU): ncu --set full captures of the hot kernels of every bench block and
# the launch list of the default bench command.  Outputs under gpurun_out/; tools/ncu_summary.py condenses them into
# profiles/r2_*.csv.   Usage: bash tools/prof_round2.sh [all|k1]
out=gpurun_out/r2_final
what=${1:-all}
B="python bench.py --steps 3 --warmup 3"
N="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
$B > ${out}_plain.log 2>&1 || { echo "plain run failed"; tail -5 ${out}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ${out}_launches.csv $B > ${out}_launches.log 2>&1
# C2 walk: the <32, true, ...> instantiation (the <8, ...> launches are the insert pre-pass), 5th match batch
$N -k 'regex:rm_match_rung_kernel<\(int\)32, \(bool\)1' -s 4 -c 1 -o ${out}_k1_c2 $B --no-blocks --no-cpu > ${out}_ncu1.log 2>&1
# target walk: same kernel, after the C2 part's 6 device + 6 host-buffer batches and 3 warm-up batches of the block
$N -k 'regex:rm_match_rung_kernel<\(int\)32, \(bool\)1' -s 16 -c 1 -o ${out}_k1_target $B > ${out}_ncu3.log 2>&1
$N -k 'regex:rm_route_(render|finish)_kernel' -s 4 -c 2 -o ${out}_k8 $B > ${out}_ncu6.log 2>&1
if [ "$what" = "all" ]; then
  $N -k 'regex:rm_split_kernel' -s 4 -c 1 -o ${out}_k3_c2 $B --no-blocks --no-cpu > ${out}_ncu2.log 2>&1
  $N -k 'regex:rm_insert_kernel<\(int\)32, \(int\)0' -s 200 -c 1 -o ${out}_k4 $B > ${out}_ncu4.log 2>&1
  $N -k 'regex:rm_evict_(keys|climb|apply|count|sel_hist|sel_compact)_kernel' -s 8 -c 8 -o ${out}_k6 $B > ${out}_ncu5.log 2>&1
fi
ls -la gpurun_out | grep r2_final
