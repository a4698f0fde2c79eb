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
# router hash kernels on one rank's 65,536-request shard (launches 0-5 are the 8 K-request shard's)
S="python tools/bench_route_shard.py --sizes 8192,65536 --steps 2"
$N -k regex:rm_route_finish_kernel --launch-skip 6 --launch-count 1 -o ${out}_route_finish -f $S > ${out}_ncu6.log 2>&1
$N -k regex:rm_route_render_kernel --launch-skip 6 --launch-count 1 -o ${out}_route_render -f $S > ${out}_ncu7.log 2>&1
if [ "$what" = "all" ]; then
  $N -k 'regex:rm_split_kernel' -s 4 -c 1 -o ${out}_k3_c2 $B --no-blocks --no-cpu > ${out}_ncu2.log 2>&1
  $N -k 'regex:rm_insert_kernel<\(int\)32, \(int\)0' -s 200 -c 1 -o ${out}_k4 $B > ${out}_ncu4.log 2>&1
  # one rm_evict call (9 kernel launches) on the 1 M-node tree, after three warm calls
  $N -k regex:rm_evict_ --launch-skip 27 --launch-count 9 -o ${out}_k6 -f python tools/bench_evict.py --batches 3 > ${out}_ncu5.log 2>&1
fi
ls -la gpurun_out | grep r2_final
