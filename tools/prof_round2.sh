# Round-2 profiling pass (run under gpurun, one GPU): ncu --set full captures of the hot kernels of every bench block and
# the launch list of the default bench command.  Outputs under gpurun_out/; the summaries are copied to profiles/.
out=gpurun_out/r2_final
B="python bench.py --steps 3 --warmup 3"
$B > ${out}_plain.log 2>&1 || { echo "plain run failed"; tail -5 ${out}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ${out}_launches.csv $B > ${out}_launches.log 2>&1
N="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
# C2 walk (after 18 insert pre-passes use the <32,...> instantiation), split stage
$N -k 'regex:rm_match_rung_kernel<32, true' -s 4 -c 1 -o ${out}_k1_c2 $B --no-blocks --no-cpu > ${out}_ncu1.log 2>&1
$N -k 'regex:rm_split_kernel' -s 4 -c 1 -o ${out}_k3_c2 $B --no-blocks --no-cpu > ${out}_ncu2.log 2>&1
# target walk: the same kernel name, launches after the C2 part (3+3 device + 3+3 host-buffer steps = 12) and the mesh block's none
$N -k 'regex:rm_match_rung_kernel<32, true' -s 16 -c 1 -o ${out}_k1_target $B > ${out}_ncu3.log 2>&1
# insert pipeline, evict, route kernels (any launch in steady state)
$N -k 'regex:rm_insert_kernel<32, 0' -s 200 -c 1 -o ${out}_k4 $B > ${out}_ncu4.log 2>&1
$N -k 'regex:rm_evict_(keys|climb|apply|count)_kernel' -s 4 -c 4 -o ${out}_k6 $B > ${out}_ncu5.log 2>&1
$N -k 'regex:rm_route_(render|finish)_kernel' -s 4 -c 2 -o ${out}_k8 $B > ${out}_ncu6.log 2>&1
ls -la gpurun_out | grep r2_final
