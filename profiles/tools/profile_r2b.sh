# end-of-round-2 profiling session (one gpurun call): ncu --set full captures of the two fused
# instantiations that carry the numbers + launch lists; every command first exits 0 without ncu.
set -x
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-configs --no-transport"
C3="python profiles/tools/bench_configs.py --configs 3 --backends B200F --steps 3"
O=gpurun_out
$B > $O/r2q_bench_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_fused_step -s 12 -c 2 -f -o $O/r2q_fused_golovin $B > $O/r2q_ncu1.log 2>&1
$B > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r2q_bench_launches.csv $B > $O/r2q_ncu1b.log 2>&1
$C3 > $O/r2q_c3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_fused_step -s 12 -c 3 -f -o $O/r2q_fused_adaptive_geometric $C3 > $O/r2q_ncu2.log 2>&1
$C3 > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2q_c3_launches.csv $C3 > $O/r2q_ncu2b.log 2>&1
ls -la $O/*.ncu-rep
tail -n 2 $O/r2q_ncu1.log; tail -n 2 $O/r2q_ncu2.log
