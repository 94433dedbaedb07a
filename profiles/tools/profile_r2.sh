# round-2 profiling session (one gpurun call): ncu --set full captures of the kernels of the
# round + launch lists of the workloads; every command first exits 0 without ncu.
set -x
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-configs"
C3="python profiles/tools/bench_configs.py --configs 3 --backends B200F --steps 3"
C2="python profiles/tools/bench_configs.py --configs 2 --backends B200F --steps 3"
C01="python profiles/tools/bench_configs.py --configs 0,1 --backends B200F --steps 3"
O=gpurun_out
$B > $O/r2p_bench_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_fused_step -s 12 -c 2 -f -o $O/r2p_fused_golovin $B > $O/r2p_ncu1.log 2>&1
$B > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r2p_bench_launches.csv $B > $O/r2p_ncu1b.log 2>&1
$C3 > $O/r2p_c3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_fused_step -s 12 -c 3 -f -o $O/r2p_fused_adaptive_geometric $C3 > $O/r2p_ncu2.log 2>&1
$C3 > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2p_c3_launches.csv $C3 > $O/r2p_ncu2b.log 2>&1
$C2 > $O/r2p_c2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:k_stream_step|k_fused_breakup|k_shuffle_walk' -s 9 -c 5 -f -o $O/r2p_stream_breakup $C2 > $O/r2p_ncu3.log 2>&1
$C2 > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2p_c2_launches.csv $C2 > $O/r2p_ncu3b.log 2>&1
$C01 > $O/r2p_c01_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2p_c01_launches.csv $C01 > $O/r2p_ncu4b.log 2>&1
ls -la $O/*.ncu-rep
tail -2 $O/r2p_ncu1.log $O/r2p_ncu2.log $O/r2p_ncu3.log
