#!/bin/bash
# Round-2 profile captures (run on the GPU box through gpurun; outputs under gpurun_out/).  Each ncu pass runs only after the
# same command has exited 0 without ncu.
set -u
O=gpurun_out
B="python bench.py --transitions 4194304 --steps 2 --warmup 3 --no-extras --no-check"
$B > $O/r2_plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/r2_launches_n4M.csv $B > $O/r2_ncu_launches.log 2>&1
echo launches_rc=$?
S="python bench.py --scan-only"
$S > $O/r2_plain_scan.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_chunk_kernel -s 6 -c 1 -o $O/r2_scan $S > $O/r2_ncu_scan.log 2>&1
echo scan_rc=$?
A="python tools/run_actor.py 4"
$A > $O/r2_plain_actor.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:actor_tc_kernel -s 2 -c 1 -o $O/r2_actor $A > $O/r2_ncu_actor.log 2>&1
echo actor_rc=$?
V="python bench.py --transitions 1048576 --steps 1 --warmup 1 --no-extras --no-check"
$V > $O/r2_plain_value.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mlp_tc_kernel -s 20 -c 1 -o $O/r2_value $V > $O/r2_ncu_value.log 2>&1
echo value_rc=$?
ls -la $O/*.ncu-rep
