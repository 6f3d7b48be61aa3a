#!/bin/bash
# Round-2 profiling pass (one GPU, under gpurun): every ncu command runs only after the same command
# exited 0 without ncu.  Outputs land in gpurun_out/; summaries are copied into profiles/ afterwards.
set -u
O=gpurun_out
run() { "$@" > /dev/null 2>&1 || { echo "plain run failed: $*"; return 1; }; }
# 1. launch list of the bench command (cold-cache, serialised times: shares, not absolutes)
run python bench.py --steps 2 --warmup 1 --no-cpu-baseline &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/r02_launches_bench.csv \
      python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/r02_ncu_bench.log 2>&1
# 2. the dominant kernel at config-3 size
run python tools/prof_workload.py r3 2 &&
  ncu --set full --clock-control none --import-source on -k regex:k_scaff_score -s 1 -c 1 -o $O/r02_k_scaff_score \
      python tools/prof_workload.py r3 2 > $O/r02_ncu_r3.log 2>&1
# 3. G-profile at config-3 size
run python tools/prof_workload.py g3 2 &&
  ncu --set full --clock-control none --import-source on -k regex:k_scaffg2r_score -s 1 -c 1 -o $O/r02_k_scaffg2_score \
      python tools/prof_workload.py g3 2 > $O/r02_ncu_g3.log 2>&1
# 4. streaming regime, one and two orderings
run python tools/prof_workload.py s1 5 &&
  ncu --set full --clock-control none --import-source on -k regex:k_scaff_stream1 -s 3 -c 1 -o $O/r02_k_scaff_stream1 \
      python tools/prof_workload.py s1 5 > $O/r02_ncu_s1.log 2>&1
S1_P=2 run python tools/prof_workload.py s1 5 &&
  S1_P=2 ncu --set full --clock-control none --import-source on -k regex:k_scaff_stream1p -s 3 -c 1 -o $O/r02_k_scaff_stream1p_p2 \
      python tools/prof_workload.py s1 5 > $O/r02_ncu_s1p.log 2>&1
# 5. GA generation on the pair-table path: launch list (operators vs scoring)
run python tools/prof_workload.py ga 6 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_ga.csv \
      python tools/prof_workload.py ga 6 > $O/r02_ncu_ga.log 2>&1
ls -la $O/r02_*
