#!/bin/bash
# ncu --set full of the finest-level float tile kernel launches (regex $1, output name $2)
RE=${1:-k_apply_tile_v5}; OUT=${2:-r2_v5}
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$RE -c 5 -f -o gpurun_out/$OUT python tools/prof_v5.py > gpurun_out/${OUT}.log 2>&1
tail -2 gpurun_out/${OUT}.log
ncu -i gpurun_out/$OUT.ncu-rep --page raw --csv > gpurun_out/${OUT}_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${OUT}_raw.csv > gpurun_out/${OUT}_summary.txt
python - <<PY >> gpurun_out/${OUT}_summary.txt
import csv
rows=list(csv.reader(open("gpurun_out/${OUT}_raw.csv")))
h=rows[0]; idx={k:i for i,k in enumerate(h)}
more=["l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
"l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
"smsp__sass_l1tex_data_pipe_lsu_wavefronts_mem_shared_op_ldgsts.sum","l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
"l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ldgsts.sum","smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
"smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio","smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
"smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio","smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
"smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio","smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
"smsp__average_warps_issue_stalled_wait_per_issue_active.ratio","smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
"dram__throughput.avg.pct_of_peak_sustained_elapsed","launch__occupancy_limit_warps","sm__maximum_warps_per_active_cycle_pct","l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum","l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum","lts__t_sectors_srcunit_tex_op_read.sum"]
for r in rows[2:]:
    print("---- more counters:", r[idx["Kernel Name"]][:80])
    for k in more:
        if k in idx: print("  %-82s %s"%(k,r[idx[k]]))
PY
for id in 1 3; do
ncu -i gpurun_out/$OUT.ncu-rep --page source --csv --print-kernel-base function --launch-skip $id --launch-count 1 > gpurun_out/${OUT}_src$id.csv 2>/dev/null
echo "==== stalls launch $id" >> gpurun_out/${OUT}_summary.txt
python tools/ncu_stalls.py gpurun_out/${OUT}_src$id.csv 30 >> gpurun_out/${OUT}_summary.txt 2>&1
done
grep -v '^----$' gpurun_out/${OUT}_summary.txt | awk '/Kernel Name/{k++} k<=1 || /more counters/{m++} m>=1' | head -20; sed -n '/==== stalls launch 1/,\$p' gpurun_out/${OUT}_summary.txt | head -120
