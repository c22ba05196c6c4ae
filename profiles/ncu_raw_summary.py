import csv, sys, subprocess
rep=sys.argv[1]
out=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[0]; units=rows[1]
want=["gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed","smsp__inst_executed.sum","smsp__issue_active.avg.pct_of_peak_sustained_active","sm__warps_active.avg.pct_of_peak_sustained_active","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum","smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio","smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio","smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio","smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio","smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio","smsp__average_warps_issue_stalled_wait_per_issue_active.ratio","smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio","launch__registers_per_thread","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
      "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed","l1tex__data_pipe_lsu_wavefronts.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum","SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg","l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ldgsts.sum","l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum","l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum","smsp__inst_executed_op_ldgsts.sum","smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"]
idx={h:i for i,h in enumerate(hdr)}
for r in rows[2:]:
    print("-----", r[idx["Kernel Name"]][:50], r[idx["Grid Size"]])
    for w in want:
        if w in idx: print(f"  {w[:80]:80s} {r[idx[w]]:>18s} {units[idx[w]]}")
