"""Print the interesting columns of an `ncu --page raw --csv` file, one block per kernel launch."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_op_ldgsts.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ldgsts.sum",
        "launch__grid_size", "launch__block_size", "smsp__warps_eligible.avg.per_cycle_active", "local_load", "local_store"]
stall = [h for h in hdr if "average_warp" in h and "issue_stalled" in h and h.endswith("_ratio") ] or [h for h in hdr if "warp_issue_stalled" in h and "per_warp_active" in h]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("=" * 100)
    for w in want:
        for h in hdr:
            if h == w or (w in h and w in ("local_load", "local_store")):
                print("%-75s %s %s" % (h, d[h], units[hdr.index(h)]))
    st = sorted(((float(d[h].replace(",", "")), h) for h in stall if d[h] not in ("", "n/a")), reverse=True)[:8]
    for v, h in st:
        print("   stall %-70s %.2f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("smsp__average_warp_latency_issue_stalled_", ""), v))
