"""Key metrics of an ncu report (--set full) as CSV: python tools/ncu_summary.py rep.ncu-rep [title]"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor_op_hmma.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_l1tex2xbar_write_bytes.sum", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_membar_per_warp_active.pct",
    "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct", "smsp__warp_issue_stalled_wait_per_warp_active.pct",
    "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct",
    # the names this ncu (2025.2, sm_100) actually reports: tensor pipe (tcgen05 = the hmma sub-pipe) and warp stalls
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.sum.per_second",
    "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
] + ["smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % r for r in (
    "long_scoreboard", "short_scoreboard", "barrier", "membar", "wait", "sleeping", "lg_throttle", "mio_throttle",
    "math_pipe_throttle", "tex_throttle", "dispatch_stall", "branch_resolving", "no_instruction", "not_selected",
    "selected", "drain", "misc")]
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
for h, i in list(idx.items()):      # some columns carry a section prefix ("TPC.TriageCompute.sm__pipe_tensor_...")
    idx.setdefault(h.split(".Triage")[-1].split(".", 1)[-1] if ".Triage" in h else h, i)
print("# " + (sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]))
names = [r[idx["Kernel Name"]].split("(")[0].replace(", ", " ") for r in data]
print("metric,unit," + ",".join("%s#%d" % (n, i) for i, n in enumerate(names)))
for w in dict.fromkeys(WANT):
    if w in idx:
        print("%s,%s,%s" % (w, units[idx[w]], ",".join(r[idx[w]].replace(",", "") for r in data)))
# derived: tensor-pipe utilisation = cycles the (four per SM) tensor sub-pipes were active / (4 x elapsed SM cycles).
# Cross-check on fb::dwpw_kernel: 0.455 x (148 SMs x 8192 dense bf16 flop/clk x 1.676 GHz) = 925 TFLOP/s =
# the executed 3 x 54.3 GFLOP / 175 us.  (ncu's own pct_of_peak for this counter reads ~2 % for every
# tcgen05 kernel: its denominator is not the tcgen05 rate.)
ha, ce = "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "sm__cycles_elapsed.max"
if ha in idx and ce in idx:
    f = lambda v: float(v.replace(",", ""))
    print("derived__tensor_pipe_utilisation,fraction,%s" % ",".join(
        "%.4f" % (f(r[idx[ha]]) / (4.0 * f(r[idx[ce]]))) for r in data))
    if "dram__bytes_read.sum" in idx:
        ub = {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}
        ut = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}
        gb = lambda r, m: f(r[idx[m]]) * ub[units[idx[m]]]      # each column carries its own unit
        print("derived__dram_GBps,GB/s,%s" % ",".join("%.1f" % (
            (gb(r, "dram__bytes_read.sum") + gb(r, "dram__bytes_write.sum")) /
            (f(r[idx["gpu__time_duration.sum"]]) * ut[units[idx["gpu__time_duration.sum"]]])) for r in data))
