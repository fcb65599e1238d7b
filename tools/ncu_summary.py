"""Text summary of an .ncu-rep (read in the build container): the counters DESIGN.md / profiles/README.md quote, per kernel.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [--source] > profiles/x_ncu.txt"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
    "smsp__average_warp_latency_issue_stalled_barrier.pct", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg",
]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units = rows[0], rows[1]
    return head, units, rows[2:]


def main():
    rep = sys.argv[1]
    head, units, rows = raw(rep)
    print(f"# {rep}: ncu --set full --clock-control none --import-source on (counters per launch; times are under the profiler)")
    idx = {k: i for i, k in enumerate(head)}
    for r in rows:
        print(f"\n== {r[idx['Kernel Name']]}   grid {r[idx['Grid Size']]} block {r[idx['Block Size']]}")
        for k in KEYS:
            if k in idx and r[idx[k]] != "":
                print(f"  {k:86s} {r[idx[k]]:>18s} {units[idx[k]]}")
    if "--source" in sys.argv:
        out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True,
                             text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))

        def num(v):
            try:
                return float(v.replace(",", ""))
            except ValueError:
                return 0.0
        # per function: rows with a line number carry the totals of that CUDA source line (the SASS rows under it are skipped)
        fn, head, lines = None, None, []

        def flush():
            if not fn or not lines:
                return
            c_ins, c_smp = head.index("Instructions Executed"), head.index("# Samples")
            print(f"\n-- {fn[:140]}\n   hottest source lines by executed instructions (warp-level) | stall samples")
            for r in sorted(lines, key=lambda r: -num(r[c_ins]))[:22]:
                print(f"  {num(r[c_ins]):13.0f} | {num(r[c_smp]):7.0f}  L{r[0]:>4s}  {r[1].strip()[:140]}")
            print("   hottest source lines by stall samples")
            for r in sorted(lines, key=lambda r: -num(r[c_smp]))[:12]:
                print(f"  {num(r[c_smp]):7.0f}  L{r[0]:>4s}  {r[1].strip()[:150]}")
        for r in rows:
            if r and r[0] == "Function Name":
                flush()
                fn, head, lines = r[1], None, []
            elif r and r[0] == "Line No":
                head = r
            elif head and len(r) >= len(head) - 1 and r[0].strip().isdigit():
                lines.append(r)
        flush()


if __name__ == "__main__":
    main()
