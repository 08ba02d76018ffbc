"""Prints the metrics worth reading from an .ncu-rep (run in the build container: `python tests/ncu_summary.py rep`)."""
import csv
import subprocess
import sys

WANT = """gpu__time_duration.sum launch__grid_size launch__registers_per_thread launch__occupancy_limit_shared_mem
launch__occupancy_limit_registers sm__warps_active.avg.pct_of_peak_sustained_active
smsp__issue_active.avg.pct_of_peak_sustained_active sm__throughput.avg.pct_of_peak_sustained_elapsed
sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
dram__bytes_read.sum dram__bytes_write.sum gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed
lts__t_sector_hit_rate.pct lts__throughput.avg.pct_of_peak_sustained_elapsed smsp__inst_executed.sum
smsp__thread_inst_executed_per_inst_executed.ratio l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum
l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum l1tex__t_requests_pipe_lsu_mem_global_op_st.sum
l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum""".split()


def main(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    names = [r[hdr.index("Kernel Name")] for r in rows[2:]]
    print("kernels:", names)
    cols = WANT + [h for h in hdr if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    for w in cols:
        if w in hdr:
            i = hdr.index(w)
            vals = [r[i] for r in rows[2:]]
            if all(v in ("0", "0.000000", "") for v in vals):
                continue
            print(f"{w.replace('smsp__average_warps_issue_stalled_', 'stall:').replace('_per_issue_active.ratio', ''):70s} {vals} {rows[1][i]}")


if __name__ == "__main__":
    main(sys.argv[1])
