"""Summarise an .ncu-rep (read here, no GPU) into the handful of metrics DESIGN.md / bench.py quote.
    python profiles/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/<name>.txt"""
import csv, io, subprocess, sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__inst_executed_op_global_red.sum", "smsp__inst_executed_op_global_atom.sum",
        "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum"]
STALL = "smsp__average_warps_issue_stalled_"

def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print("kernel:", d.get("Kernel Name", ""), "| id", d.get("ID", ""))
        for k in KEYS:
            if k in d:
                print(f"  {k:62s} {d[k]:>16s} {units[hdr.index(k)]}")
        st = sorted(((float(d[k]), k) for k in hdr if k.startswith(STALL) and k.endswith("_per_issue_active.ratio")
                     and "not_issued" not in k), reverse=True)[:6]
        for v, k in st:
            print(f"  stall {k[len(STALL):-len('_per_issue_active.ratio')]:28s} {v:8.3f} warps/issue")

if __name__ == "__main__":
    main(sys.argv[1])
