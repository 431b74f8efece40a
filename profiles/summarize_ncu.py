"""Key metrics of an `ncu --set full` report as text (what profiles/*_summary.txt files hold).
Usage: python profiles/summarize_ncu.py REPORT.ncu-rep [launch index] > profiles/NAME_summary.txt"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes.sum.per_second",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__grid_size", "launch__block_size", "launch__cluster_size", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max"]


def main():
    rep = sys.argv[1]
    which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2 + which]
    name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print(f"-- launch {which}: {name}")
    stalls = []
    for h, u, v in zip(hdr, units, vals):
        if h in WANT:
            print(f"{h:70s} {u:18s} {v}")
        if "smsp__average_warps_issue_stalled" in h and "not_issued" not in h:
            try:
                stalls.append((float(v), h.split("stalled_")[1].split("_per")[0]))
            except ValueError:
                pass
    tot = sum(v for v, _ in stalls) or 1.0
    print("warp stall samples: " + ", ".join(f"{n} {100 * v / tot:.1f}%" for v, n in sorted(stalls, reverse=True)[:8]))


if __name__ == "__main__":
    main()
