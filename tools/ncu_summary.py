"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`) into a small text file for profiles/.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/x_summary.txt
"""
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_read\.sum|dram__bytes_write\.sum|"
    r"dram__throughput\.avg\.pct_of_peak_sustained_elapsed|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"lts__t_bytes\.sum|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum(\.pct_of_peak_sustained_elapsed)?|"
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
    r"sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)|"
    r"sm__inst_executed_pipe_tensor.*\.avg\.pct_of_peak_sustained_active|"
    r"sm__pipe_tensor_subpipe_.*cycles_active.*pct.*|"
    r"smsp__inst_executed\.sum|sm__issue_active\.avg\.pct_of_peak_sustained_elapsed|"
    r"smsp__issue_active\.avg\.pct_of_peak_sustained_active|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed|sm__cycles_elapsed\.max|"
    r"launch__(grid_size|block_size|cluster_size|registers_per_thread|shared_mem_per_block_dynamic|waves_per_multiprocessor|occupancy_limit_.*)|"
    r"smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio)$")


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    lines = [f"# ncu --set full summary of {rep} (one block per captured launch)"]
    for r in rows[2:]:
        lines.append("")
        lines.append("kernel: " + r[hdr.index("Kernel Name")])
        for h, u, v in zip(hdr, units, r):
            if KEEP.match(h) and v not in ("", "0") or h in ("dram__bytes_write.sum",):
                if KEEP.match(h):
                    lines.append(f"  {h:88s} {v} {u}")
    with open(out, "w") as fh:
        fh.write("\n".join(lines) + "\n")
    print("\n".join(lines[:60]))


if __name__ == "__main__":
    main()
