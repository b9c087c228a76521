"""Condense an .ncu-rep (ncu --set full) into one CSV row per kernel launch with the metrics the
roofline discussion uses.  Usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.csv"""
import csv
import io
import subprocess
import sys

KEEP = [
    "Kernel Name", "Grid Size", "Block Size",
    "gpu__time_duration.sum",
    "sm__cycles_elapsed.max",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
]


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(k) for k in KEEP if k in hdr]
    w = csv.writer(sys.stdout)
    w.writerow([hdr[i] + (f" [{units[i]}]" if units[i] else "") for i in idx])
    for r in rows[2:]:
        w.writerow([r[i][:90] for i in idx])


if __name__ == "__main__":
    main(sys.argv[1])
