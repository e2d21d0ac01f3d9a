"""Per-kernel summary of an .ncu-rep (time, DRAM bytes, tensor / issue activity) as a table."""
import csv
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
           "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv", "--metrics", ",".join(METRICS)],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
print("kernel | us | dram_rd_MB | dram_wr_MB | regs | tensor% | sm% | issue% | dram%")
for r in rows[2:]:
    name = r[ix["Kernel Name"]].replace("void ", "").replace("bqa::<unnamed>::", "").replace("unnamed>::", "")
    name = name.split("(")[0][:44]
    vals = [r[ix[m]] if m in ix else "" for m in METRICS]
    print(name, "|", " | ".join(f"{float(v):.1f}" if v else "-" for v in vals))
