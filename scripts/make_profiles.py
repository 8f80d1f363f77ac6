#!/usr/bin/env python
"""Turn the .ncu-rep files brought back in gpurun_out/ into the text summaries committed under profiles/."""
import csv, glob, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
           "launch__occupancy_limit_shared_mem", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
           "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum", "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum.per_second",
           "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum.pct_of_peak_sustained_elapsed",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]


def summarise(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    if len(rows) < 3:
        return
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on  ({os.path.basename(rep)})\n")
        f.write("# one block per profiled launch; values as reported by `ncu -i ... --page raw --csv`\n\n")
        for r in rows[2:]:
            f.write(r[hdr.index("Kernel Name")][:150] + "\n")
            for m in METRICS:
                if m in hdr:
                    i = hdr.index(m)
                    f.write(f"    {m:72s} {r[i]:>18s} {units[i]}\n")
            f.write("\n")
    print("wrote", out)


if __name__ == "__main__":
    for rep in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "*.ncu-rep"))):
        name = os.path.splitext(os.path.basename(rep))[0]
        summarise(rep, os.path.join(ROOT, "profiles", f"ncu_{name}.txt"))
