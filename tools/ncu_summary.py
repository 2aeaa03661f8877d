"""Text summary of one `ncu --set full` report for profiles/: selected raw metrics of the first kernel in the report plus the
per-CUDA-line top list of tools/ncu_lines.py. Usage: python tools/ncu_summary.py <report.ncu-rep> "<header line>" > profiles/x.txt"""
import csv
import subprocess
import sys

rep, title = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
PREFIX = ("gpu__time", "dram__bytes", "dram__throughput", "launch__registers", "launch__occupancy", "launch__grid_size", "launch__block_size",
          "launch__shared_mem", "sm__warps_active", "sm__throughput", "l1tex__throughput", "lts__throughput", "sm__inst_executed_pipe_tex",
          "l1tex__t_sector_hit", "lts__t_sector_hit", "smsp__thread_inst_executed_per_inst", "sm__inst_issued.avg", "l1tex__m_xbar2l1tex_read_bytes.sum",
          "sm__cycles_active", "smsp__inst_executed.sum", "l1tex__texin_sm2tex_req_cycles_active.avg", "smsp__pcsamp_warps_issue_stalled",
          "l1tex__t_sectors_pipe_tex_mem_texture", "l1tex__t_requests_pipe_tex_mem_texture.sum", "lts__t_sectors_srcunit_tex.sum",
          "smsp__cycles_active.avg", "sm__maximum_warps")
print("# " + title)
print("# kernel: " + vals[hdr.index("Kernel Name")] + "  grid " + vals[hdr.index("Grid Size")] + " block " + vals[hdr.index("Block Size")])
print()
for k in hdr:
    if any(k.startswith(p) for p in PREFIX) and not k.endswith("_not_issued"):
        i = hdr.index(k)
        print(f"{k} [{units[i]}] = {vals[i]}")
print()
print("# per CUDA source line, sorted by stall samples (tools/ncu_lines.py)")
print(subprocess.run([sys.executable, "tools/ncu_lines.py", rep, "30"], capture_output=True, text=True).stdout)
