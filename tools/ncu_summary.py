"""Text summary of an .ncu-rep (raw page): python tools/ncu_summary.py rep.ncu-rep > profiles/x.txt"""
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "sm__inst_executed_pipe_lsu.sum",
    "smsp__cycles_active.avg", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum", "lts__t_requests_srcunit_tex_op_red.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio" ,
]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
print(f"# {sys.argv[1]}: ncu --set full --clock-control none (per launch; cold-cache, serialised)")
for r in rows[2:]:
    print(r[hdr.index("Kernel Name")], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
    for w in WANT:
        if w in hdr:
            print(f"    {w:72s} {r[hdr.index(w)]:>16s} {units[hdr.index(w)]}")
    if "dram__bytes_read.sum" in hdr:
        def val(n):
            v, u = float(r[hdr.index(n)]), units[hdr.index(n)]
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        print(f"    traffic (dram read + write)  {val('dram__bytes_read.sum') + val('dram__bytes_write.sum'):.0f} B")
