"""Extracts the judged metrics from an ncu report (ncu -i X.ncu-rep --page raw --csv) into a small text table.
usage: python profiles/summarize.py gpurun_out/prof.ncu-rep > profiles/NAME.txt"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    print(f"# {path}: one block per profiled launch (ncu --set full --clock-control none)")
    for r in rows[2:]:
        print("\nkernel:", r[hdr.index("Kernel Name")])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:75s} {r[i]:>18s} {units[i]}")
        rd = float(r[hdr.index("dram__bytes_read.sum")].replace(",", ""))
        wr = float(r[hdr.index("dram__bytes_write.sum")].replace(",", ""))
        ur, uw = units[hdr.index("dram__bytes_read.sum")], units[hdr.index("dram__bytes_write.sum")]
        print(f"  traffic = dram read + write = {rd} {ur} + {wr} {uw}")


def hot_spots(path, top=12):
    """Per kernel, the instructions that collect the most warp-stall samples (ncu --page source, needs --import-source on)."""
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    kern, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "rows": []}
            kern.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None:
            cur["rows"].append(r)
    seen = set()
    for k in kern:
        if k["name"] in seen or not k["rows"]:
            continue
        seen.add(k["name"])
        h = k["hdr"]
        si, ie, src, lsb = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source"), h.index("stall_long_sb")
        tot = sum(int(r[si]) for r in k["rows"])
        toti = sum(int(r[ie]) for r in k["rows"])
        print(f"\nhot spots: {k['name'][:90]}\n  {len(k['rows'])} SASS instructions, {tot} warp samples, {toti} warp-instructions executed")
        idx = sorted(range(len(k["rows"])), key=lambda i: -int(k["rows"][i][si]))[:top]
        for i in sorted(idx):
            r = k["rows"][i]
            print(f"  #{i:5d} {100 * int(r[si]) / max(tot, 1):5.1f} % of samples ({int(r[lsb])} long-scoreboard)  executed {int(r[ie]):9d}  {r[src].strip()[:90]}")


if __name__ == "__main__":
    main(sys.argv[1])
    if len(sys.argv) > 2 and sys.argv[2] == "--hot":
        hot_spots(sys.argv[1])
