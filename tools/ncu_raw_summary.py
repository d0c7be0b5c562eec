"""Markdown table of the per-launch headline metrics of an `ncu --set full` report.
usage: python tools/ncu_raw_summary.py report.ncu-rep > table.md"""
import csv
import io
import subprocess
import sys

COLS = [("ms", "gpu__time_duration.sum"), ("grid", "Grid Size"), ("block", "Block Size"),
        ("DRAM rd MB", "dram__bytes_read.sum"), ("DRAM wr MB", "dram__bytes_write.sum"),
        ("DRAM %", "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("tensor %", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active"),
        ("tensor(all) %", "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active"),
        ("smem %", "l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed"),
        ("warps %", "sm__warps_active.avg.pct_of_peak_sustained_active"), ("regs", "launch__registers_per_thread")]


def main():
    raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]

    def find(name):
        for i, h in enumerate(head):
            if h == name or h.endswith("." + name) or h == name.split(".")[-1]:
                return i
        for i, h in enumerate(head):
            if name in h:
                return i
        return None

    idx = [(lab, find(n)) for lab, n in COLS]
    idx = [(lab, i) for lab, i in idx if i is not None]
    kn = head.index("Kernel Name")
    print("| kernel | " + " | ".join(f"{lab} [{units[i]}]" if units[i] else lab for lab, i in idx) + " |")
    print("|---|" + "---|" * len(idx))
    for r in body:
        name = r[kn].split("(")[0].replace("void ", "")
        vals = []
        for lab, i in idx:
            v = r[i]
            try:
                v = f"{float(v.replace(',', '')):.3f}".rstrip("0").rstrip(".")
            except ValueError:
                pass
            vals.append(v)
        print(f"| `{name}` | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main()
