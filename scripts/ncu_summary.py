#!/usr/bin/env python3
"""ncu_summary.py REPORT.ncu-rep [OUT.md] — the per-launch numbers of an `ncu --set full` capture that the roofline
discussion in DESIGN.md / profiles/README.md quotes, as a markdown table (one row per captured launch)."""
import csv
import subprocess
import sys

WANT = [
    ("Kernel Name", "kernel"),
    ("Grid Size", "grid"),
    ("Block Size", "block"),
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor inst"),
    ("sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active", "imma %"),
    ("smsp__inst_executed.sum", "inst"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    cols = [(hdr.index(k), name) for (k, name) in WANT if k in hdr]
    out = ["| " + " | ".join(f"{name} [{units[i]}]" if units[i] else name for i, name in cols) + " |",
           "|" + "---|" * len(cols)]
    for r in rows[2:]:
        cells = []
        for i, _ in cols:
            v = r[i]
            try:
                f = float(v)
                v = f"{f:.3f}".rstrip("0").rstrip(".") if "." in v else v
            except ValueError:
                v = v.replace("|", "/")[:70]
            cells.append(v)
        out.append("| " + " | ".join(cells) + " |")
    text = "\n".join(out) + "\n"
    if len(sys.argv) > 2:
        with open(sys.argv[2], "w") as f:
            f.write(f"<!-- ncu -i {rep.split('/')[-1]} --page raw --csv, reduced by scripts/ncu_summary.py -->\n" + text)
    else:
        print(text)


if __name__ == "__main__":
    main()
