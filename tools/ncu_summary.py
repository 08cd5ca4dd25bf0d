"""One line per profiled launch from `ncu -i X.ncu-rep --page raw --csv` output: the figures profiles/README.md quotes.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv [more.csv ...]

Columns are matched by metric name (the section prefix differs between ncu versions); missing ones print as '-'.
"""
import csv
import sys

WANTED = [
    ("time us", "gpu__time_duration.sum", 1e-3),                      # ns in the raw page
    ("dram rd MB", "dram__bytes_read.sum", 1e-6),
    ("dram wr MB", "dram__bytes_write.sum", 1e-6),
    ("dram GB/s", "dram__bytes.sum.per_second", 1.0),
    ("L2 hit %", "lts__t_sector_hit_rate.pct", 1.0),
    ("warps act %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    ("issue %", "sm__inst_issued.avg.pct_of_peak_sustained_active", 1.0),
    ("fp64 pipe %", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", 1.0),
    ("regs", "launch__registers_per_thread", 1.0),
    ("smem/blk KB", "launch__shared_mem_per_block_dynamic", 1e-3),
]


def column(header, metric):
    """index of the column whose name ends with the metric (after an optional section prefix)"""
    exact = [i for i, h in enumerate(header) if h == metric]
    prefixed = [i for i, h in enumerate(header) if h.endswith("." + metric)]
    return (exact + prefixed)[0] if exact or prefixed else None


def number(text):
    try:
        return float(text.replace(",", ""))
    except ValueError:
        return None


def main():
    for path in sys.argv[1:]:
        with open(path, newline="") as f:
            rows = [r for r in csv.reader(f) if r]
        header = rows[0]
        body = [r for r in rows[1:] if r[0].strip().isdigit()]  # the row after the header holds the units
        units = rows[1] if len(rows) > 1 and not rows[1][0].strip().isdigit() else None
        name_col = header.index("Kernel Name")
        cols = [(label, column(header, metric), scale) for label, metric, scale in WANTED]
        print(path)
        print("  " + " | ".join(["kernel".ljust(34), "grid x block".ljust(16)] + [label.rjust(11) for label, _, _ in cols]))
        for r in body:
            cells = []
            for label, c, scale in cols:
                v = number(r[c]) if c is not None else None
                if v is not None and units is not None and label == "time us":
                    unit = units[c].strip().lower()
                    v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
                if v is not None and units is not None and label.startswith("dram") and label.endswith("MB"):
                    unit = units[c].strip().lower()
                    v *= {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(unit, 1.0)
                if v is not None and units is not None and label == "dram GB/s":
                    v *= {"byte/s": 1e-9, "kbyte/s": 1e-6, "mbyte/s": 1e-3, "gbyte/s": 1.0, "tbyte/s": 1e3}.get(units[c].strip().lower(), 1.0)
                cells.append(("%.2f" % (v * scale)).rjust(11) if v is not None else "-".rjust(11))
            grid = r[header.index("Grid Size")].replace(" ", "") + " x " + r[header.index("Block Size")].replace(" ", "")
            print("  " + " | ".join([r[name_col][:34].ljust(34), grid[:16].ljust(16)] + cells))


if __name__ == "__main__":
    main()
