"""Markdown table + JSON of the per-kernel metrics of an .ncu-rep: python tools/ncu_table.py rep [out.json]"""
import csv
import json
import subprocess
import sys

COLS = [("us", "gpu__time_duration.sum", 1.0), ("rd MB", "dram__bytes_read.sum", 1.0), ("wr MB", "dram__bytes_write.sum", 1.0),
        ("dram %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
        ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
        ("occ %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
        ("regs", "launch__registers_per_thread", 1.0),
        ("fma %", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", 1.0),
        ("alu %", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", 1.0),
        ("xu %", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", 1.0),
        ("lsu %", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", 1.0),
        ("tensor %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
        ("L2 hit %", "lts__t_sector_hit_rate.pct", 1.0), ("Minst", "smsp__inst_executed.sum", 1e-6)]
# older reports name the tensor pipe differently: first column found wins
ALIASES = {"tensor %": ["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
                        "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
                        "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active"]}


def main(rep, out=None):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    print("| kernel | " + " | ".join(c[0] for c in COLS) + " |")
    print("|---|" + "---|" * len(COLS))
    js = []
    for r in rows[2:]:
        name = r[ik].split("(")[0].replace("csb::", "").replace("void ", "")
        vals = {}
        for label, m, scale in COLS:
            if m not in hdr:
                m = next((a for a in ALIASES.get(label, []) if a in hdr), None)
            if m is None or m not in hdr:
                vals[label] = None
                continue
            i = hdr.index(m)
            v = float(r[i].replace(",", "")) * scale
            u = units[i]
            if label.endswith("MB"):
                v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
            if label == "us":
                v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(u, 1.0)
            vals[label] = v
        print(f"| `{name}` | " + " | ".join("-" if vals[c[0]] is None else f"{vals[c[0]]:.1f}" for c in COLS) + " |")
        js.append({"kernel": name, **vals})
    if out:
        json.dump(js, open(out, "w"), indent=1)


if __name__ == "__main__":
    main(*sys.argv[1:])
