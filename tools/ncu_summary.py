#!/usr/bin/env python
"""Summarise .ncu-rep captures (read offline with `ncu -i ... --page raw --csv`) into profiles/ncu/<name>.md + .raw.csv."""
import csv
import io
import os
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "lts__t_bytes.sum", "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum", "smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
]


def main():
    out_dir = sys.argv[1]
    os.makedirs(out_dir, exist_ok=True)
    for rep in sys.argv[2:]:
        name = os.path.basename(rep).replace(".ncu-rep", "")
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        if len(rows) < 3:
            print("no data in", rep)
            continue
        hdr, units, vals = rows[0], rows[1], rows[-1]
        if len(rows) > 3:  # several kernels in one report: one table row per launch
            col = {h: i for i, h in enumerate(hdr)}
            pick = ["gpu__time_duration.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
                    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
                    "launch__grid_size", "dram__bytes_read.sum", "dram__bytes_write.sum"]
            with open(os.path.join(out_dir, name + ".md"), "w") as f:
                f.write(f"# ncu --set full, every launch of `{os.path.basename(rep)}`\n\n| kernel | time | dram % | sm % | tensor % | warps active % | regs | grid | dram read | dram write |\n|---|---|---|---|---|---|---|---|---|---|\n")
                for r in rows[2:]:
                    kn = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("b2::", "")
                    cells = []
                    for k in pick:
                        if k in col:
                            v = r[col[k]]
                            try:
                                v = f"{float(v.replace(',', '')):.2f}"
                            except ValueError:
                                pass
                            cells.append(f"{v} {units[col[k]]}".strip())
                        else:
                            cells.append("-")
                    f.write(f"| `{kn}` | " + " | ".join(cells) + " |\n")
            with open(os.path.join(out_dir, name + ".raw.csv"), "w") as f:
                f.write(raw)
            print("wrote", name, f"({len(rows) - 2} launches)")
            continue
        with open(os.path.join(out_dir, name + ".raw.csv"), "w") as f:
            w = csv.writer(f)
            w.writerow(["metric", "unit", "value"])
            for h, u, v in zip(hdr, units, vals):
                w.writerow([h, u, v])
        d = dict(zip(hdr, zip(vals, units)))
        with open(os.path.join(out_dir, name + ".md"), "w") as f:
            f.write(f"# ncu --set full: `{d.get('Kernel Name', ('?',))[0]}`\n\n")
            f.write(f"capture: `{os.path.basename(rep)}` (`--clock-control none --import-source on`), full raw page in `{name}.raw.csv`\n\n| metric | value |\n|---|---|\n")
            for k in KEYS:
                if k in d:
                    f.write(f"| `{k}` | {d[k][0]} {d[k][1]} |\n")
        print("wrote", name)


if __name__ == "__main__":
    main()
