"""Pulls the headline metrics of every kernel in an `ncu --set full` report into a markdown table.
usage: python profiles/ncu_extract.py gpurun_out/x.ncu-rep > profiles/ncu_rNN_x.md  (needs ncu on PATH)"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("sm__inst_executed.sum", "warp inst"),
    ("sm__inst_executed.avg.per_cycle_elapsed", "ipc"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
    ("smsp__inst_executed_op_shared_atom.sum", "shared atomics (inst)"),
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units = rows[0], dict(zip(rows[0], rows[1]))
    print("| kernel | " + " | ".join(n for _, n in KEYS) + " |")
    print("|---|" + "---:|" * len(KEYS))
    for r in rows[2:]:
        d = dict(zip(head, r))
        cells = []
        for k, _ in KEYS:
            v = d.get(k, "")
            cells.append((v + " " + units.get(k, "")).strip() if v else "-")
        print("| `" + d["Kernel Name"].split("(")[0] + "` | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
