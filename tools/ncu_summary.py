"""Markdown summary of an .ncu-rep: the metrics the roofline discussion uses plus the warp-stall mix.

    python tools/ncu_summary.py profiles/k_darlin_pass_v4.ncu-rep
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]


def main(path: str) -> int:
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    print("| metric | value |\n|---|---|")
    for k in KEYS:
        for h, u, v in zip(hdr, units, vals):
            if h == k:
                print(f"| {h} | {v} {u} |")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = rows[1]
    st = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = {hdr[i]: 0 for i in st}
    for r in rows[2:]:
        for i in st:
            try:
                tot[hdr[i]] += int(r[i])
            except (ValueError, IndexError):
                pass
    s = sum(tot.values()) or 1
    top = sorted(tot.items(), key=lambda x: -x[1])[:6]
    print("| warp stall sampling | " + ", ".join(f"{k[6:]} {100 * v / s:.1f} %" for k, v in top) + " |")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1]))
