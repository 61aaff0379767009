"""Table of one Nsight Compute report: per captured launch its duration, DRAM bytes read / written, % of the DRAM peak,
tensor-pipe activity, registers, grid, achieved occupancy and the top stall reason.
    python bench/ncu_extract.py gpurun_out/prof_families.ncu-rep [algorithmic-bytes log]  > profiles/ncu_families.md
The optional second argument is the text bench/ncu_target_families.py prints (name, bytes): DRAM bytes are then set
against the bytes the operation has to move."""
import csv
import io
import subprocess
import sys


def raw_page(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main():
    rep = sys.argv[1]
    head, units, rows = raw_page(rep)
    ix = {n: i for i, n in enumerate(head)}

    def get(row, name, default=""):
        i = ix.get(name)
        return row[i] if i is not None and row[i] != "" else default

    def num(row, name):
        try:
            return float(get(row, name, "nan").replace(",", ""))
        except ValueError:
            return float("nan")

    stalls = [n for n in head if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")]
    if not stalls:
        stalls = [n for n in head if n.startswith("smsp__average_warp") and "stalled" in n]
    print("| # | kernel | grid x block | regs | time us | DRAM rd MB | DRAM wr MB | DRAM % of peak | tensor pipe % | "
          "occupancy % | top stall |")
    print("|---|---|---|---|---|---|---|---|---|---|---|")
    for n, row in enumerate(rows):
        t_unit = units[ix["gpu__time_duration.sum"]]
        t = num(row, "gpu__time_duration.sum") * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(t_unit, 1.0)

        def mb(name):
            u = units[ix[name]]
            return num(row, name) * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
        best = ("", 0.0)
        for s in stalls:
            v = num(row, s)
            if v == v and v > best[1]:
                best = (s.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v)
        name = get(row, "Kernel Name")
        name = name.split("(")[0][-60:]
        print(f"| {n} | `{name}` | {get(row, 'Grid Size')} x {get(row, 'Block Size')} | "
              f"{get(row, 'launch__registers_per_thread')} | {t:.1f} | {mb('dram__bytes_read.sum'):.1f} | "
              f"{mb('dram__bytes_write.sum'):.1f} | {num(row, 'dram__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
              f"{num(row, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed'):.1f} | "
              f"{num(row, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | {best[0]} ({best[1]:.1f}) |")
    if len(sys.argv) > 2:
        print("\nAlgorithmic bytes (or FLOPs where marked) of the launches, in launch order, as printed by the target script:\n")
        print("```")
        print(open(sys.argv[2]).read().rstrip())
        print("```")


if __name__ == "__main__":
    main()
