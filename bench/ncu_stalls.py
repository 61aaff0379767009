"""Per-instruction stall samples of one captured launch, from `ncu --page source --csv`:
    python bench/ncu_stalls.py report.ncu-rep <kernel regex> [launch index among the matches] [min samples]
Prints the instructions with at least `min samples` samples plus every barrier / TMA / tensor-memory / store
instruction, with the running total, so that a stall can be attributed to a phase of the kernel."""
import csv
import io
import re
import subprocess
import sys


def launches(rep, regex):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{regex}"],
                         capture_output=True, text=True, check=True).stdout
    cur = None
    res = []
    for row in csv.reader(io.StringIO(out)):
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1], "head": None, "rows": []}
            res.append(cur)
        elif cur is not None and cur["head"] is None:
            cur["head"] = row
        elif cur is not None and len(row) == len(cur["head"]):
            cur["rows"].append(row)
    return res


def main():
    rep, regex = sys.argv[1], sys.argv[2]
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    floor = int(sys.argv[4]) if len(sys.argv) > 4 else 12
    k = launches(rep, regex)[which]
    ix = {n: i for i, n in enumerate(k["head"])}
    stall_cols = [n for n in k["head"] if n.startswith("stall_") and "Not Issued" not in n]
    tot = sum(int(r[ix["# Samples"]]) for r in k["rows"])
    print(f"# {k['name'][:90]}: {tot} samples, {len(k['rows'])} instructions")
    key = re.compile(r"UTMALDG|UTMASTG|UTCBAR|SYNCS\.PHASECHK|BAR\.|EXIT|UTCATOM|LDTM|STTM|STG|STS|LDS|UCGABAR|REDG|MEMBAR")
    acc = 0
    for i, r in enumerate(k["rows"]):
        s = int(r[ix["# Samples"]])
        acc += s
        src = r[ix["Source"]].strip()
        ex = r[ix["Instructions Executed"]]
        if s >= floor or (key.search(src) and ex not in ("", "0")):
            top = sorted(((int(r[ix[c]]), c[6:]) for c in stall_cols), reverse=True)[0]
            print(f"{i:5d} {s:5d} {ex:>9} cum={acc:6d} {src[:64]:64s} {top[1] if top[0] else ''}")


if __name__ == "__main__":
    main()
