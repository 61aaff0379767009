"""SASS evidence per kernel (W28): counts of the Blackwell-specific instructions in every kernel of libddpx_kernels.so
(cuobjdump -sass on the object files), written as a markdown table.   python bench/sass_summary.py > profiles/sass_summary.md"""
import re
import subprocess
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
OBJ = REPO / "build" / "obj"
MNEMONICS = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UTCBAR", "UTCATOM", "LDTM", "STTM", "SYNCS", "UCGABAR", "REDG",
             "LDGMC", "ATOMS.CAST.SPIN"]


def demangle(n):
    try:
        return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip() or n
    except Exception:
        return n


def main():
    print("# SASS summary (cuobjdump -sass, sm_100a)\n")
    print("Counts of selected instructions per kernel; `UTCHMMA` = tcgen05.mma, `UTMALDG` = TMA tensor load, `UTCBAR` = "
          "tcgen05.commit, `LDTM` / `STTM` = tcgen05.ld / tcgen05.st, `SYNCS` = mbarrier ops, `UTMASTG` = TMA tensor store (absent: "
          "epilogues store through registers / shared-memory transposes), `LDGMC` = multimem.ld_reduce (the in-switch NVLS "
          "reduction), `ATOMS.CAST.SPIN` = shared-memory float atomic as a compare-and-swap loop (should be 0), `.2CTA` forms "
          "appear in the CTA-pair flavours (`kCG = 2`).  Regenerate with `python bench/sass_summary.py`.\n")
    for obj in sorted(OBJ.glob("*.o")):
        out = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
        if "Function :" not in out:
            continue
        res = subprocess.run(["cuobjdump", "--dump-resource-usage", str(obj)], capture_output=True, text=True).stdout
        regs = {}
        cur = None
        for line in res.splitlines():
            m = re.search(r"Function (\S+):", line)
            if m:
                cur = m.group(1)
            m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
            if m and cur:
                regs[cur] = (m.group(1), m.group(2))
        print(f"## {obj.name}\n")
        print("| kernel | regs | instr | " + " | ".join(MNEMONICS) + " | 2CTA forms |")
        print("|---|---|---|" + "---|" * (len(MNEMONICS) + 1))
        parts = out.split("Function : ")[1:]
        for part in parts:
            name = part.split("\n", 1)[0].strip()
            body = part
            ninstr = len(re.findall(r"/\*[0-9a-f]{4,}\*/\s+\S", body))
            counts = [len(re.findall(r"\b" + re.escape(m), body)) for m in MNEMONICS]
            two = len(re.findall(r"\.2CTA", body))
            short = demangle(name).replace("(anonymous namespace)::", "").replace("void ", "")
            short = re.sub(r"\(.*", "", short)[:70]
            print(f"| `{short}` | {regs.get(name, ('?', '?'))[0]} | {ninstr} | " + " | ".join(str(c) for c in counts) + f" | {two} |")
        print()


if __name__ == "__main__":
    main()
