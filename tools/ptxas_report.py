"""Registers, spills, shared memory and barriers of every sm_100a kernel, from the `-Xptxas -v` lines the build keeps in
build/obj/*.log (`python -m b200flow.build` compiles with it).  usage: python tools/ptxas_report.py > profiles/ptxas_summary.txt"""
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LOGS = sorted((ROOT / "build" / "obj").glob("*_*.log"))


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
        return dict(zip(names, out))
    except OSError:
        return {n: n for n in names}


def main():
    rows = []
    for log in LOGS:
        text = log.read_text(errors="ignore")
        for m in re.finditer(r"Compiling entry function '([^']+)' for 'sm_100a'\n(?:.*\n)*?.*?(\d+) bytes stack frame, (\d+) bytes spill stores, "
                             r"(\d+) bytes spill loads\n.*?Used (\d+) registers(?:, used (\d+) barriers)?(?:, (\d+) bytes cumulative stack size)?"
                             r"(?:, (\d+) bytes smem)?", text):
            rows.append((log.stem, m.group(1), int(m.group(5)), int(m.group(3)), int(m.group(4)), int(m.group(2)),
                         int(m.group(8) or 0)))
    names = demangle(sorted({r[1] for r in rows}))
    print("# ptxas -v summary of the sm_100a build (b200flow/lib/libb200flow_cuda.so); spills are bytes per thread.")
    print("# Kernels that use setmaxnreg (attention, fused GEMMs) show the launch-time register count; dynamic shared")
    print("# memory is set at launch and not listed here.")
    print(f"{'translation unit':<22}{'regs':>5}{'spill st':>9}{'spill ld':>9}{'stack':>6}{'smem':>7}  kernel")
    for unit, mangled, regs, st, ld, stack, smem in sorted(rows, key=lambda r: (r[0], names[r[1]])):
        short = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", names[mangled])
        short = re.sub(r"\((?:[^()]|\([^()]*\))*\)$", "", short)
        print(f"{unit:<22}{regs:>5}{st:>9}{ld:>9}{stack:>6}{smem:>7}  {short[:110]}")
    worst = [r for r in rows if r[3] or r[4]]
    print(f"# {len(rows)} kernels; {len(worst)} with any spill; largest spill store {max((r[3] for r in rows), default=0)} B")


if __name__ == "__main__":
    sys.exit(main())
