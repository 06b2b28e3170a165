"""Per-kernel SASS evidence for profiles/: which Blackwell instructions each b200flow kernel uses.

Runs `cuobjdump -sass` on the in-tree objects (no GPU needed) and counts, per kernel, the
mnemonics that prove the sm_100a paths: UTCHMMA / UTCQMMA (tcgen05.mma), UTMALDG / UTMASTG (TMA),
LDTM / STTM (TMEM), UTCBAR (tcgen05.commit), SYNCS (mbarrier), UBLKCP (bulk copy), multimem
(LDGMC/REDG... .MC), *.SYS loads / stores (peer flags), MUFU, FFMA2.  Writes a markdown table and
a short listing excerpt around the first tensor instruction of every kernel that has one.
"""
from __future__ import annotations

import re
import subprocess
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OBJ = ROOT / "build" / "obj"
KEYS = ["UTCHMMA", "UTCQMMA", "UTCOMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "SYNCS",
        "MULTIMEM", ".SYS", "MUFU", "FFMA2", "ATOM", "RED"]


def demangle(name: str) -> str:
    try:
        out = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip()
        out = re.sub(r"\(.*", "", out)
        out = out.replace("(anonymous namespace)::", "").replace("void ", "")
        return out or name
    except FileNotFoundError:
        return name


def main(out_md: Path, out_listing: Path) -> None:
    rows = OrderedDict()
    listing = []
    for obj in sorted(OBJ.glob("*.o")):
        sass = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
        cur, lines = None, []
        def flush():
            if cur is None:
                return
            cnt = Counter()
            for ln in lines:
                m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
                if not m:
                    continue
                op = m.group(1)
                for k in KEYS:
                    if k == ".SYS":
                        if ".SYS" in op and (op.startswith("LD") or op.startswith("ST") or op.startswith("RED") or op.startswith("ATOM")):
                            cnt[k] += 1
                    elif k == "MULTIMEM":
                        if ".MC" in op or "MULTIMEM" in op or op.startswith("LDGMC") or op.startswith("STGMC") or op.startswith("REDGMC"):
                            cnt[k] += 1
                    elif op.startswith(k):
                        cnt[k] += 1
            rows[(obj.stem, demangle(cur))] = (cnt, len(lines))
            for i, ln in enumerate(lines):
                if "UTCHMMA" in ln or "UTCQMMA" in ln or "LDGMC" in ln or ".MC" in ln:
                    listing.append(f"### {obj.stem} :: {demangle(cur)}\n```\n" + "\n".join(
                        l.split("/*", 2)[0].rstrip() + " /*" + l.split("/*", 2)[1] if False else l[:118].rstrip()
                        for l in lines[max(0, i - 6): i + 10]) + "\n```\n")
                    break
        for ln in sass.splitlines():
            m = re.search(r"Function : (\S+)", ln)
            if m:
                flush()
                cur, lines = m.group(1), []
            elif cur and re.search(r"/\*[0-9a-f]{4,}\*/", ln):      # addresses grow past 4 hex digits after 64 KB of code
                lines.append(ln)
        flush()
    with out_md.open("w") as f:
        f.write("# SASS evidence per kernel (cuobjdump -sass of build/obj/*.o, sm_100a)\n\n")
        f.write("| object | kernel | instrs | " + " | ".join(KEYS) + " |\n|---|---|---|" + "---|" * len(KEYS) + "\n")
        for (o, k), (cnt, n) in rows.items():
            f.write(f"| {o} | `{k[:70]}` | {n} | " + " | ".join(str(cnt.get(x, 0) or "") for x in KEYS) + " |\n")
    out_listing.write_text("# SASS excerpts around the first tensor-core / multimem instruction\n\n" + "\n".join(listing))


FULL = {   # complete listings committed for the kernels the fused paths rest on (K0, K3/K4, K5, K7)
    "kernels_gemm2": "k0_gemm2", "comm_fused_gemm": "k3_k4_fused_gemm", "kernels_attention": "k7_attention",
    "kernels_gemm_mx": "k0_gemm_mxfp8", "comm_fused_moe": "k5_fused_moe", "comm_collectives": "k1_k2_collectives",
}


def full_listings(out_dir: Path) -> list[str]:
    """`cuobjdump -sass` of whole objects, instruction text only (encodings stripped to keep the files readable)."""
    out_dir.mkdir(parents=True, exist_ok=True)
    done = []
    for stem, name in FULL.items():
        obj = OBJ / f"{stem}.o"
        if not obj.exists():
            continue
        sass = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
        keep = []
        for ln in sass.splitlines():
            if "Function :" in ln:
                keep.append("\n" + ln.strip())
            else:
                m = re.search(r"/\*([0-9a-f]{4,})\*/\s+(.*?)\s*/\* 0x", ln)
                if m:
                    keep.append(f"/*{m.group(1)}*/ {m.group(2)}")
        (out_dir / f"{name}.sass").write_text("\n".join(keep) + "\n")
        done.append(name)
    return done


if __name__ == "__main__":
    main(ROOT / "profiles" / "sass_summary.md", ROOT / "profiles" / "sass_excerpts.md")
    print("wrote profiles/sass_summary.md, profiles/sass_excerpts.md")
    print("full listings:", full_listings(ROOT / "profiles" / "sass"))
