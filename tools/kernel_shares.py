"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel time shares."""
import collections
import csv
import re
import sys


def main(path, out=None, title=""):
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[hi]
    kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[hi + 1:]:
        if len(r) <= mv:
            continue
        name = re.sub(r"\(.*", "", r[kn])
        name = re.sub(r"^void ", "", name)[:80]
        try:
            v = float(r[mv].replace(",", ""))
        except ValueError:
            continue
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    lines = [f"# {title}", f"# total {tot / 1e6:.2f} ms over {sum(v[0] for v in agg.values())} launches"]
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:32]:
        lines.append(f"{100 * t / tot:5.1f}%  {t / 1e3:9.1f} us  n={n:4d}  avg={t / n / 1e3:8.1f} us  {k}")
    text = "\n".join(lines) + "\n"
    print(text)
    if out:
        open(out, "w").write(text)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None, sys.argv[3] if len(sys.argv) > 3 else "")
