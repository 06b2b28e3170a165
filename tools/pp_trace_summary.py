"""Per-stage time accounting of the pipeline timelines written by B2F_PP_TRACE (device timestamps of every schedule
instruction; "x" = the exchange that follows an F / B, i.e. p2p plus whatever the stage waits for its neighbour).
usage: python tools/pp_trace_summary.py profiles/r2_call4_pp2_trace_1f1b profiles/r2_call4_pp2_trace_zb"""
import json
import sys
from pathlib import Path


def load(prefix):
    out = {}
    for f in sorted(Path(prefix).parent.glob(Path(prefix).name + ".stage*.jsonl")):
        stage = int(f.name.split(".stage")[1].split(".")[0])
        out[stage] = json.loads(f.read_text().strip().splitlines()[-1])      # the last recorded batch
    return out


def summarise(prefix):
    stages = load(prefix)
    print(f"## {Path(prefix).name}")
    for s, ev in sorted(stages.items()):
        tot = ev[-1][2] - ev[0][2]
        acc = {}
        for (op, mb, t), (_, _, t1) in zip(ev[:-1], ev[1:]):
            if op not in ("start",):
                acc[op] = acc.get(op, 0.0) + (t1 - t)
        busy = sum(v for k, v in acc.items() if k != "x")
        line = "  ".join(f"{k} {acc.get(k, 0.0):7.1f} ms ({100 * acc.get(k, 0.0) / tot:4.1f} %)" for k in ("F", "B", "W", "x"))
        print(f"stage {s}: batch {tot:7.1f} ms   {line}   compute {100 * busy / tot:4.1f} %")
        # one character per 1/100 of the batch: the instruction that owns that instant
        cells, k = [], 0
        for i in range(100):
            t = ev[0][2] + (i + 0.5) * tot / 100
            while k + 1 < len(ev) - 1 and ev[k + 1][2] <= t:
                k += 1
            op = ev[k][0]
            cells.append({"F": "F", "B": "B", "W": "w", "x": ".", "start": " "}.get(op, "?"))
        print("         " + "".join(cells))
    print()


if __name__ == "__main__":
    print("# F forward, B backward (dgrad only under the split-backward schedule), w deferred weight gradients, . exchange")
    print("# (send / receive plus waiting for the neighbouring stage).  Llama-2 7B, PP2, 16 micro-batches, 2 GPUs.\n")
    for p in sys.argv[1:]:
        summarise(p)
