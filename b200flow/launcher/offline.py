"""Offline shims: make the reference's values files runnable with no network.

The reference's scripts `git clone` an upstream repo, `pip install` pins and sometimes
`wget` data (e.g. examples/accelerate/bert-glue-mrpc/pretrain.yaml:20-27).  On the box there
is no network, so the launcher prepends a directory of tiny executables to `PATH`:

* `git clone <url> <dir>` of a known upstream URL copies b200flow's drop-in workload tree
  (same relative entry-point paths, e.g. `examples/complete_nlp_example.py`,
  `pretrain_gpt.py`, `tools/preprocess_data.py`) into `<dir>`; `checkout` / `fetch` /
  `reset` inside such a tree succeed as no-ops; anything else goes to the real git.
* `pip` / `pip3 install …` log and succeed (`offline: skip-pip`).
* `yum` / `apt-get` log and succeed without doing anything.
* `wget -O <file> <url>` leaves a synthetic stand-in for the download (`data/synthetic.py::write_standin`: a generated
  text corpus for `.jsonl`, a placeholder for backbone `.npz` files, a small marker for archives) and `unzip <marker>
  -d <dir>` expands a marker into the tree the real archive holds (COCO 2017 images / annotations); so the download
  scripts of the `coco-data` and `redpajama-data` charts run unchanged.  Real archives still go to the real `unzip`.
* absolute script paths baked into images (`/tensorpack/examples/FasterRCNN/train.py`,
  `/mask-rcnn-tensorflow/MaskRCNN/train.py`) are resolved by a `python3` shim.
"""
from __future__ import annotations

import stat
import sys
from pathlib import Path

WORKLOADS = Path(__file__).resolve().parent.parent / "workloads"
STANDIN_KEY = "b200flow_offline_standin"      # first key of the JSON marker that stands in for a downloaded archive

# upstream URL fragment -> local tree (relative to b200flow/workloads/trees)
REPO_TREES = {
    "huggingface/accelerate": "accelerate",
    "microsoft/Megatron-DeepSpeed": "megatron_deepspeed",
    "ray-project/kuberay": "kuberay",
}

# image-baked absolute paths -> local entry points
ABSOLUTE_SCRIPTS = {
    "/tensorpack/examples/FasterRCNN/train.py": WORKLOADS / "maskrcnn" / "train.py",
    "/mask-rcnn-tensorflow/MaskRCNN/train.py": WORKLOADS / "maskrcnn" / "train.py",
}


def resolve_repo(url: str) -> Path | None:
    for frag, tree in REPO_TREES.items():
        if frag.lower() in url.lower():
            return WORKLOADS / "trees" / tree
    return None


_GIT = r'''#!{py}
import os, shutil, subprocess, sys
sys.path.insert(0, {repo!r})
from b200flow.launcher.offline import resolve_repo
args = sys.argv[1:]
if args[:1] == ["clone"]:
    rest = [a for a in args[1:] if not a.startswith("-")]
    url, dst = rest[0], (rest[1] if len(rest) > 1 else os.path.basename(rest[0]).removesuffix(".git"))
    tree = resolve_repo(url)
    if tree is not None and tree.is_dir():
        shutil.copytree(tree, dst, dirs_exist_ok=True)
        open(os.path.join(dst, ".b200flow_tree"), "w").write(url + "\n")
        print(f"[b200flow offline] git clone {{url}} -> local tree {{tree}}")
        sys.exit(0)
    print(f"[b200flow offline] no local tree for {{url}}", file=sys.stderr)
    sys.exit(128)
if args[:1] and args[0] in ("checkout", "fetch", "reset", "pull", "submodule") and os.path.exists(".b200flow_tree"):
    print("[b200flow offline] git " + " ".join(args) + " (no-op in local tree)")
    sys.exit(0)
real = next((os.path.join(d, "git") for d in os.environ.get("PATH", "").split(os.pathsep)[1:]
             if os.access(os.path.join(d, "git"), os.X_OK)), None)
sys.exit(subprocess.call([real] + args) if real else 127)
'''

_NOOP = r'''#!/bin/bash
echo "[b200flow offline] skipped: {name} $*"
exit 0
'''

_WGET = r'''#!{py}
import json, os, sys
args = sys.argv[1:]
out, url, i = None, None, 0
while i < len(args):
    if args[i] in ("-O", "--output-document") and i + 1 < len(args):
        out, i = args[i + 1], i + 2
        continue
    if not args[i].startswith("-"):
        url = args[i]
    i += 1
if out is None or url is None:
    print("[b200flow offline] skipped: wget " + " ".join(args))
    sys.exit(0)
name = url.rstrip("/").rsplit("/", 1)[-1]
if name.endswith((".zip", ".tar", ".tar.gz", ".tgz")):      # an archive: a marker the unzip shim expands (no heavy imports)
    os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
    with open(out, "w") as f:
        f.write(json.dumps({{{key!r}: url, "name": name}}) + "\n")
    print(f"[b200flow offline] wget {{url}} -> {{out}}: archive marker")
    sys.exit(0)
sys.path.insert(0, {repo!r})
from b200flow.data.synthetic import main
main(["standin", "--url", url, "--out", out])
'''

_UNZIP = r'''#!{py}
import os, subprocess, sys
sys.path.insert(0, {repo!r})
args = sys.argv[1:]
files = [a for i, a in enumerate(args) if not a.startswith("-") and (i == 0 or args[i - 1] != "-d")]
dest = args[args.index("-d") + 1] if "-d" in args and args.index("-d") + 1 < len(args) else "."
from b200flow.data.synthetic import main, read_standin
if files and read_standin(files[0]) is not None:
    main(["unpack", "--archive", files[0], "--out", dest])
    sys.exit(0)
real = next((os.path.join(d, "unzip") for d in os.environ.get("PATH", "").split(os.pathsep)[1:]
             if os.access(os.path.join(d, "unzip"), os.X_OK)), None)
sys.exit(subprocess.call([real] + args) if real else 127)
'''

_PY = r'''#!{py}
import os, sys
sys.path.insert(0, {repo!r})
from b200flow.launcher.offline import ABSOLUTE_SCRIPTS
args = sys.argv[1:]
for i, a in enumerate(args):
    if a in ABSOLUTE_SCRIPTS:
        args[i] = str(ABSOLUTE_SCRIPTS[a])
        break
    if not a.startswith("-"):
        break
os.execv({py!r}, [{py!r}] + args)
'''


def write_shims(dirpath: Path) -> None:
    dirpath.mkdir(parents=True, exist_ok=True)
    repo = str(Path(__file__).resolve().parent.parent.parent)
    py = sys.executable
    files = {"git": _GIT.format(py=py, repo=repo),
             "python3": _PY.format(py=py, repo=repo),
             "wget": _WGET.format(py=py, repo=repo, key=STANDIN_KEY),
             "unzip": _UNZIP.format(py=py, repo=repo)}
    for name in ("pip", "pip3", "yum", "apt-get", "apt"):
        files[name] = _NOOP.format(name=name)
    for name, body in files.items():
        p = dirpath / name
        p.write_text(body)
        p.chmod(p.stat().st_mode | stat.S_IXUSR | stat.S_IXGRP | stat.S_IXOTH)
