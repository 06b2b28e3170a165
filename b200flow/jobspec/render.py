"""Values -> rendered job: the bash script a "pod" runs plus the process-topology plan.

The script keeps the reference's semantics (the ConfigMap at
charts/machine-learning/training/pytorchjob-distributed/templates/train.yaml:7-47 and its
four siblings): optional git clone into `$HOME/tmp/$HOSTNAME` (exported as `GIT_CLONE_DIR`,
cwd moved there), `pre_script` lines verbatim in the same shell, the command + args joined by
line continuations and terminated by `&& echo "Training script done"`, `post_script`, clone
cleanup.  Two deliberate, opt-in fixes of reference holes: `strict_exit` adds
`set -o pipefail` and propagates the command's status (the reference's `| tee` masks
failures), and `offline` lets `git clone` / `pip install` resolve locally.

The plan (`JobPlan`) is what the reference expresses as a PyTorchJob / elastic PyTorchJob /
MPIJob / RayJob / Pod manifest: replicas, processes per replica, environment contract
(`PET_*`, `OMPI_*`), mounts, restart budget.
"""
from __future__ import annotations

import datetime as _dt
import shlex
from dataclasses import dataclass, field
from pathlib import Path

from . import values as V
from .charts import Chart, default_values, get_chart
from .template import Renderer, TemplateError


@dataclass
class JobPlan:
    release: str
    chart: str
    kind: str
    nnodes: int = 1
    nproc_per_node: int | None = None
    gpus_per_node: int | None = None
    env: dict = field(default_factory=dict)
    mounts: dict = field(default_factory=dict)        # container path -> local path
    backoff_limit: int = 0
    elastic: dict | None = None                       # rdzv_backend, rdzv_port, min/max replicas
    mpirun_args: list = field(default_factory=list)   # templated args before the script
    mpirun_np: int | None = None
    launcher_command: list = field(default_factory=list)  # maskrcnn charts: python3 train.py ...
    ray: dict | None = None
    image: str | None = None
    script_name: str = "train-script.sh"
    shell: str = "bash"
    script: str = ""
    working_dir: str | None = None
    annotations: dict = field(default_factory=dict)


def _ctx(values: dict, release: str, when: _dt.datetime, chart: Chart) -> Renderer:
    return Renderer({"Values": values,
                     "Release": {"Name": release, "Service": "b200flow", "Namespace": "local", "Time": when},
                     "Chart": {"Name": chart.name, "AppVersion": "1.0.0", "Version": "0.1.0"}})


def _script_lines(values: dict, workload_key: str, strict_exit: bool) -> list[str]:
    git = values.get("git") or {}
    out = ["#!/bin/bash"]
    if strict_exit:
        out.append("set -o pipefail")
    cloned = bool(git.get("repo_url"))
    if cloned:
        out += ["mkdir -p $HOME/tmp",
                "GIT_CLONE_DIR=$HOME/tmp/$HOSTNAME",
                "[[ -d $GIT_CLONE_DIR ]] && rm -rf $GIT_CLONE_DIR",
                f"git clone {git['repo_url']} $GIT_CLONE_DIR",
                "cd $GIT_CLONE_DIR"]
        if git.get("branch"):
            out.append(f"git checkout {git['branch']}")
        if git.get("commit"):
            out += [f"git fetch origin {git['commit']}", f"git reset --hard {git['commit']}"]
    out += [str(line) for line in (values.get("pre_script") or [])]
    work = values.get(workload_key) or {}
    for item in list(work.get("command") or []) + list(work.get("args") or []):
        out.append(f"{item} \\")
    # charts/machine-learning/data-prep/data-process/templates/data-process.yaml:38 says "Processing", the four
    # training charts say "Training" (…/templates/train.yaml:38)
    out.append('&& echo "Processing script done"' if workload_key == "process" else '&& echo "Training script done"')
    if strict_exit:
        out.append("B2F_STATUS=$?")
    out += [str(line) for line in (values.get("post_script") or [])]
    if cloned:
        out += ["cd $HOME", "rm -rf $GIT_CLONE_DIR"]
    if strict_exit:
        out.append("exit $B2F_STATUS")
    return out


def _env(values_env, r: Renderer) -> dict:
    env = {}
    for e in values_env or []:
        env[str(e["name"])] = r.render(str(e.get("value", "")))
    return env


def _mounts(pvcs, root: Path) -> dict:
    """pvc mount paths -> directories under the local root (`/efs` -> <root>/efs)."""
    if isinstance(pvcs, dict):
        pvcs = [pvcs]
    return {str(p["mount_path"]): str(root / str(p["mount_path"]).strip("/")) for p in (pvcs or [])}


# What the two download charts fetch (charts/machine-learning/data-prep/coco-data/templates/coco-data.yaml:19-35,
# redpajama-data/templates/redpajama-data.yaml:15).  Offline the launcher's `wget` / `unzip` / `yum` shims stand in.
COCO_HOST = "http://images.cocodataset.org"
COCO_ARCHIVES = ("zips/train2017.zip", "zips/val2017.zip", "zips/test2017.zip",
                 "annotations/annotations_trainval2017.zip")
BACKBONE_HOST = "http://models.tensorpack.com/FasterRCNN"
BACKBONES = ("ImageNet-R50-AlignPadding.npz", "ImageNet-R101-AlignPadding.npz")
REDPAJAMA_BOOK = "https://data.together.xyz/redpajama-data-1T/v1.0.0/book/book.jsonl"


def _data_script(kind: str) -> list[str]:
    """The script of the coco-data / redpajama-data chart: install the tools, leave an existing $DATA_DIR alone,
    download (and unpack) into it."""
    tools = ("unzip", "wget") if kind == "coco" else ("wget",)
    out = ["#!/bin/bash"] + [f"yum install -y {t}" for t in tools]
    out += ['[[ -d $DATA_DIR ]] && echo "$DATA_DIR already exists" && exit 0', "mkdir -p $DATA_DIR",
            'echo "Data dir: $DATA_DIR"']
    if kind != "coco":
        return out + [f"wget -O $DATA_DIR/data.jsonl {REDPAJAMA_BOOK}"]
    out += ["DOWNLOAD_DIR=$DATA_DIR.$(cat /dev/urandom | tr -dc 'a-z0-9' | fold -w 16 | head -n 1)",
            "mkdir -p $DOWNLOAD_DIR", 'echo "Download dir: $DOWNLOAD_DIR"']
    for rel in COCO_ARCHIVES:
        name = rel.rsplit("/", 1)[-1]
        out += [f"wget -O $DOWNLOAD_DIR/{name} {COCO_HOST}/{rel}", f"unzip $DOWNLOAD_DIR/{name} -d $DATA_DIR"]
    out.append("mkdir $DATA_DIR/pretrained-models")
    out += [f"wget -O $DATA_DIR/pretrained-models/{b} {BACKBONE_HOST}/{b}" for b in BACKBONES]
    return out + ["rm -rf $DOWNLOAD_DIR"]


def render_job(chart_name: str, release: str, value_files=(), sets=(), root: str | Path = "/tmp/b200flow",
               strict_exit: bool = False, when: _dt.datetime | None = None,
               extra_values: dict | None = None, strict_required: bool = False) -> JobPlan:
    chart = get_chart(chart_name)
    vals = default_values(chart)
    for f in value_files:
        vals = V.deep_merge(vals, V.load_yaml(f) if not isinstance(f, dict) else f)
    if extra_values:
        vals = V.deep_merge(vals, extra_values)
    for s in sets:
        vals = V.apply_set(vals, s)
    when = when or _dt.datetime.now()
    r = _ctx(vals, release, when, chart)
    root = Path(root)
    plan = JobPlan(release=release, chart=chart.name, kind=chart.kind, image=vals.get("image"),
                   script_name=chart.script_name, shell=chart.shell,
                   annotations={"app.kubernetes.io/instance": release,
                                "app.kubernetes.io/managed-by": "b200flow"})

    if chart.kind in ("static", "elastic", "rayjob", "pod") and chart.workload_key:
        res = vals.get("resources") or {}
        plan.nnodes = int(res.get("nnodes") or 1)
        plan.gpus_per_node = V.gpus_per_node(vals)
        npn = res.get("nproc_per_node")
        plan.nproc_per_node = int(npn) if npn else plan.gpus_per_node
        plan.backoff_limit = int(vals.get("backoff_limit") or 0)
        plan.env = _env((vals.get(chart.workload_key) or {}).get("env"), r)
        plan.mounts = _mounts(vals.get("pvc"), root)
        plan.script = "\n".join(_script_lines(vals, chart.workload_key, strict_exit)) + "\n"
        if chart.kind == "elastic":
            ep = vals.get("elastic_policy") or {}
            plan.elastic = {"rdzv_backend": ep.get("rdzv_backend", "c10d"),
                            "rdzv_port": int(ep.get("rdzv_port", 44000)),
                            "min_replicas": int(ep.get("min_replicas", 1)),
                            "max_replicas": int(ep.get("max_replicas", 1)), "rdzv_id": release}
        if chart.kind == "rayjob":
            ray = vals.get("ray") or {}
            plan.ray = {"version": ray.get("version"), "dashboard_host": (ray.get("dashboard") or {}).get("host"),
                        "ports": ray.get("ports") or [], "runtime_env_yaml": ray.get("runtime_env_yaml"),
                        "workers": plan.nnodes, "shutdown_after_job_finishes": True}
        return plan

    if chart.kind == "mpijob":
        res = vals.get("resources") or {}
        plan.nnodes = int(res.get("gpu_nodes") or 1)
        plan.gpus_per_node = int(res.get("gpus_per_node") or 1)
        plan.nproc_per_node = plan.gpus_per_node
        plan.mpirun_np = plan.nnodes * plan.gpus_per_node
        plan.backoff_limit = int(vals.get("backoff_limit") or 0)
        mp = vals.get("mpirun") or {}
        plan.env = _env(mp.get("env"), r)
        plan.mpirun_args = [r.render(str(a)) for a in (mp.get("args") or [])]
        plan.mounts = _mounts(vals.get("pvc"), root)
        plan.script = "\n".join(_script_lines(vals, "train", strict_exit)) + "\n"
        return plan

    if chart.kind == "maskrcnn":
        g, m = vals["global"], vals["maskrcnn"]
        fs = g["shared_fs"]
        stamp = r.render('{{ date "2006-01-02-15-04-05" .Release.Time }}')
        logroot = f"/{fs}/logs/{release}-{stamp}"
        plan.nnodes, plan.gpus_per_node = int(m["gpu_nodes"]), int(m["gpus_per_node"])
        plan.nproc_per_node = plan.gpus_per_node
        plan.mpirun_np = int(m["gpus"])
        plan.backoff_limit = int(m.get("backoff_limit") or 0)
        plan.image = m.get("image")
        plan.working_dir = m.get("working_dir")
        plan.mounts = {f"/{fs}": str(root / fs), "/ebs": str(root / "ebs")}
        plan.env = {"HOROVOD_AUTOTUNE": str(m["horovod_autotune"]), "HOROVOD_LOG_LEVEL": str(m["horovod_log_level"]),
                    "NCCL_SOCKET_IFNAME": str(m["nccl_socket_ifname"]), "NCCL_DEBUG": str(m["nccl_debug"]),
                    "TF_DEVICE_MIN_SYS_MEMORY_IN_MB": str(m["tf_device_min_sys_mem_mb"]),
                    "TF_ENABLE_AUTO_MIXED_PRECISION": str(m["tf_enable_auto_mixed_precision"])}
        optimized = "images_per_epoch" in m       # the -optimized chart (aws-samples fork of tensorpack's trainer)
        # the launcher container's mpirun line, argument for argument (templates/maskrcnn.yaml:52-117 of both charts)
        exports = ["HOROVOD_AUTOTUNE", "HOROVOD_HIERARCHICAL_ALLREDUCE=0", "HOROVOD_HIERARCHICAL_ALLGATHER=0",
                   "HOROVOD_TORUS_ALLREDUCE=0", "HOROVOD_LOG_LEVEL", "NCCL_SOCKET_IFNAME", "NCCL_DEBUG",
                   "TF_DEVICE_MIN_SYS_MEMORY_IN_MB", "TF_CPP_MIN_LOG_LEVEL", "TF_GPU_ALLOCATOR", "TF_AUTOTUNE_THRESHOLD",
                   "TF_ENABLE_AUTO_MIXED_PRECISION", "TF_ENABLE_CUBLAS_TENSOR_OP_MATH_FP32",
                   "TF_ENABLE_CUDNN_TENSOR_OP_MATH_FP32", "LD_LIBRARY_PATH", "PATH"]
        plan.mpirun_args = (["--verbose"] if optimized else []) + [
            "--output-filename", logroot, "--allow-run-as-root", "-np", str(m["gpus"]), "-bind-to", "none",
            "-map-by", "slot", "-mca", "btl_tcp_if_exclude", "lo,docker0", "-mca", "oob_tcp_if_exclude", "lo,docker0",
            "-mca", "plm_rsh_no_tree_spawn", "1"]
        for e in exports:
            plan.mpirun_args += ["-x", e]
        plan.mpirun_args += ["-mca", "pml", "ob1", "-mca", "btl", "^openib", "--display-map", "--tag-output",
                             "--timestamp-output"]
        cfg = [f"MODE_MASK={m['mode_mask']}", f"MODE_FPN={m['mode_fpn']}",
               f"DATA.BASEDIR=/{m['data_fs']}/{m['data_dir']}", f"DATA.TRAIN={m['data_train']}",
               f"DATA.VAL={m['data_val']}"]
        cmd = ["python3", str(m["train_script"]), "--logdir", f"{logroot}/train_log/maskrcnn"]
        weights = [f"TRAIN.BASE_LR={m['base_lr']}", f"BACKBONE.WEIGHTS=/{m['data_fs']}/{m['backbone_weights']}",
                   f"BACKBONE.NORM={m['backbone_norm']}"]
        if optimized:                     # templates/maskrcnn.yaml:121-135 of maskrcnn-optimized
            cmd += ["--images_per_epoch", str(m["images_per_epoch"])]
            cfg += [f"TRAIN.BATCH_SIZE_PER_GPU={m['batch_size_per_gpu']}",
                    f"TRAIN.EVAL_PERIOD={m['eval_period_in_epochs']}",
                    f"TRAIN.LR_EPOCH_SCHEDULE={m['lr_epoch_schedule']}"] + weights + [
                    "PREPROC.PREDEFINED_PADDING=True", "TRAINER=horovod"]
        else:                             # tensorpack chart (templates/maskrcnn.yaml:119-132 of maskrcnn)
            cfg += [f"TRAIN.EVAL_PERIOD={m['eval_period_in_epochs']}", f"TRAIN.STEPS_PER_EPOCH={m['steps_per_epoch']}",
                    f"TRAIN.LR_SCHEDULE={m['lr_schedule']}"] + weights + ["TRAINER=horovod"]
        if m.get("extra_config"):
            cfg += str(m["extra_config"]).split()
        plan.launcher_command = cmd + ["--config"] + cfg
        plan.script = "#!/bin/bash\n" + " ".join(shlex.quote(c) for c in plan.launcher_command) + "\n"
        return plan

    if chart.name in ("coco-data", "redpajama-data"):
        pvc = vals["pvc"]
        plan.mounts = _mounts(pvc, root)
        plan.env = {"DATA_DIR": f"{pvc['mount_path']}/{pvc['data_path']}"}
        plan.script = "\n".join(_data_script("coco" if chart.name == "coco-data" else "redpajama")) + "\n"
        return plan

    if chart.kind == "deployment":        # *-jupyter: inference + visualisation + TensorBoard
        g = vals["global"]
        for key, where in (("log_dir", "global.log_dir"),):
            if not g.get(key):
                raise TemplateError(f"{where} is required (charts/machine-learning/testing/"
                                    "maskrcnn-jupyter/templates/jupyter.yaml `required` guard)")
        if strict_required:
            if not g.get("source_cidr"):
                raise TemplateError("global.source_cidr is required")
            if not (vals.get("nginx") or {}).get("htpasswd"):
                raise TemplateError("nginx.htpasswd is required")
        fs = g["shared_fs"]
        plan.mounts = {f"/{fs}": str(root / fs)}
        logdir = f"/{fs}/{g['log_dir']}"
        plan.env = {"LOGDIR": logdir}
        plan.gpus_per_node = plan.nproc_per_node = 1
        plan.script = ("#!/bin/bash\n"
                       f"python -m b200flow.workloads.maskrcnn.visualize --logdir $LOGDIR "
                       f"--out $LOGDIR/visualization\n")
        return plan

    raise TemplateError(f"chart kind {chart.kind!r} has no renderer")
