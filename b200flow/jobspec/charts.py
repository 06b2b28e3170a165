"""Chart registry: the job templates of the reference's machine-learning charts, re-expressed
for a local launcher.

Each `Chart` carries (a) the chart's default values, restated as data from the reference's
`values.yaml` (cited per chart), (b) the *kind* of process topology the chart stands for
(PyTorchJob static / elastic, MPIJob, RayJob, Pod, Deployment) and (c) which values key
holds the workload (`train` or `process`).  Rendering lives in `render.py`; running in
`b200flow.launcher`.

SURVEY.md §2A/§2B list the charts and their keys.
"""
from __future__ import annotations

import copy
from dataclasses import dataclass, field

_COMMON_PVC = [{"name": "pv-fsx", "mount_path": "/fsx"}, {"name": "pv-efs", "mount_path": "/efs"}]
_GIT = {"repo_url": None, "branch": None, "commit": None}


@dataclass(frozen=True)
class Chart:
    name: str
    kind: str            # static | elastic | mpijob | rayjob | pod | maskrcnn | deployment
    workload_key: str    # "train" | "process" | ""
    defaults: dict
    source: str          # reference file the defaults restate
    script_name: str = "train-script.sh"
    shell: str = "bash"
    aliases: tuple = field(default_factory=tuple)


# charts/machine-learning/training/pytorchjob-distributed/values.yaml:1-24
_PYTORCHJOB_DISTRIBUTED = {
    "image": None, "backoff_limit": 2000,
    "resources": {"requests": {}, "limits": {}, "nnodes": 2, "nproc_per_node": None, "node_type": None},
    "tolerations": [], "pvc": _COMMON_PVC, "git": _GIT, "pre_script": [], "post_script": [],
    "train": {"env": [], "command": [], "args": []},
}

# charts/machine-learning/training/pytorchjob-elastic/values.yaml:1-27
_PYTORCHJOB_ELASTIC = {
    "image": None, "backoff_limit": 2000,
    "resources": {"nnodes": 1, "nproc_per_node": None, "node_type": None, "requests": {}, "limits": {}},
    "tolerations": [],
    "elastic_policy": {"rdzv_backend": "c10d", "rdzv_port": 44000, "min_replicas": 1, "max_replicas": 1},
    "pvc": _COMMON_PVC, "git": _GIT, "pre_script": [], "post_script": [],
    "train": {"env": [], "command": [], "args": []},
}

_MPI_EXPORTS = ["HOME", "HOROVOD_AUTOTUNE", "HOROVOD_HIERARCHICAL_ALLREDUCE=0",
                "HOROVOD_HIERARCHICAL_ALLGATHER=0", "HOROVOD_TORUS_ALLREDUCE=0", "HOROVOD_LOG_LEVEL",
                "NCCL_SOCKET_IFNAME", "NCCL_DEBUG", "TF_DEVICE_MIN_SYS_MEMORY_IN_MB",
                "TF_CPP_MIN_LOG_LEVEL", "TF_GPU_ALLOCATOR", "TF_AUTOTUNE_THRESHOLD",
                "TF_ENABLE_AUTO_MIXED_PRECISION", "TF_ENABLE_CUBLAS_TENSOR_OP_MATH_FP32",
                "TF_ENABLE_CUDNN_TENSOR_OP_MATH_FP32", "LD_LIBRARY_PATH", "PATH"]


def _mpirun_args() -> list[str]:
    # charts/machine-learning/training/mpijob-horovod-tensorflow-gpu/values.yaml:60-121
    a = ["--output-filename",
         '$HOME/logs/{{ .Release.Name }}-{{ date "2006-01-02-15-04-05" .Release.Time }}',
         "--allow-run-as-root", "-np",
         "{{ mul .Values.resources.gpu_nodes .Values.resources.gpus_per_node }}",
         "-bind-to", "none", "-map-by", "slot",
         "-mca", "btl_tcp_if_exclude", "lo,docker0", "-mca", "oob_tcp_if_exclude", "lo,docker0",
         "-mca", "plm_rsh_no_tree_spawn", "1"]
    for e in _MPI_EXPORTS:
        a += ["-x", e]
    a += ["-mca", "pml", "ob1", "-mca", "btl", "^openib", "--display-map", "--tag-output",
          "--timestamp-output"]
    return a


# charts/machine-learning/training/mpijob-horovod-tensorflow-gpu/values.yaml:1-59
_MPIJOB = {
    "image": None, "image_pull_policy": "IfNotPresent", "backoff_limit": 2000,
    "resources": {"gpu_nodes": 2, "gpus_per_node": 8, "gpu_instance_type": "p3dn.24xlarge"},
    "pvc": _COMMON_PVC, "git": _GIT, "pre_script": [], "post_script": [],
    "train": {"command": [], "args": []},
    "horovod": {"autotune": "1", "log_level": "INFO"},
    "nccl": {"ifname": "^lo,docker0", "debug_level": "INFO"},
    "tensorflow": {"sys_memory_mb": "4096", "autotune_threshold": "1", "cpp_min_log_level": "2",
                   "mixed_precision": "0", "gpu_allocator": "cuda_malloc_async"},
    "mpirun": {
        "env": [
            {"name": "HOME", "value": "/efs/home/{{ .Release.Name }}"},
            {"name": "HOROVOD_AUTOTUNE", "value": "{{ .Values.horovod.autotune }}"},
            {"name": "HOROVOD_LOG_LEVEL", "value": "{{ .Values.horovod.log_level }}"},
            {"name": "NCCL_SOCKET_IFNAME", "value": "{{ .Values.nccl.ifname }}"},
            {"name": "NCCL_DEBUG", "value": "{{ .Values.nccl.debug_level }}"},
            {"name": "TF_DEVICE_MIN_SYS_MEMORY_IN_MB", "value": "{{ .Values.tensorflow.sys_memory_mb }}"},
            {"name": "TF_CPP_MIN_LOG_LEVEL", "value": "{{ .Values.tensorflow.cpp_min_log_level }}"},
            {"name": "TF_GPU_ALLOCATOR", "value": "{{ .Values.tensorflow.gpu_allocator }}"},
            {"name": "TF_AUTOTUNE_THRESHOLD", "value": "{{ .Values.tensorflow.autotune_threshold }}"},
            {"name": "TF_ENABLE_AUTO_MIXED_PRECISION", "value": "{{ .Values.tensorflow.mixed_precision }}"},
            {"name": "TF_ENABLE_CUBLAS_TENSOR_OP_MATH_FP32", "value": "{{ .Values.tensorflow.mixed_precision }}"},
            {"name": "TF_ENABLE_CUDNN_TENSOR_OP_MATH_FP32", "value": "{{ .Values.tensorflow.mixed_precision }}"},
        ],
        "args": _mpirun_args(),
    },
}

# charts/machine-learning/training/raytrain/values.yaml:1-32
_RAYTRAIN = {
    "ray": {"version": "2.9.0", "dashboard": {"host": "0.0.0.0"}, "ports": [],
            "resources": {"requests": {}, "limits": {}}, "runtime_env_yaml": None},
    "image": None, "image_pull_policy": "IfNotPresent",
    "resources": {"nnodes": 1, "node_type": None, "requests": {}, "limits": {}},
    "tolerations": [], "pvc": _COMMON_PVC, "git": _GIT, "pre_script": [], "post_script": [],
    "train": {"env": [], "command": [], "args": []},
}

# charts/machine-learning/data-prep/data-process/values.yaml:1-19
_DATA_PROCESS = {
    "image": None, "pvc": _COMMON_PVC, "git": _GIT, "pre_script": [], "post_script": [],
    "process": {"env": [], "command": [], "args": []},
    "resources": {"requests": {}, "limits": {}},
}

# charts/machine-learning/data-prep/coco-data/values.yaml:1-4, redpajama-data/values.yaml:1-4
_COCO = {"pvc": {"name": "pv-fsx", "mount_path": "/fsx", "data_path": "data/coco2017"}}
_REDPAJAMA = {"pvc": {"name": "pv-fsx", "mount_path": "/fsx", "data_path": "data/redpajama"}}

# charts/machine-learning/training/maskrcnn/values.yaml:1-34
_MASKRCNN = {
    "global": {"namespace": "kubeflow", "shared_fs": "fsx", "shared_pvc": "pv-fsx"},
    "maskrcnn": {
        "name": "maskrcnn", "gpus": 16, "gpu_nodes": 2, "gpus_per_node": 8,
        "gpu_instance_type": "p3dn.24xlarge", "image": None,
        "train_script": "/tensorpack/examples/FasterRCNN/train.py",
        "data_fs": "fsx", "data_dir": "data/coco2017", "steps_per_epoch": 7500,
        "lr_schedule": "[240000,320000,360000]", "base_lr": 0.01, "eval_period_in_epochs": 1,
        "data_train": '["coco_train2017"]', "data_val": '("coco_val2017")',
        "mode_fpn": "True", "mode_mask": "True", "backbone_norm": "FreezeBN",
        "backbone_weights": "data/coco2017/pretrained-models/ImageNet-R50-AlignPadding.npz",
        "image_pull_policy": "Always", "nccl_socket_ifname": "^lo,docker0", "if_exclude": "lo,docker0",
        "tf_device_min_sys_mem_mb": 4096, "tf_enable_auto_mixed_precision": 0, "nccl_debug": "INFO",
        "horovod_autotune": 1, "horovod_log_level": "INFO",
        "extra_config": "TRAIN.CHECKPOINT_PERIOD=2", "backoff_limit": 2000,
    },
}

# charts/machine-learning/training/maskrcnn-optimized/values.yaml:1-36
_MASKRCNN_OPT = {
    "global": {"namespace": "kubeflow", "shared_fs": "fsx", "shared_pvc": "pv-fsx"},
    "maskrcnn": {
        "name": "mask-rcnn-tensorflow", "gpus": 16, "gpu_nodes": 2, "gpus_per_node": 8,
        "gpu_instance_type": "p3dn.24xlarge", "image": None,
        "train_script": "/mask-rcnn-tensorflow/MaskRCNN/train.py", "batch_size_per_gpu": 4,
        "data_fs": "fsx", "data_dir": "data/coco2017", "working_dir": "/mask-rcnn-tensorflow",
        "images_per_epoch": 120000, "lr_epoch_schedule": "[(16, 0.1), (20, 0.01), (24, None)]",
        "base_lr": 0.0015625, "eval_period_in_epochs": 1,
        "data_train": '["train2017"]', "data_val": '("val2017")',
        "mode_fpn": "True", "mode_mask": "True", "backbone_norm": "FreezeBN",
        "backbone_weights": "data/coco2017/pretrained-models/ImageNet-R50-AlignPadding.npz",
        "image_pull_policy": "Always", "extra_config": "TRAIN.GRADIENT_CLIP=0.36",
        "nccl_socket_ifname": "^lo,docker0", "if_exclude": "lo,docker0",
        "tf_device_min_sys_mem_mb": 4096, "tf_enable_auto_mixed_precision": 0, "nccl_debug": "INFO",
        "horovod_autotune": 1, "horovod_log_level": "INFO", "backoff_limit": 2000,
    },
}


def _jupyter(name: str) -> dict:
    # charts/machine-learning/testing/maskrcnn-jupyter/values.yaml:1-24 (the -optimized- twin
    # differs only in global.name)
    return {
        "global": {"name": name, "namespace": "kubeflow", "shared_fs": "fsx", "shared_pvc": "pv-fsx",
                   "source_cidr": None, "log_dir": None},
        "image": None, "image_pull_policy": "Always",
        "jupyter": {"name": "jupyter", "port": 443, "target_port": 8888, "gpu_instance_type": "g5.xlarge"},
        "tensorboard": {"name": "tensorboard", "port": 6443, "target_port": 6443, "upstream_port": 6006},
        "nginx": {"name": "nginx", "ssl_certificate": "/etc/ssl/domain.crt",
                  "ssl_certificate_key": "/etc/ssl/domain.key", "user": "tensorboard", "htpasswd": None},
    }


_T = "charts/machine-learning/training/"
_D = "charts/machine-learning/data-prep/"
_J = "charts/machine-learning/testing/"

CHARTS: dict[str, Chart] = {c.name: c for c in [
    Chart("pytorchjob-distributed", "static", "train", _PYTORCHJOB_DISTRIBUTED,
          _T + "pytorchjob-distributed/values.yaml"),
    Chart("pytorchjob-elastic", "elastic", "train", _PYTORCHJOB_ELASTIC,
          _T + "pytorchjob-elastic/values.yaml"),
    Chart("mpijob-horovod-tensorflow-gpu", "mpijob", "train", _MPIJOB,
          _T + "mpijob-horovod-tensorflow-gpu/values.yaml"),
    Chart("raytrain", "rayjob", "train", _RAYTRAIN, _T + "raytrain/values.yaml"),
    Chart("maskrcnn", "maskrcnn", "", _MASKRCNN, _T + "maskrcnn/values.yaml"),
    Chart("maskrcnn-optimized", "maskrcnn", "", _MASKRCNN_OPT, _T + "maskrcnn-optimized/values.yaml"),
    Chart("data-process", "pod", "process", _DATA_PROCESS, _D + "data-process/values.yaml",
          script_name="data-process.sh", shell="sh"),
    Chart("coco-data", "pod", "", _COCO, _D + "coco-data/values.yaml", script_name="coco-data.sh"),
    Chart("redpajama-data", "pod", "", _REDPAJAMA, _D + "redpajama-data/values.yaml",
          script_name="redpajama-data.sh"),
    Chart("maskrcnn-jupyter", "deployment", "", _jupyter("maskrcnn-jupyter"),
          _J + "maskrcnn-jupyter/values.yaml"),
    Chart("maskrcnn-optimized-jupyter", "deployment", "", _jupyter("maskrcnn-optimized-jupyter"),
          _J + "maskrcnn-optimized-jupyter/values.yaml"),
]}


def get_chart(name_or_path: str) -> Chart:
    """Accepts a bare chart name or a path ending in it (`charts/.../pytorchjob-elastic`)."""
    key = str(name_or_path).rstrip("/").split("/")[-1]
    if key not in CHARTS:
        raise KeyError(f"unknown chart {name_or_path!r}; known: {sorted(CHARTS)}")
    return CHARTS[key]


def default_values(chart: Chart) -> dict:
    return copy.deepcopy(chart.defaults)
