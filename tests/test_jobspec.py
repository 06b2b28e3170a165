"""Job-spec layer: template subset, values handling, rendering of every reference values file."""
import datetime as dt
from pathlib import Path

import pytest

from b200flow.jobspec import values as V
from b200flow.jobspec.charts import CHARTS, get_chart
from b200flow.jobspec.render import render_job
from b200flow.jobspec.template import TemplateError, render

REF = Path("/root/reference")
EX = REF / "examples"
needs_ref = pytest.mark.skipif(not EX.exists(), reason="reference tree not mounted")
T0 = dt.datetime(2026, 1, 2, 3, 4, 5)


def test_template_subset():
    vals = {"resources": {"gpu_nodes": 2, "gpus_per_node": 8}, "horovod": {"autotune": "1"},
            "items": ["a", "b"], "git": {"repo_url": None}}
    assert render("{{ mul .Values.resources.gpu_nodes .Values.resources.gpus_per_node }}", vals, "r") == "16"
    assert render("/efs/home/{{ .Release.Name }}", vals, "rel") == "/efs/home/rel"
    assert render('x-{{ date "2006-01-02-15-04-05" .Release.Time }}', vals, "r", release_time=T0) == \
        "x-2026-01-02-03-04-05"
    assert render("{{- range .Values.items }}\n{{ . }} \\\n{{- end }}\nend", vals, "r") == "\na \\\nb \\\nend"
    assert render("{{ if .Values.git.repo_url }}yes{{ else }}no{{ end }}", vals, "r") == "no"
    assert render("{{ if gt (int .Values.resources.gpu_nodes) 1 }}w={{ sub .Values.resources.gpu_nodes 1 }}{{ end }}",
                  vals, "r") == "w=1"
    assert render("{{ tpl .Values.horovod.autotune . }}|{{ .Values.missing | default \"d\" }}", vals, "r") == "1|d"
    assert render("{{ add 1 2 3 }} {{ .Chart.AppVersion }}", vals, "r", chart={"AppVersion": "9"}) == "6 9"
    with pytest.raises(TemplateError):
        render('{{ required "need it" .Values.git.repo_url }}', vals, "r")


def test_values_merge_set_and_duplicate_keys():
    base = {"a": {"b": 1, "c": [1, 2]}, "d": 3}
    assert V.deep_merge(base, {"a": {"c": [9]}, "e": 1}) == {"a": {"b": 1, "c": [9]}, "d": 3, "e": 1}
    v = V.apply_set(base, "a.b=7")
    v = V.apply_set(v, 'resources.limits."nvidia\\.com/gpu"=4')
    assert v["a"]["b"] == 7 and v["resources"]["limits"]["nvidia.com/gpu"] == 4
    assert V.load_yaml("train:\n  args: [1]\n  args: [2]\n")["train"]["args"] == [2]
    assert V.gpus_per_node({"resources": {"limits": {"nvidia.com/gpu": 4}}}) == 4


@needs_ref
@pytest.mark.parametrize("chart,rel,kind,nnodes,nproc", [
    ("pytorchjob-elastic", "accelerate/bert-glue-mrpc/pretrain.yaml", "elastic", 1, 4),
    ("pytorchjob-distributed", "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml", "static", 2, 4),
    ("pytorchjob-distributed", "megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml", "static", 2, 4),
    ("data-process", "megatron-deepspeed/gpt2_345m/wikicorpus.yaml", "pod", 1, 1),
    ("raytrain", "ray/lightning-bert/fine-tune.yaml", "rayjob", 1, 1),
    ("mpijob-horovod-tensorflow-gpu", "maskrcnn/train-maskrcnn-tensorpack.yaml", "mpijob", 2, 8),
    ("mpijob-horovod-tensorflow-gpu", "maskrcnn/train-maskrcnn-aws.yaml", "mpijob", 2, 8),
])
def test_reference_values_files_render(chart, rel, kind, nnodes, nproc, tmp_path):
    plan = render_job(chart, "rel1", [EX / rel], root=tmp_path, when=T0)
    assert (plan.kind, plan.nnodes, plan.nproc_per_node) == (kind, nnodes, nproc)
    lines = plan.script.splitlines()
    assert lines[0] == "#!/bin/bash"
    done = '&& echo "Processing script done"' if kind == "pod" else '&& echo "Training script done"'
    assert done in lines                                     # data-process.yaml:38 vs train.yaml:38
    assert plan.env["HOME"].endswith("/home/rel1")           # tpl of {{ .Release.Name }}
    if kind in ("static", "elastic", "pod", "rayjob"):
        i = lines.index(done)
        assert lines[i - 1].endswith("\\")                   # command/args joined by continuations
    if kind == "mpijob":
        assert plan.mpirun_np == 16 and plan.mpirun_args[3:5] == ["-np", "16"]
        assert plan.mpirun_args[1] == "$HOME/logs/rel1-2026-01-02-03-04-05"
        assert plan.env["NCCL_SOCKET_IFNAME"] == "^lo,docker0"


@needs_ref
def test_script_semantics_and_typos(tmp_path):
    plan = render_job("pytorchjob-distributed", "gpt", [EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml"],
                      root=tmp_path)
    s = plan.script
    # clone block, exported GIT_CLONE_DIR, cleanup at the end; `$$OUTPUT_LOG` typo kept verbatim
    assert "GIT_CLONE_DIR=$HOME/tmp/$HOSTNAME" in s and s.rstrip().endswith("rm -rf $GIT_CLONE_DIR")
    assert "git reset --hard a9856ce0e75dbe69c96d4e241e8a191b344118d7" in s
    assert "$$OUTPUT_LOG" in s
    assert plan.backoff_limit == 2000
    strict = render_job("pytorchjob-distributed", "gpt", [EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml"],
                        root=tmp_path, strict_exit=True)
    assert "set -o pipefail" in strict.script and strict.script.rstrip().endswith("exit $B2F_STATUS")


def test_chart_defaults_and_special_charts(tmp_path):
    assert set(CHARTS) >= {"pytorchjob-distributed", "pytorchjob-elastic", "mpijob-horovod-tensorflow-gpu",
                           "maskrcnn", "maskrcnn-optimized", "raytrain", "data-process", "coco-data",
                           "redpajama-data", "maskrcnn-jupyter", "maskrcnn-optimized-jupyter"}
    assert get_chart("charts/machine-learning/training/pytorchjob-elastic/").name == "pytorchjob-elastic"
    p = render_job("pytorchjob-elastic", "e", [], root=tmp_path)
    assert p.elastic == {"rdzv_backend": "c10d", "rdzv_port": 44000, "min_replicas": 1, "max_replicas": 1,
                         "rdzv_id": "e"}
    m = render_job("maskrcnn", "m", [], root=tmp_path, when=T0)
    assert m.mpirun_np == 16 and m.launcher_command[:2] == ["python3", "/tensorpack/examples/FasterRCNN/train.py"]
    assert "TRAIN.LR_SCHEDULE=[240000,320000,360000]" in m.launcher_command
    assert "/fsx/logs/m-2026-01-02-03-04-05/train_log/maskrcnn" in m.launcher_command
    o = render_job("maskrcnn-optimized", "o", [], root=tmp_path, sets=["maskrcnn.gpus=8"])
    assert "TRAIN.BATCH_SIZE_PER_GPU=4" in o.launcher_command and "TRAIN.GRADIENT_CLIP=0.36" in o.launcher_command
    assert o.mpirun_np == 8
    c = render_job("coco-data", "c", [], root=tmp_path)
    assert c.env["DATA_DIR"] == "/fsx/data/coco2017" and "already exists" in c.script
    with pytest.raises(TemplateError):
        render_job("maskrcnn-jupyter", "j", [], root=tmp_path)
    j = render_job("maskrcnn-jupyter", "j", [], root=tmp_path, sets=["global.log_dir=logs/run1"])
    assert j.env["LOGDIR"] == "/fsx/logs/run1"


def test_template_variables_with_and_keyed_range():
    """The Go-template features the reference's chart templates use beyond the values files: variables with block
    scope (`:=` declares, `=` assigns the enclosing one), `range $v :=` / `range $k, $v :=` (maps in sorted key order),
    `with`, `$` for the root inside loops (charts/machine-learning/training/pytorchjob-distributed/templates/train.yaml:86-98,126-133)."""
    from b200flow.jobspec.template import TemplateError, render

    vals = {"pvc": [{"name": "efs"}, {"name": "fsx"}], "requests": {"nvidia.com/gpu": 4, "cpu": 2},
            "tol": [{"key": "{{ .Values.k }}", "operator": "Exists"}], "k": "gpu", "empty": [], "opt": {"x": 1}}
    # a counter declared outside the loop and advanced inside it (`$pv_index = add $pv_index 1`)
    t = "{{- $i := 1 }}{{- range $pv := .Values.pvc }}[pv-{{ $i }}:{{ $pv.name }}]{{- $i = add $i 1 }}{{- end }} next={{ $i }}"
    assert render(t, vals, "r") == "[pv-1:efs][pv-2:fsx] next=3"
    # a declaration inside the block shadows and does not leak
    t = "{{ $x := 1 }}{{ if true }}{{ $x := 2 }}{{ $x }}{{ end }}{{ $x }}"
    assert render(t, vals, "r") == "21"
    # maps iterate in sorted key order; key and value variables
    t = "{{ range $k, $v := .Values.requests }}{{ $k }}={{ $v }};{{ end }}"
    assert render(t, vals, "r") == "cpu=2;nvidia.com/gpu=4;"
    # `$` is the root inside a loop whose dot is the element; tpl evaluates a value as a template
    t = "{{ range $v := .Values.tol }}{{ tpl $v.key $ }}/{{ $.Release.Name }}/{{ $v.operator }}{{ end }}"
    assert render(t, vals, "rel") == "gpu/rel/Exists"
    # index variable over a list, else branch of an empty range, with / else
    assert render("{{ range $i, $p := .Values.pvc }}{{ $i }}{{ $p.name }}{{ end }}", vals, "r") == "0efs1fsx"
    assert render("{{ range .Values.empty }}x{{ else }}none{{ end }}", vals, "r") == "none"
    assert render("{{ with .Values.opt }}{{ .x }}{{ else }}no{{ end }}|{{ with .Values.missing }}y{{ else }}no{{ end }}", vals, "r") == "1|no"
    with pytest.raises(TemplateError):
        render("{{ $undeclared = 1 }}", vals, "r")
    with pytest.raises(TemplateError):
        render("{{ $nope }}", vals, "r")
