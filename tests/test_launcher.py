"""Local launcher: env contract per job kind, path remap, restart budget, pipeline semantics."""
import json
import os
import sys
from pathlib import Path


from b200flow.jobspec.render import render_job
from b200flow.launcher import runner as R
from b200flow.pipeline.runner import run_pipeline


def _vals(cmd, **extra):
    v = {"train": {"env": [{"name": "HOME", "value": "/efs/home/{{ .Release.Name }}"}],
                   "command": ["bash -c"], "args": [json.dumps(cmd)]}}
    v.update(extra)
    return v


def _run(chart, rel, vals, tmp_root, sets=(), **kw):
    plan = render_job(chart, rel, [vals], sets, root=tmp_root)
    rc = R.Runner(plan, tmp_root, **kw).run(timeout=120)
    return rc, R.logs(tmp_root, rel), R.status(tmp_root, rel)


def test_static_env_contract_and_path_remap(tmp_root):
    cmd = "echo N=$PET_NNODES R=$PET_NODE_RANK P=$PET_NPROC_PER_NODE H=$HOSTNAME M=$PET_MASTER_ADDR; echo hi > /fsx/out.$PET_NODE_RANK; echo $HOME"
    rc, logs, st = _run("pytorchjob-distributed", "s1", _vals(cmd, resources={"nnodes": 2, "nproc_per_node": 3}), tmp_root)
    assert rc == 0 and st["phase"] == "Succeeded"
    assert "N=2 R=0 P=3 H=s1-master-0 M=127.0.0.1" in logs["s1-master-0"]
    assert "N=2 R=1 P=3 H=s1-worker-0" in logs["s1-worker-0"]
    assert (tmp_root / "fsx" / "out.0").exists() and (tmp_root / "fsx" / "out.1").exists()
    assert str(tmp_root / "efs" / "home" / "s1") in logs["s1-master-0"]
    assert "Training script done" in logs["s1-master-0"]


def test_elastic_and_mpirun_and_ray_env(tmp_root):
    rc, logs, _ = _run("pytorchjob-elastic", "e1", _vals("echo ID=$PET_RDZV_ID EP=$PET_RDZV_ENDPOINT B=$PET_RDZV_BACKEND"),
                       tmp_root)
    assert rc == 0 and "ID=e1 EP=127.0.0.1:" in logs["e1-worker-0"] and "B=c10d" in logs["e1-worker-0"]
    v = {"train": {"command": ["bash -c"], "args": [json.dumps("echo RANK=$OMPI_COMM_WORLD_RANK/$OMPI_COMM_WORLD_SIZE L=$OMPI_COMM_WORLD_LOCAL_RANK A=$HOROVOD_AUTOTUNE")]},
         "resources": {"gpu_nodes": 1, "gpus_per_node": 3}}
    rc, logs, _ = _run("mpijob-horovod-tensorflow-gpu", "m1", v, tmp_root)
    assert rc == 0 and len(logs) == 3
    assert any("RANK=2/3 L=2 A=1" in t for t in logs.values())
    rc, logs, _ = _run("raytrain", "r1", _vals("echo W=$B2F_RAY_WORKERS A=$RAY_ADDRESS", resources={"nnodes": 2}), tmp_root)
    assert rc == 0 and "W=2 A=local" in logs["r1-submitter"]


def test_restart_budget_and_strict_exit(tmp_root):
    # fails on the first incarnation only -> one restart, then success
    cmd = "if [ $B2F_RESTART_COUNT = 0 ]; then exit 3; fi; echo recovered"
    plan = render_job("pytorchjob-distributed", "f1", [_vals(cmd, resources={"nnodes": 1})], root=tmp_root,
                      strict_exit=True)
    run = R.Runner(plan, tmp_root)
    assert run.run(timeout=60) == 0 and run.restarts == 1
    assert "recovered" in R.logs(tmp_root, "f1")["f1-master-0"]
    # budget exhausted -> failure with the command's status
    plan = render_job("pytorchjob-distributed", "f2", [_vals("exit 5", resources={"nnodes": 1}, backoff_limit=1)],
                      root=tmp_root, strict_exit=True)
    run = R.Runner(plan, tmp_root)
    assert run.run(timeout=60) == 5 and run.restarts == 1
    assert R.status(tmp_root, "f2")["phase"] == "Failed"
    # reference semantics (no strict_exit): a failing command is masked, the pod "succeeds"
    plan = render_job("pytorchjob-distributed", "f3", [_vals("false | tee /dev/null", resources={"nnodes": 1})], root=tmp_root)
    assert R.Runner(plan, tmp_root).run(timeout=60) == 0


def test_offline_git_and_pip_shims(tmp_root):
    v = _vals("ls examples/complete_nlp_example.py && pwd",
              git={"repo_url": "https://github.com/huggingface/accelerate.git", "branch": "main", "commit": "abc"},
              pre_script=["pip3 install transformers==4.38.1"], resources={"nnodes": 1})
    rc, logs, _ = _run("pytorchjob-distributed", "g1", v, tmp_root)
    txt = logs["g1-master-0"]
    assert rc == 0 and "examples/complete_nlp_example.py" in txt and "skipped: pip3 install" in txt
    assert "/tmp/g1-master-0" in txt          # cwd is $HOME/tmp/$HOSTNAME
    assert not (tmp_root / "efs/home/g1/tmp/g1-master-0").exists()   # clone dir removed at the end


def test_pipeline_stops_at_first_failure_and_always_uninstalls(tmp_root):
    ok = {"release_name": "p-ok", "chart": "data-process",
          "values": {"process": {"command": ["echo"], "args": ["step1"]}}}
    bad = {"release_name": "p-bad", "chart": "data-process", "strict_exit": True, "max_restarts": 0,
           "values": {"process": {"command": ["bash -c"], "args": ["'exit 9'"]}}}
    never = {"release_name": "p-never", "chart": "data-process",
             "values": {"process": {"command": ["echo"], "args": ["nope"]}}}
    msgs = []
    assert run_pipeline([ok, bad, never], root=str(tmp_root), log=msgs.append) == "Failure"
    assert R.status(tmp_root, "p-ok")["phase"] == "Uninstalled"
    assert R.status(tmp_root, "p-bad")["phase"] == "Uninstalled"
    assert R.status(tmp_root, "p-never") == {}
    assert any("uninstalled p-bad" in m for m in msgs)
    assert run_pipeline([ok], root=str(tmp_root), log=msgs.append) == "Success"


def test_data_charts_run_offline_and_are_idempotent(tmp_root):
    """`coco-data` and `redpajama-data` (charts/machine-learning/data-prep/*): offline they write synthetic
    stand-ins in the layout the training charts expect; like the reference's pod, a second run leaves an
    existing data directory alone (coco-data.yaml:13 exits when the directory exists)."""
    import json
    import subprocess
    import sys

    v = tmp_root / "empty.yaml"
    v.write_text("{}\n")

    def install(release, chart):
        cmd = [sys.executable, "-m", "b200flow.launcher", "install", release, chart, "-f", str(v),
               "--root", str(tmp_root), "--max-restarts", "0"]
        res = subprocess.run(cmd, cwd=Path(__file__).resolve().parent.parent, capture_output=True, text=True,
                             timeout=600, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
        return res.returncode, json.loads(res.stdout.strip().splitlines()[-1])

    rc, st = install("d-coco", "coco-data")
    assert rc == 0 and st["phase"] == "Succeeded" and st["kind"] == "pod"
    coco = tmp_root / "fsx/data/coco2017"
    assert {"train2017", "val2017", "test2017", "annotations", "pretrained-models"} <= {p.name for p in coco.iterdir()}
    assert {p.name for p in (coco / "pretrained-models").iterdir()} == {"ImageNet-R50-AlignPadding.npz",
                                                                        "ImageNet-R101-AlignPadding.npz"}
    assert not list(coco.parent.glob("coco2017.*"))           # the script removed its download directory
    marker = coco / "annotations" / "keep.me"
    marker.write_text("x")
    rc, st = install("d-coco", "coco-data")
    assert rc == 0 and marker.exists()            # untouched on the second run

    rc, st = install("d-rp", "redpajama-data")
    assert rc == 0 and st["phase"] == "Succeeded"
    jl = tmp_root / "fsx/data/redpajama/data.jsonl"
    first = json.loads(jl.read_text().splitlines()[0])
    assert isinstance(first.get("text"), str) and len(first["text"]) > 0


def test_smoke_stage_data_attach_commands(tmp_path):
    """The reference's manual cluster checks (eks-cluster/tests/test-gpu*.yaml) and volume utilities
    (eks-cluster/utils/stage-data.yaml, attach-pvc.yaml) as launcher commands."""
    import json
    import subprocess
    import sys
    from pathlib import Path

    repo = Path(__file__).resolve().parent.parent
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    run = lambda *a: subprocess.run([sys.executable, "-m", "b200flow.launcher", *a, "--root", str(tmp_path)],  # noqa: E731
                                    cwd=repo, env=env, capture_output=True, text=True, timeout=300)
    r = run("smoke", "--gpus", "2")
    rep = json.loads(r.stdout)
    assert r.returncode == 0 and rep["ok"], r.stdout + r.stderr
    assert rep["collective"]["backend"] == "gloo" and all(x["allreduce_ok"] for x in rep["collective"]["ranks"])
    assert rep["mounts"]["/fsx"]["writable"] and rep["mounts"]["/efs"]["writable"]
    src = tmp_path / "bucket" / "my-bucket" / "ml-platform" / "data"
    src.mkdir(parents=True)
    (src / "a.txt").write_text("hello")
    (src / "sub").mkdir()
    (src / "sub" / "b.bin").write_bytes(b"\\0" * 10)
    env["B2F_S3_ROOT"] = str(tmp_path / "bucket")
    r = run("stage-data", "s3://my-bucket/ml-platform/data", "/efs/data/coco2017")
    out = json.loads(r.stdout)
    assert r.returncode == 0 and out["files"] == 2 and (tmp_path / "efs/data/coco2017/sub/b.bin").exists()
    assert run("stage-data", str(src), "/tmp/elsewhere").returncode != 0       # only the shared mounts
    r = run("attach")
    listing = json.loads(r.stdout)
    assert "/efs/data/coco2017/a.txt" in listing["/efs"]["entries"]


def test_install_no_wait_then_status_logs_uninstall(tmp_path):
    """`helm install` without --wait returns at once; the job is followed with status / logs and stopped by uninstall."""
    import json
    import subprocess
    import sys
    import time
    from pathlib import Path

    repo = Path(__file__).resolve().parent.parent
    v = tmp_path / "v.yaml"
    v.write_text("process:\n  command: [bash, -c]\n  args:\n    - \"'for i in $(seq 1 600); do echo tick $i; sleep 0.2; done'\"\n")
    run = lambda *a: subprocess.run([sys.executable, "-m", "b200flow.launcher", *a, "--root", str(tmp_path)], cwd=repo,  # noqa: E731
                                    capture_output=True, text=True, timeout=120)
    t0 = time.time()
    r = run("install", "bg", "data-process", "-f", str(v), "--no-wait")
    assert r.returncode == 0 and time.time() - t0 < 60, r.stdout + r.stderr
    st = json.loads(r.stdout.strip().splitlines()[-1])
    assert st["phase"] == "Running" and st["pods"] and st.get("supervisor_pid")
    time.sleep(1.5)
    assert "tick" in run("logs", "bg", "--tail", "3").stdout
    assert json.loads(run("status", "bg").stdout)["phase"] == "Running"
    follower = subprocess.Popen([sys.executable, "-m", "b200flow.launcher", "logs", "bg", "-f", "--tail", "1", "--root", str(tmp_path)],
                                cwd=repo, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    time.sleep(2.0)
    assert run("uninstall", "bg").returncode == 0
    out, _ = follower.communicate(timeout=30)                     # the follower ends when the release stops
    assert follower.returncode == 0 and out.count("[bg-pod] tick") >= 3, out[-500:]
    time.sleep(0.5)
    st2 = json.loads(run("status", "bg").stdout)
    assert st2["phase"] == "Uninstalled"
    listed = json.loads(run("list").stdout)
    assert [(r["name"], r["chart"], r["phase"]) for r in listed] == [("bg", "data-process", "Uninstalled")]
    n1 = run("logs", "bg").stdout.count("tick")
    time.sleep(1.0)
    assert run("logs", "bg").stdout.count("tick") == n1          # nothing is running any more


def test_elastic_membership_changes_between_min_and_max_replicas(tmp_root):
    """`elastic_policy.min_replicas < max_replicas`: the job starts on one worker pod, is scaled out to two while it
    runs (the pods re-rendezvous through torchrun's c10d backend and the world doubles), then scaled back in; the
    training continues across every change of membership and the release succeeds
    (charts/machine-learning/training/pytorchjob-elastic/templates/train.yaml:58-71)."""
    import re
    import threading
    import time

    stub = Path(__file__).resolve().parent / "elastic_train_stub.py"
    prog = tmp_root / "progress.txt"
    cmd = (f"torchrun --nnodes $PET_NNODES --nproc_per_node 1 --rdzv_id $PET_RDZV_ID --rdzv_backend $PET_RDZV_BACKEND "
           f"--rdzv_endpoint $PET_RDZV_ENDPOINT --rdzv_conf last_call_timeout=2,join_timeout=120 --max_restarts 10 "
           f"--monitor_interval 0.5 {stub} {prog} 10000")
    vals = _vals(cmd, resources={"nnodes": 1, "nproc_per_node": 1})
    vals["elastic_policy"] = {"rdzv_backend": "c10d", "rdzv_port": 44000, "min_replicas": 1, "max_replicas": 2}
    plan = render_job("pytorchjob-elastic", "el", [vals], root=tmp_root, strict_exit=True)
    assert plan.env is not None
    run = R.Runner(plan, tmp_root, max_restarts=5)
    out = {}
    th = threading.Thread(target=lambda: out.setdefault("rc", run.run(timeout=400)))
    th.start()

    def worlds():
        txt = "".join(R.logs(tmp_root, "el").values())
        return [int(m) for m in re.findall(r"step \d+ world (\d+)", txt)]

    def wait_for(pred, secs):
        end = time.time() + secs
        while time.time() < end and not pred():
            time.sleep(0.5)
        return pred()

    try:
        assert wait_for(lambda: len(worlds()) >= 3, 120), R.logs(tmp_root, "el")
        assert set(worlds()) == {1}
        assert "PET_NNODES" not in "".join(R.logs(tmp_root, "el").values())
        assert R.scale(tmp_root, "el", 5)                                   # clamped to max_replicas = 2
        assert wait_for(lambda: 2 in worlds(), 180), R.logs(tmp_root, "el")
        st = R.status(tmp_root, "el")
        assert st["desired_replicas"] == 2 and st["replicas"] == 2 and st["membership_changes"] >= 1
        n2 = len(worlds())
        assert R.scale(tmp_root, "el", 1)
        assert wait_for(lambda: 1 in worlds()[n2:], 180), R.logs(tmp_root, "el")
        steps = [int(m) for m in re.findall(r"step (\d+) world", R.logs(tmp_root, "el")["el-worker-0"])]
        assert steps == sorted(steps) and len(set(steps)) >= len(steps) - 4     # progress carried across worlds
        Path(str(prog) + ".stop").write_text("1")                           # let it finish
    finally:
        if not wait_for(lambda: not th.is_alive(), 120):
            run.stop()
        th.join(timeout=60)
    assert out.get("rc") == 0, (out, R.logs(tmp_root, "el"))
    assert R.status(tmp_root, "el")["phase"] == "Succeeded"


def test_pipeline_pending_bound_and_phase_polling(tmp_root, monkeypatch):
    """KFP `wait_for_pods` (kfp/components/src/helm-charts-component/helm_charts_component.py:135-197): a release
    whose GPUs are held by another running release stays Pending for at most `pod_error_timeout`; a running one is
    polled every `pod_check_secs`; `pod_complete_timeout` bounds the run."""
    import threading
    import time

    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "0,1")            # a two-GPU box as far as the scheduler is concerned
    hold = render_job("pytorchjob-distributed", "holder",
                      [_vals("sleep 30", resources={"nnodes": 1, "limits": {"nvidia.com/gpu": 2}})], root=tmp_root)
    holder = R.Runner(hold, tmp_root)
    th = threading.Thread(target=lambda: holder.run(timeout=60))
    th.start()
    try:
        end = time.time() + 30
        while time.time() < end and R.status(tmp_root, "holder").get("phase") != "Running":
            time.sleep(0.1)
        assert R.status(tmp_root, "holder")["gpus_claimed"] == 2
        msgs = []
        want_gpu = {"release_name": "p-pend", "chart": "pytorchjob-distributed", "pod_error_timeout": 2, "pod_check_secs": 1,
                    "values": _vals("echo ran", resources={"nnodes": 1, "limits": {"nvidia.com/gpu": 1}})}
        t0 = time.time()
        assert run_pipeline([want_gpu], root=str(tmp_root), log=msgs.append) == "Failure"
        assert 2 <= time.time() - t0 < 20
        assert any("phase: Pending" in m for m in msgs) and "ran" not in "".join(R.logs(tmp_root, "p-pend").values())
        cpu_only = {"release_name": "p-cpu", "chart": "data-process", "pod_check_secs": 0.2,
                    "values": {"process": {"command": ["bash -c"], "args": ["'sleep 1; echo fine'"]}}}
        msgs.clear()
        assert run_pipeline([cpu_only], root=str(tmp_root), log=msgs.append) == "Success"     # needs no GPU: admitted at once
        assert sum("phase: Running" in m for m in msgs) >= 2                                   # polled while running
    finally:
        holder.stop()
        th.join(timeout=30)
    slow = {"release_name": "p-slow", "chart": "data-process", "pod_complete_timeout": 1, "max_restarts": 0,
            "values": {"process": {"command": ["sleep"], "args": ["30"]}}}
    t0 = time.time()
    assert run_pipeline([slow], root=str(tmp_root), log=lambda m: None) == "Failure" and time.time() - t0 < 25


def test_more_ranks_than_the_pods_gpu_limit_fails_before_spawning(tmp_root, monkeypatch):
    """A pod only sees the GPUs of its `nvidia.com/gpu` limit; nproc_per_node above it is refused with the --set
    that fixes it, instead of dying in NCCL after the pre-script (what the reference's 4-GPU values files did when
    only nproc_per_node was raised to 8)."""
    monkeypatch.setattr(R, "visible_gpu_count", lambda: 8)
    vals = _vals("echo should-not-run", resources={"nnodes": 1, "nproc_per_node": 8, "limits": {"nvidia.com/gpu": 4}})
    rc, logs, st = _run("pytorchjob-distributed", "lim", vals, tmp_root)
    assert rc == 2 and st["phase"] == "Failed" and "nvidia\\.com/gpu=8" in st["message"]
    assert not any("should-not-run" in v for v in logs.values())
    rc, logs, st = _run("pytorchjob-distributed", "lim2", vals, tmp_root, sets=[r"resources.limits.nvidia\.com/gpu=8"])
    assert st.get("message") is None and "should-not-run" in "".join(logs.values())


def test_readme_command_lines_of_the_reference_work_with_the_cli_swapped_in(tmp_path):
    """`helm install --debug <release> ./charts/machine-learning/training/<chart>/ -f <values> -n <namespace>`,
    `helm uninstall <release> -n <namespace>`, `kubectl logs -f <pod>` (examples/accelerate/bert-glue-mrpc/README.md:23-34):
    the same argument lists are accepted by `python -m b200flow.launcher`."""
    import subprocess

    repo = Path(__file__).resolve().parent.parent
    vals = tmp_path / "v.yaml"
    vals.write_text(json.dumps(_vals("echo hello-from-$HOSTNAME", resources={"nnodes": 1, "nproc_per_node": 1})))
    run = lambda *a: subprocess.run([sys.executable, "-m", "b200flow.launcher", *a, "--root", str(tmp_path)],  # noqa: E731
                                    cwd=repo, capture_output=True, text=True, timeout=300)
    r = run("install", "--debug", "accel-bert", "./charts/machine-learning/training/pytorchjob-distributed/",
            "-f", str(vals), "-n", "kubeflow-user-example-com")
    assert r.returncode == 0 and json.loads(r.stdout.strip().splitlines()[-1])["phase"] == "Succeeded", r.stderr[-800:]
    assert '"chart": "pytorchjob-distributed"' in r.stderr                 # --debug prints the plan
    r = run("logs", "accel-bert", "-n", "kubeflow-user-example-com")
    assert "hello-from-accel-bert-master-0" in r.stdout
    r = run("logs", "-f", "pytorchjob-accel-bert-master-0", "-n", "kubeflow-user-example-com")      # a pod name, as kubectl takes
    assert "hello-from-accel-bert-master-0" in r.stdout and r.returncode == 0
    r = run("uninstall", "accel-bert", "-n", "kubeflow-user-example-com")
    assert r.returncode == 0 and "uninstalled" in r.stdout
