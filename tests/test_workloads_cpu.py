"""The reference's example values files, unchanged, through the local launcher on CPU/gloo
(BASELINE.json config 1 and the plumbing of the others at test size)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
EX = Path("/root/reference/examples")
pytestmark = pytest.mark.skipif(not EX.exists(), reason="reference tree not mounted")

TINY_GPT = ("--num-layers 2 --hidden-size 64 --num-attention-heads 4 --seq-length 64 "
            "--max-position-embeddings 64 --micro-batch-size 2 --log-interval 2 --eval-interval 1000")


def _install(release, chart, values, root, sets=(), env=None, extra=()):
    cmd = [sys.executable, "-m", "b200flow.launcher", "install", release, chart, "-f", str(values),
           "--root", str(root), "--max-restarts", "0", *extra]
    for s in sets:
        cmd += ["--set", s]
    e = dict(os.environ, **(env or {}))
    e.pop("CUDA_VISIBLE_DEVICES", None)
    e["CUDA_VISIBLE_DEVICES"] = ""          # plumbing tests run on CPU even on a GPU box
    res = subprocess.run(cmd, cwd=REPO, env=e, capture_output=True, text=True, timeout=900)
    logs = "\n".join(p.read_text(errors="replace") for p in
                     (Path(root) / ".b200flow/releases" / release / "pods").glob("*/stdout.log"))
    return res.returncode, json.loads(res.stdout.strip().splitlines()[-1]), logs


def test_accelerate_bert_mrpc_gloo_ws2(tmp_root):
    """BASELINE.json config 1."""
    rc, st, logs = _install("accel-bert", "pytorchjob-elastic", EX / "accelerate/bert-glue-mrpc/pretrain.yaml",
                            tmp_root, sets=["resources.nproc_per_node=2"],
                            env={"B2F_MODEL": "bert-tiny", "B2F_EPOCHS": "2", "B2F_TRAIN_SAMPLES": "96",
                                 "B2F_EVAL_SAMPLES": "40"})
    assert rc == 0 and st["phase"] == "Succeeded", logs[-2000:]
    home = tmp_root / "efs/home/accel-bert"
    ck = home / "checkpoints/accel-bert-worker-0"
    for ep in ("epoch_0", "epoch_1"):
        names = {p.name for p in (ck / ep).iterdir()}
        assert {"model.safetensors", "optimizer.bin", "scheduler.bin", "random_states_0.pkl",
                "random_states_1.pkl"} <= names
    log = (home / "logs/accel-bert-worker-0/pretrain.log").read_text()
    assert "epoch 1:" in log and "accuracy" in log
    assert list((home / "project/accel-bert-worker-0").rglob("metrics.jsonl"))
    assert "git clone https://github.com/huggingface/accelerate.git -> local tree" in logs


def test_accelerate_elastic_chart_on_two_nodes_rendezvous(tmp_root):
    """`pytorchjob-elastic` with two worker pods: both run `torchrun --rdzv_backend c10d` against the shared
    `PET_RDZV_ENDPOINT` and form one job (charts/machine-learning/training/pytorchjob-elastic/templates/train.yaml:58-71)."""
    rc, st, logs = _install("accel2", "pytorchjob-elastic", EX / "accelerate/bert-glue-mrpc/pretrain.yaml", tmp_root,
                            sets=["resources.nnodes=2", "resources.nproc_per_node=1", "elastic_policy.min_replicas=2",
                                  "elastic_policy.max_replicas=2"],
                            env={"B2F_MODEL": "bert-tiny", "B2F_EPOCHS": "1", "B2F_TRAIN_SAMPLES": "64",
                                 "B2F_EVAL_SAMPLES": "32"})
    assert rc == 0 and st["phase"] == "Succeeded" and sorted(st["pods"]) == ["accel2-worker-0", "accel2-worker-1"], logs[-2000:]
    assert logs.count("--nnodes 2 --nproc_per_node 1 --rdzv_id accel2 --rdzv_backend c10d") == 2
    assert logs.count("epoch 0: {'accuracy'") == 1          # one job: the main process alone prints the metrics


def test_megatron_tp2_pp2_zero1_and_resume(tmp_root):
    """W3: the TP=2 x PP=2 values file (test-size model), Megatron-DS checkpoint layout, resume."""
    values = EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml"
    env = {"B2F_TRAIN_ITERS": "4",
           "B2F_OVERRIDE_ARGS": TINY_GPT + " --save-interval 2 --eval-interval 2 --eval-iters 2"}
    sets = ["resources.nnodes=1", "resources.nproc_per_node=4"]
    rc, st, logs = _install("mds", "pytorchjob-distributed", values, tmp_root, sets, env)
    assert rc == 0 and st["phase"] == "Succeeded", logs[-3000:]
    assert "iteration        4/       4" in logs and "lm loss" in logs, logs[-4000:]
    # the values file trains in fp16: DeepSpeed's loss-scale state and skipped-step count are logged
    assert "loss scale: 65536.0" in logs and "number of skipped iterations:   0" in logs
    # --eval-interval under PP=2: forward-only pipeline schedule
    assert "validation loss at iteration 2" in logs and "validation loss at iteration 4" in logs
    ck = tmp_root / "fsx/home/mds/checkpoints/0"
    assert (ck / "latest").read_text() == "global_step4"
    names = {p.name for p in (ck / "global_step4").iterdir()}
    assert {"layer_00-model_00-model_states.pt", "layer_01-model_01-model_states.pt",
            "layer_03-model_00-model_states.pt", "mp_rank_00_model_states.pt", "mp_rank_01_model_states.pt",
            "zero_pp_rank_0_mp_rank_00_optim_states.pt", "zero_pp_rank_1_mp_rank_01_optim_states.pt"} <= names
    # resume: --load picks up `latest` and continues to iteration 6
    env2 = {"B2F_TRAIN_ITERS": "6",
            "B2F_OVERRIDE_ARGS": TINY_GPT + f" --load {ck} --save-interval 100"}
    rc, st, logs = _install("mds", "pytorchjob-distributed", values, tmp_root, sets, env2)
    assert rc == 0, logs[-3000:]
    assert "successfully loaded checkpoint" in logs and "at iteration 4" in logs
    assert "iteration        6/       6" in logs


def test_megatron_moe_flags_top2_expert_parallel(tmp_root):
    """DeepSpeed-MoE flags of the pinned Megatron-DeepSpeed (--num-experts / --topk /
    --moe-expert-parallel-size) on the DP + ZeRO-1 values file: 4 experts over EP=2, top-2 gating,
    per-expert checkpoint files."""
    values = EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml"
    moe = (" --num-experts 4 --topk 2 --moe-expert-parallel-size 2 --moe-loss-coeff 0.01 "
           "--moe-train-capacity-factor 1.0 --moe-min-capacity 4 --create-moe-param-group")
    env = {"B2F_TRAIN_ITERS": "4", "B2F_OVERRIDE_ARGS": TINY_GPT + moe}
    rc, st, logs = _install("mds-moe", "pytorchjob-distributed", values, tmp_root,
                            ["resources.nnodes=1", "resources.nproc_per_node=2"], env)
    assert rc == 0 and st["phase"] == "Succeeded", logs[-3000:]
    assert "iteration        4/       4" in logs and "unknown arguments" not in logs
    ck = tmp_root / "fsx/home/mds-moe/checkpoints/0/global_step4"
    names = {p.name for p in ck.iterdir()}
    assert {f"layer_1_expert_{e}_mp_rank_00_model_states.pt" for e in range(4)} <= names, names
    assert "expp_rank_0_mp_rank_00_optim_states.pt" in names and "expp_rank_1_mp_rank_00_optim_states.pt" in names


def test_llama_flags_on_the_tp2_pp2_dp2_topology(tmp_root):
    """BASELINE.json config 4 at test size: Llama architecture flags of the pinned Megatron-DeepSpeed (SwiGLU, RMSNorm,
    rotary positions, untied output layer, no biases, grouped-query attention) with TP2 x PP2 x DP2, ZeRO-1, sequence
    parallelism and bf16, through the unchanged TP / PP values file on 8 ranks."""
    values = EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml"
    llama = (" --swiglu --normalization rmsnorm --use-rotary-position-embeddings --untie-embeddings-and-output-weights"
             " --disable-bias-linear --num-key-value-heads 2 --ffn-hidden-size 128 --sequence-parallel --bf16"
             " --hidden-dropout 0 --attention-dropout 0 --global-batch-size 16")
    env = {"B2F_TRAIN_ITERS": "4", "B2F_OVERRIDE_ARGS": TINY_GPT + llama}
    rc, st, logs = _install("llama", "pytorchjob-distributed", values, tmp_root,
                            ["resources.nnodes=2", "resources.nproc_per_node=4"], env)
    assert rc == 0 and st["phase"] == "Succeeded" and len(st["pods"]) == 2, logs[-3000:]
    assert "iteration        4/       4" in logs and "unknown arguments" not in logs, logs[-3000:]
    assert "global batch size:    16" in logs
    # the values file saves to $CKPT_ROOT/$PET_NODE_RANK: node 0 holds the first pipeline stage, node 1 the second,
    # and both directories name their newest checkpoint
    ck = tmp_root / "fsx/home/llama/checkpoints"
    n0 = {p.name for p in (ck / "0/global_step4").iterdir()}
    n1 = {p.name for p in (ck / "1/global_step4").iterdir()}
    assert "layer_00-model_00-model_states.pt" in n0 and "layer_03-model_01-model_states.pt" in n1
    assert any("zero_pp_rank_3_mp_rank_01" in n for n in n1), n1
    assert (ck / "0/latest").read_text() == (ck / "1/latest").read_text() == "global_step4"
    # resume on both nodes
    env2 = {"B2F_TRAIN_ITERS": "6", "B2F_OVERRIDE_ARGS": TINY_GPT + llama + " --save-interval 100"}
    rc, st, logs = _install("llama", "pytorchjob-distributed", values, tmp_root,
                            ["resources.nnodes=2", "resources.nproc_per_node=4"], dict(env2), extra=["--auto-resume"])
    assert rc == 0 and "at iteration 4" in logs and "iteration        6/       6" in logs, logs[-3000:]


def test_megatron_ulysses_sequence_parallel_flag(tmp_root):
    """`--ds-sequence-parallel-size 2` (DeepSpeed-Ulysses of the pinned Megatron-DeepSpeed) on the DP + ZeRO-1 values
    file: 4 ranks = 2 replicas x 2 sequence slices; the global batch counts replicas, not ranks."""
    values = EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml"
    env = {"B2F_TRAIN_ITERS": "4", "B2F_OVERRIDE_ARGS": TINY_GPT + " --ds-sequence-parallel-size 2"}
    rc, st, logs = _install("mds-sp", "pytorchjob-distributed", values, tmp_root,
                            ["resources.nnodes=1", "resources.nproc_per_node=4"], env)
    assert rc == 0 and st["phase"] == "Succeeded", logs[-3000:]
    assert "iteration        4/       4" in logs and "unknown arguments" not in logs, logs[-3000:]
    assert "global batch size:     8" in logs          # ds_config micro batch 4 x 2 replicas (4 ranks / SP 2)


def test_wikicorpus_preprocess_then_ddp_zero1_pipeline(tmp_root):
    """KFP-style pipeline: data-process(wikicorpus) then pytorchjob-distributed (W4 -> W2)."""
    from b200flow.jobspec.values import load_yaml
    from b200flow.pipeline.runner import run_pipeline

    os.environ.update(B2F_SYNTH_DOCS="200", B2F_TRAIN_ITERS="3", B2F_OVERRIDE_ARGS=TINY_GPT,
                      CUDA_VISIBLE_DEVICES="")
    try:
        v1 = load_yaml(EX / "megatron-deepspeed/gpt2_345m/wikicorpus.yaml")
        v2 = load_yaml(EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml")
        res = run_pipeline([
            {"release_name": "mds-gpt2-345m", "chart": "charts/machine-learning/data-prep/data-process",
             "values": v1, "max_restarts": 0},
            {"release_name": "mds-gpt2-345m", "chart": "charts/machine-learning/training/pytorchjob-distributed",
             "values": v2, "set": ["resources.nnodes=1", "resources.nproc_per_node=2"], "max_restarts": 0},
        ], root=str(tmp_root), log=lambda m: None)
    finally:
        for k in ("B2F_SYNTH_DOCS", "B2F_TRAIN_ITERS", "B2F_OVERRIDE_ARGS", "CUDA_VISIBLE_DEVICES"):
            os.environ.pop(k, None)
    assert res == "Success"
    data = tmp_root / "fsx/home/mds-gpt2-345m/data/wikicorpus"
    assert (data / "gpt2_text_document.bin").stat().st_size > 0 and (data / "gpt2_text_document.idx").exists()
    log = next((tmp_root / "efs/home/mds-gpt2-345m/logs").rglob("*.log")).read_text()
    assert "iteration        3/       3" in log and "synthetic" not in log   # trained on the preprocessed corpus
    # text generation from the checkpoint that run saved (model geometry from the stored arguments)
    ck = tmp_root / "fsx/home/mds-gpt2-345m/checkpoints/0"
    res = subprocess.run([sys.executable, "-m", "b200flow.workloads.generate_gpt", "--load", str(ck), "--prompt-ids", "5", "6", "7",
                          "--max-new-tokens", "5"], cwd=REPO, env=dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=str(REPO)),
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, (res.stdout + res.stderr)[-2000:]
    ids = json.loads(res.stdout.split("token ids:")[1].strip().splitlines()[0])
    assert ids[:3] == [5, 6, 7] and len(ids) == 8


def test_ray_lightning_bert(tmp_root):
    rc, st, logs = _install("ray-bert", "raytrain", EX / "ray/lightning-bert/fine-tune.yaml", tmp_root,
                            sets=["resources.nnodes=2"],
                            env={"B2F_MODEL": "bert-tiny", "B2F_EPOCHS": "3", "B2F_TRAIN_SAMPLES": "64",
                                 "B2F_EVAL_SAMPLES": "32"})
    assert rc == 0 and st["phase"] == "Succeeded", logs[-2000:]
    out = tmp_root / "efs/home/ray-bert/ray_results/ptl-sent-classification"
    cks = sorted(p.name for p in out.glob("checkpoint_*"))
    assert len(cks) == 2                                   # CheckpointConfig(num_to_keep=2)
    assert (out / cks[0] / "checkpoint.ckpt").exists()
    assert len((out / "result.json").read_text().splitlines()) == 3
    assert "matthews_correlation" in (tmp_root / "efs/home/ray-bert/logs/ray-bert-submitter/fine-tune.log").read_text()


def test_fault_injection_restarts_the_job(tmp_root):
    values = EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml"
    cmd = [sys.executable, "-m", "b200flow.launcher", "install", "flt", "pytorchjob-distributed", "-f", str(values),
           "--root", str(tmp_root), "--max-restarts", "2", "--strict-exit", "--inject-fault",
           "rank=0,step=2,kind=exit", "--set", "resources.nnodes=1", "--set", "resources.nproc_per_node=2"]
    env = dict(os.environ, B2F_TRAIN_ITERS="3", B2F_OVERRIDE_ARGS=TINY_GPT, CUDA_VISIBLE_DEVICES="")
    res = subprocess.run(cmd, cwd=REPO, env=env, capture_output=True, text=True, timeout=900)
    st = json.loads(res.stdout.strip().splitlines()[-1])
    assert res.returncode == 0 and st["phase"] == "Succeeded" and st["restarts"] == 1, res.stdout[-2000:]
    logs = (tmp_root / ".b200flow/releases/flt/pods/flt-master-0/stdout.log").read_text()
    assert "injected fault exit at rank 0 step 2" in logs and "iteration        3/       3" in logs


def test_auto_resume_continues_from_checkpoint_after_a_fault(tmp_root):
    """The reference's values files pass --save but no --load, so a restarted job would
    start over; `--auto-resume` makes the restarted incarnation continue from the newest checkpoint."""
    values = EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml"
    cmd = [sys.executable, "-m", "b200flow.launcher", "install", "arz", "pytorchjob-distributed", "-f", str(values),
           "--root", str(tmp_root), "--max-restarts", "2", "--strict-exit", "--auto-resume", "--inject-fault",
           "rank=0,step=3,kind=exit", "--set", "resources.nnodes=1", "--set", "resources.nproc_per_node=2"]
    env = dict(os.environ, B2F_TRAIN_ITERS="5", B2F_OVERRIDE_ARGS=TINY_GPT + " --save-interval 2",
               CUDA_VISIBLE_DEVICES="")
    res = subprocess.run(cmd, cwd=REPO, env=env, capture_output=True, text=True, timeout=900)
    st = json.loads(res.stdout.strip().splitlines()[-1])
    assert res.returncode == 0 and st["phase"] == "Succeeded" and st["restarts"] == 1, res.stdout[-2000:]
    logs = (tmp_root / ".b200flow/releases/arz/pods/arz-master-0/stdout.log").read_text()
    assert "injected fault exit at rank 0 step 3" in logs
    assert "successfully loaded checkpoint" in logs and "at iteration 2" in logs      # saved at 2, died at 3
    assert "iteration        5/       5" in logs


def test_hang_is_detected_and_the_group_restarted(tmp_root):
    values = EX / "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml"
    cmd = [sys.executable, "-m", "b200flow.launcher", "install", "hng", "pytorchjob-distributed", "-f", str(values),
           "--root", str(tmp_root), "--max-restarts", "2", "--strict-exit", "--hang-timeout", "12", "--inject-fault",
           "rank=1,step=2,kind=hang", "--set", "resources.nnodes=1", "--set", "resources.nproc_per_node=2"]
    env = dict(os.environ, B2F_TRAIN_ITERS="3", B2F_OVERRIDE_ARGS=TINY_GPT, CUDA_VISIBLE_DEVICES="")
    res = subprocess.run(cmd, cwd=REPO, env=env, capture_output=True, text=True, timeout=900)
    st = json.loads(res.stdout.strip().splitlines()[-1])
    assert res.returncode == 0 and st["phase"] == "Succeeded" and st["restarts"] == 1, res.stdout[-2000:]
    logs = (tmp_root / ".b200flow/releases/hng/pods/hng-master-0/stdout.log").read_text()
    assert "injected fault hang at rank 1 step 2" in logs and "iteration        3/       3" in logs


def test_accelerate_auto_resume_from_epoch_checkpoint(tmp_root):
    values = EX / "accelerate/bert-glue-mrpc/pretrain.yaml"
    env = {"B2F_MODEL": "bert-tiny", "B2F_TRAIN_SAMPLES": "64", "B2F_EVAL_SAMPLES": "32"}
    rc, st, logs = _install("acr", "pytorchjob-elastic", values, tmp_root, sets=["resources.nproc_per_node=2"],
                            env=dict(env, B2F_EPOCHS="1"))
    assert rc == 0 and st["phase"] == "Succeeded", logs[-2000:]
    ck = tmp_root / "efs/home/acr/checkpoints/acr-worker-0"
    assert (ck / "epoch_0").is_dir()
    # second install of the same release: with --auto-resume it loads epoch_0 and only trains epoch 1
    rc, st, logs = _install("acr", "pytorchjob-elastic", values, tmp_root, sets=["resources.nproc_per_node=2"],
                            env=dict(env, B2F_EPOCHS="2"), extra=["--auto-resume"])
    assert rc == 0 and st["phase"] == "Succeeded", logs[-2000:]
    assert "Resumed from checkpoint" in logs and (ck / "epoch_1").is_dir()
    log = (tmp_root / "efs/home/acr/logs/acr-worker-0/pretrain.log").read_text()
    assert "epoch 1:" in log
