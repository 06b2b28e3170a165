"""BERT pre-training entry point (BASELINE.json config 2 as a user program) on CPU: sentence corpus -> C++ sample
map -> masked-LM / NSP batches -> Accelerate-shaped loop, checkpoint + resume, two ranks over gloo."""
import os
import re
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def _corpus(tmp_path):
    from b200flow.data.indexed_dataset import MMapIndexedDatasetBuilder

    rng = np.random.RandomState(0)
    prefix = str(tmp_path / "wiki_text_sentence")
    b = MMapIndexedDatasetBuilder(prefix, dtype=np.int32)
    for _ in range(80):
        for _ in range(int(rng.randint(2, 9))):
            b.add_item(rng.randint(5, 1000, size=int(rng.randint(4, 24))))
        b.end_document()
    b.finalize()
    return prefix


def _run(args, nproc=1, timeout=600):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=str(ROOT), OMP_NUM_THREADS="1")
    if nproc == 1:
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
            env.pop(k, None)
        cmd = [sys.executable, "-m", "b200flow.workloads.pretrain_bert", *args]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
               "--master-addr", "127.0.0.1", "--master-port", str(29800 + os.getpid() % 100),
               "-m", "b200flow.workloads.pretrain_bert", *args]
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert res.returncode == 0, (res.stdout + res.stderr)[-3000:]
    return res.stdout + res.stderr


def _losses(out):
    return [float(m) for m in re.findall(r"\| loss ([0-9.]+) \|", out)]


def test_pretrain_bert_on_sentence_corpus_checkpoints_and_resumes(tmp_path):
    prefix = _corpus(tmp_path)
    common = ["--model", "bert-tiny", "--data-prefix", prefix, "--max-seq-length", "64", "--per-device-batch", "8",
              "--lr", "2e-3", "--warmup-steps", "2", "--dropout", "0", "--log-interval", "2", "--cpu", "--mixed_precision", "no",
              "--output_dir", str(tmp_path / "ckpt")]
    out = _run(common + ["--max-steps", "8", "--checkpointing_steps", "4", "--with_tracking",
                         "--project_dir", str(tmp_path / "logs")])
    ls = _losses(out)
    # uniform random tokens: nothing to learn beyond ln(vocab) + ln 2 = 7.62 — the loss sits at that floor
    assert len(ls) == 4 and all(7.0 < v < 8.2 for v in ls), out[-1500:]
    names = {p.name for p in (tmp_path / "ckpt" / "step_4").iterdir()}
    assert {"model.safetensors", "optimizer.bin", "scheduler.bin", "random_states_0.pkl"} <= names
    assert (tmp_path / "ckpt" / "step_8").is_dir() and list((tmp_path / "logs").rglob("metrics.jsonl"))
    # resume from step 4: same data order and state -> the loss of steps 5..8 repeats
    out2 = _run(common + ["--max-steps", "8", "--resume_from_checkpoint", str(tmp_path / "ckpt" / "step_4")])
    assert "resumed from" in out2 and "at step 4" in out2
    ls2 = _losses(out2)
    assert len(ls2) == 2 and all(abs(a - b) < 2e-3 for a, b in zip(ls2, ls[2:])), (ls, ls2)


def test_pretrain_bert_two_ranks_synthetic(tmp_path):
    out = _run(["--model", "bert-tiny", "--max-seq-length", "32", "--per-device-batch", "4", "--max-steps", "4",
                "--log-interval", "2", "--cpu", "--mixed_precision", "fp16", "--gradient_accumulation_steps", "2"],
               nproc=2)
    assert len(_losses(out)) == 2 and "step       4/4" in out


def test_text_to_wordpiece_corpus_to_pretraining(tmp_path):
    """The whole BERT data path on text: WordPiece vocabulary -> sentence-level preprocessing with the Megatron tool ->
    pre-training that takes its special tokens and vocabulary size from the same vocab.txt."""
    import json

    from b200flow.data.tokenizer import BertWordPieceTokenizer, write_wordpiece_vocab

    docs = [" ".join(f"Sentence {i} of document {d} mentions topic {(i * d) % 5} again." for i in range(3 + d % 4))
            for d in range(40)]
    vocab = tmp_path / "vocab.txt"
    write_wordpiece_vocab(vocab, docs, size=300)
    (tmp_path / "train.json").write_text("\n".join(json.dumps({"text": t}) for t in docs) + "\n")
    tool = ROOT / "b200flow/workloads/trees/megatron_deepspeed/tools/preprocess_data.py"
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=str(ROOT))
    res = subprocess.run([sys.executable, str(tool), "--input", str(tmp_path / "train.json"), "--output-prefix",
                          str(tmp_path / "wiki"), "--vocab-file", str(vocab), "--tokenizer-type", "BertWordPieceLowerCase",
                          "--split-sentences"], env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    out = _run(["--model", "bert-tiny", "--data-prefix", str(tmp_path / "wiki_text_sentence"), "--vocab-file", str(vocab),
                "--max-seq-length", "48", "--per-device-batch", "8", "--max-steps", "30", "--lr", "3e-3", "--warmup-steps", "3",
                "--dropout", "0", "--log-interval", "10", "--cpu", "--mixed_precision", "no"])
    ls = _losses(out)
    tok = BertWordPieceTokenizer(vocab)
    assert len(ls) == 3 and ls[-1] < ls[0] - 0.5, out[-1500:]          # real text has structure: the loss falls
    assert ls[0] < 1.3 * (__import__("math").log((tok.vocab_size + 127) // 128 * 128) + 0.7)
