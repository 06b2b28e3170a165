"""Accelerate `save_state` / `load_state` directory layout.

The reference checkpoints with `--checkpointing_steps epoch --output_dir $CKPT_DIR`
(examples/accelerate/bert-glue-mrpc/pretrain.yaml:47-49) and Accelerate writes, per call
(SURVEY.md §5.4): `model.safetensors` (safe serialisation; `pytorch_model.bin` otherwise),
`optimizer.bin`, `scheduler.bin`, `sampler.bin`, `scaler.pt` (fp16 only),
`random_states_<rank>.pkl` and `custom_checkpoint_<i>.pkl`, inside `epoch_<n>` / `step_<n>`
directories.  This module writes and reads exactly those names.
"""
from __future__ import annotations

import pickle
import random
from pathlib import Path

import numpy as np
import torch

MODEL_NAME, MODEL_BIN = "model.safetensors", "pytorch_model.bin"
OPTIMIZER_NAME, SCHEDULER_NAME, SAMPLER_NAME, SCALER_NAME = "optimizer.bin", "scheduler.bin", "sampler.bin", "scaler.pt"


def _suffix(i: int) -> str:
    return "" if i == 0 else f"_{i}"


def _name(base: str, i: int) -> str:
    stem, dot, ext = base.rpartition(".")
    return f"{stem}{_suffix(i)}{dot}{ext}"


def save_model_weights(model, out_dir, safe_serialization: bool = True, index: int = 0) -> Path:
    """One model's weights as `model[_i].safetensors` (or `pytorch_model[_i].bin`)."""
    out = Path(out_dir)
    sd = {k: v.detach().cpu().contiguous() for k, v in model.state_dict().items()}
    if safe_serialization:
        from safetensors.torch import save_file

        # tied tensors share storage; safetensors wants each saved once
        seen, uniq = {}, {}
        for k, v in sd.items():
            key = (v.data_ptr(), tuple(v.shape))
            if key in seen:
                continue
            seen[key] = k
            uniq[k] = v
        path = out / _name(MODEL_NAME, index)
        save_file(uniq, str(path), metadata={"format": "pt"})
    else:
        path = out / _name(MODEL_BIN, index)
        torch.save(sd, path)
    return path


def save_state(out_dir, models, optimizers, schedulers, samplers, scaler, rank: int, is_main: bool,
               custom_objects=(), safe_serialization: bool = True) -> Path:
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    if is_main:
        for i, m in enumerate(models):
            save_model_weights(m, out, safe_serialization, i)
        for i, o in enumerate(optimizers):
            torch.save(o.state_dict(), out / _name(OPTIMIZER_NAME, i))
        for i, s in enumerate(schedulers):
            torch.save(s.state_dict(), out / _name(SCHEDULER_NAME, i))
        for i, s in enumerate(samplers):
            torch.save(s, out / _name(SAMPLER_NAME, i))
        if scaler is not None:
            torch.save(scaler.state_dict(), out / SCALER_NAME)
        for i, obj in enumerate(custom_objects):
            with open(out / f"custom_checkpoint_{i}.pkl", "wb") as f:
                pickle.dump(obj.state_dict() if hasattr(obj, "state_dict") else obj, f)
    states = {"random_state": random.getstate(), "numpy_random_seed": np.random.get_state(),
              "torch_manual_seed": torch.get_rng_state()}
    if torch.cuda.is_available():
        states["torch_cuda_manual_seed"] = torch.cuda.get_rng_state_all()
    with open(out / f"random_states_{rank}.pkl", "wb") as f:
        pickle.dump(states, f)
    return out


def load_state(in_dir, models, optimizers, schedulers, scaler, rank: int, custom_objects=()) -> None:
    d = Path(in_dir)
    for i, m in enumerate(models):
        st, bn = d / _name(MODEL_NAME, i), d / _name(MODEL_BIN, i)
        if st.exists():
            from safetensors.torch import load_file

            sd = load_file(str(st))
        else:
            sd = torch.load(bn, map_location="cpu")
        m.load_state_dict(sd, strict=False)
    for i, o in enumerate(optimizers):
        o.load_state_dict(torch.load(d / _name(OPTIMIZER_NAME, i), map_location="cpu", weights_only=False))
    for i, s in enumerate(schedulers):
        s.load_state_dict(torch.load(d / _name(SCHEDULER_NAME, i), map_location="cpu", weights_only=False))
    if scaler is not None and (d / SCALER_NAME).exists():
        scaler.load_state_dict(torch.load(d / SCALER_NAME, weights_only=False))
    for i, obj in enumerate(custom_objects):
        with open(d / f"custom_checkpoint_{i}.pkl", "rb") as f:
            st = pickle.load(f)
        if hasattr(obj, "load_state_dict"):
            obj.load_state_dict(st)
    rs = d / f"random_states_{rank}.pkl"
    if rs.exists():
        with open(rs, "rb") as f:
            states = pickle.load(f)
        random.setstate(states["random_state"])
        np.random.set_state(states["numpy_random_seed"])
        torch.set_rng_state(states["torch_manual_seed"])
        if torch.cuda.is_available() and "torch_cuda_manual_seed" in states:
            try:
                torch.cuda.set_rng_state_all(states["torch_cuda_manual_seed"])
            except Exception:
                pass
