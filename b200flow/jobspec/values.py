"""Helm-values handling: tolerant YAML loading, deep merge, `--set` overrides.

Precedence is Helm's: chart defaults  <-  `-f values.yaml` files (in order)  <-  `--set k=v`.
The loader tolerates the duplicated `args:` key in the reference's
examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:74-75 (last one wins, as
Helm's own YAML parser does) — SURVEY.md §2C "known typos".
"""
from __future__ import annotations

import copy
import re
from pathlib import Path
from typing import Any

import yaml


class _LastKeyWinsLoader(yaml.SafeLoader):
    """SafeLoader that keeps the last occurrence of a duplicated mapping key."""


def _construct_mapping(loader, node, deep=False):
    loader.flatten_mapping(node)
    out = {}
    for k_node, v_node in node.value:
        out[loader.construct_object(k_node, deep=True)] = loader.construct_object(v_node, deep=True)
    return out


_LastKeyWinsLoader.add_constructor(yaml.resolver.BaseResolver.DEFAULT_MAPPING_TAG, _construct_mapping)


def load_yaml(text_or_path: str | Path) -> dict:
    p = Path(str(text_or_path))
    text = p.read_text() if ("\n" not in str(text_or_path) and p.exists()) else str(text_or_path)
    data = yaml.load(text, Loader=_LastKeyWinsLoader)
    return data or {}


def deep_merge(base: dict, over: dict) -> dict:
    """Maps merge recursively; scalars and lists are replaced (Helm semantics)."""
    out = copy.deepcopy(base)
    for k, v in (over or {}).items():
        if isinstance(v, dict) and isinstance(out.get(k), dict):
            out[k] = deep_merge(out[k], v)
        else:
            out[k] = copy.deepcopy(v)
    return out


def _parse_scalar(s: str) -> Any:
    if s in ("true", "false"):
        return s == "true"
    if s in ("null", "~", ""):
        return None if s else ""
    if re.fullmatch(r"-?\d+", s):
        return int(s)
    if re.fullmatch(r"-?\d+\.\d+", s):
        return float(s)
    return s


def apply_set(values: dict, assignment: str) -> dict:
    """`--set a.b[0].c=v` / `--set a.b=v1,v2`-less single assignment (quotes stripped)."""
    key, _, raw = assignment.partition("=")
    if not _:
        raise ValueError(f"--set needs key=value, got {assignment!r}")
    val = _parse_scalar(raw.strip().strip('"').strip("'")) if not (raw.startswith('"') or raw.startswith("'")) \
        else raw.strip().strip('"').strip("'")
    out = copy.deepcopy(values)
    cur: Any = out
    parts = re.findall(r'"[^"]+"|[^.\[\]]+|\[\d+\]', key.replace("\\.", "\0"))
    parts = [p.replace("\0", ".").strip('"') for p in parts]
    for i, part in enumerate(parts):
        last = i == len(parts) - 1
        if part.startswith("["):
            idx = int(part[1:-1])
            while len(cur) <= idx:
                cur.append({} if not last else None)
            if last:
                cur[idx] = val
            else:
                cur = cur[idx]
            continue
        nxt_is_index = (not last) and parts[i + 1].startswith("[")
        if last:
            cur[part] = val
        else:
            if part not in cur or cur[part] is None:
                cur[part] = [] if nxt_is_index else {}
            cur = cur[part]
    return out


def gpus_per_node(values: dict) -> int | None:
    """`resources.limits."nvidia.com/gpu"` (examples/accelerate/bert-glue-mrpc/pretrain.yaml:4-7)."""
    res = values.get("resources") or {}
    for sect in ("limits", "requests"):
        v = (res.get(sect) or {}).get("nvidia.com/gpu")
        if v is not None:
            return int(v)
    if res.get("gpus_per_node") is not None:
        return int(res["gpus_per_node"])
    return None
