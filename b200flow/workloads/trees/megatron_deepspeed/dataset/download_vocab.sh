#!/bin/bash
# Offline stand-in for Megatron-DeepSpeed's dataset/download_vocab.sh (called at
# examples/megatron-deepspeed/gpt2_345m/wikicorpus.yaml:23): with no network, write a
# byte-level vocab.json / merges.txt pair in GPT-2's format and id layout (50257 entries).
python - <<'PY'
from b200flow.data.tokenizer import write_byte_level_vocab
write_byte_level_vocab("gpt2-vocab.json", "gpt2-merges.txt", 50257)
print("[b200flow offline] wrote gpt2-vocab.json / gpt2-merges.txt (byte-level, 50257 entries)")
PY
