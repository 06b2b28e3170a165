"""Drop-in for Megatron-DeepSpeed `tools/preprocess_data.py`: json-lines -> mmap indexed
dataset.  Flags as used by the reference (examples/megatron-deepspeed/gpt2_345m/
wikicorpus.yaml:34-42): --input --output-prefix --vocab-file --dataset-impl mmap
--tokenizer-type GPT2BPETokenizer --merge-file --append-eod --workers N [--json-keys text].
Output: `<prefix>_<key>_document.{bin,idx}`.  For BERT-style corpora: `--split-sentences` stores one item per
sentence (`<prefix>_<key>_sentence.{bin,idx}`, document boundaries kept) and `--tokenizer-type
BertWordPieceLowerCase | BertWordPieceCase` tokenises with a local WordPiece `vocab.txt`."""
import argparse
import json
import multiprocessing as mp
import sys
import time

from b200flow.data.indexed_dataset import MMapIndexedDatasetBuilder, best_dtype
from b200flow.data.tokenizer import BertWordPieceTokenizer, GPT2BPETokenizer, split_sentences

_tok = None
_args = None


def _init(args):
    global _tok, _args
    _args = args
    _tok = _make_tokenizer(args)


def _make_tokenizer(args):
    if args.tokenizer_type.startswith("BertWordPiece"):
        return BertWordPieceTokenizer(args.vocab_file, lower_case=args.tokenizer_type.endswith("LowerCase"))
    return GPT2BPETokenizer(args.vocab_file, args.merge_file)


def _encode(line):
    line = line.strip()
    if not line:
        return {}, 0
    data = json.loads(line)
    out = {}
    for key in _args.json_keys:
        if _args.split_sentences:      # a document is a list of sentences
            sents = [ids for ids in (_tok.tokenize(s) for s in split_sentences(data[key])) if ids]
            if sents and _args.append_eod:
                sents[-1].append(_tok.eod)
            out[key] = sents
            continue
        ids = _tok.tokenize(data[key])
        if ids and _args.append_eod:
            ids.append(_tok.eod)
        out[key] = ids
    return out, len(line)


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--input", required=True)
    p.add_argument("--json-keys", nargs="+", default=["text"])
    p.add_argument("--output-prefix", required=True)
    p.add_argument("--vocab-file", required=True)
    p.add_argument("--merge-file", default=None)
    p.add_argument("--split-sentences", action="store_true")
    p.add_argument("--dataset-impl", default="mmap", choices=["mmap"])
    p.add_argument("--tokenizer-type", default="GPT2BPETokenizer")
    p.add_argument("--append-eod", action="store_true")
    p.add_argument("--workers", type=int, default=1)
    p.add_argument("--log-interval", type=int, default=1000)
    args, unknown = p.parse_known_args()
    if unknown:
        print(f"[b200flow] ignoring unknown flags: {unknown}")
    tok = _make_tokenizer(args)
    level = "sentence" if args.split_sentences else "document"
    builders = {k: MMapIndexedDatasetBuilder(f"{args.output_prefix}_{k}_{level}", best_dtype(tok.vocab_size))
                for k in args.json_keys}
    t0, nbytes, ndocs = time.time(), 0, 0
    with open(args.input, encoding="utf-8") as fin:
        if args.workers > 1:
            pool = mp.Pool(args.workers, initializer=_init, initargs=(args,))
            it = pool.imap(_encode, fin, 32)
        else:
            _init(args)
            it = map(_encode, fin)
        for doc, n in it:
            nbytes += n
            for k, ids in doc.items():
                if ids and args.split_sentences:
                    for sent in ids:
                        builders[k].add_item(sent)
                    builders[k].end_document()
                elif ids:
                    builders[k].add_item(ids)
                    builders[k].end_document()
            ndocs += 1
            if ndocs % args.log_interval == 0:
                dt = time.time() - t0
                print(f"Processed {ndocs} documents ({ndocs / dt:.1f} docs/s, {nbytes / dt / 1e6:.2f} MB/s).",
                      file=sys.stderr)
    for b in builders.values():
        b.finalize()
    dt = time.time() - t0
    print(f"done: {ndocs} documents in {dt:.1f}s ({ndocs / max(dt, 1e-9):.1f} docs/s)")


if __name__ == "__main__":
    main()
