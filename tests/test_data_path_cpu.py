"""CPU tier: the real-data path (VERDICT r1 missing 1 / 2): parallel tokenisation into memory-mapped shards with a
cache, block semantics equal to the reference's group_texts, streaming with rank sharding on the RAW stream, and
run_clm end to end on a large local text file with bounded memory.
reference: run_clm.py:316-381,463-544; sft_llama2.py:99-138."""
import json
import os
import resource
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

from distributed_lion_pytorch_b200.train.data import (ByteTokenizer, ConstantLengthDataset, StreamingBlockDataset,
                                                      build_loader, group_texts, text_file_dataset)
from distributed_lion_pytorch_b200.train.token_cache import MemmapBlockDataset, tokenize_text_file

ROOT = Path(__file__).resolve().parent.parent


def _write_corpus(path, n_lines, seed=0):
    rng = np.random.default_rng(seed)
    words = ["lion", "sign", "vote", "momentum", "bucket", "nvlink", "tile", "éléphant", "flag", "arena", ""]
    with open(path, "w", encoding="utf-8") as f:
        for i in range(n_lines):
            k = int(rng.integers(0, 12))
            f.write(" ".join(words[j] for j in rng.integers(0, len(words), k)) + "\n")


@pytest.mark.parametrize("workers", [1, 3])
def test_memmap_blocks_equal_group_texts(tmp_path, workers):
    """Shard boundaries, empty lines and multi-byte characters: the blocks must be exactly what the reference's
    tokenize + group_texts produce on the same file, whatever the number of tokeniser processes."""
    f = tmp_path / "corpus.txt"
    _write_corpus(f, 3000)
    tok = ByteTokenizer()
    lines = f.read_text(encoding="utf-8").splitlines(keepends=True)
    want = group_texts([tok(l)["input_ids"] for l in lines if l.strip()], 128)
    ds = text_file_dataset(str(f), tok, 128, num_workers=workers, cache_dir=str(tmp_path / "cache"))
    assert len(ds) == len(want) and len(ds) > 50
    for i in (0, 1, len(ds) // 2, len(ds) - 1):
        assert ds[i]["input_ids"].tolist() == want[i] and torch.equal(ds[i]["input_ids"], ds[i]["labels"])
    if workers > 1:
        assert len(ds.index["shards"]) > 1                       # blocks really straddle shard files


def test_cache_is_reused_and_overwrite_rebuilds(tmp_path):
    f = tmp_path / "c.txt"
    _write_corpus(f, 500)
    tok = ByteTokenizer()
    d1 = tokenize_text_file(str(f), tok, cache_dir=str(tmp_path / "cache"))
    stamp = (d1 / "index.json").stat().st_mtime_ns
    d2 = tokenize_text_file(str(f), tok, cache_dir=str(tmp_path / "cache"))
    assert d1 == d2 and (d2 / "index.json").stat().st_mtime_ns == stamp            # second call: cache hit
    d3 = tokenize_text_file(str(f), tok, cache_dir=str(tmp_path / "cache"), overwrite_cache=True)
    assert (d3 / "index.json").stat().st_mtime_ns != stamp
    _write_corpus(f, 600, seed=1)                                                   # the file changed: new key
    assert tokenize_text_file(str(f), tok, cache_dir=str(tmp_path / "cache")) != d1


class _CountingTok(ByteTokenizer):
    def __init__(self):
        self.calls = 0

    def encode_np(self, text):
        self.calls += 1
        return super().encode_np(text)

    def __call__(self, text, **kw):
        self.calls += 1
        return super().__call__(text, **kw)


def test_streaming_shards_the_raw_stream_before_tokenising():
    """W ranks over one stream: each raw example is tokenised by exactly one rank (1/W of the calls each), the ranks'
    blocks are disjoint, and together they cover the stream (every rank used to tokenise everything)."""
    stream = [{"text": f"example {i} " + "x" * (i % 37)} for i in range(400)]
    W = 4
    calls, seen = [], []
    for r in range(W):
        tok = _CountingTok()
        ds = StreamingBlockDataset(stream, tok, 16).shard(r, W)
        blocks = [b["input_ids"].tolist() for b in ds]
        calls.append(tok.calls)
        seen.append(blocks)
    assert calls == [100] * W
    texts = ["".join(chr(c) for blk in s for c in blk) for s in seen]
    assert "example 0 " in texts[0] and "example 1 " in texts[1] and "example 1 " not in texts[0]
    # through the loader (what the trainer calls), with sharding requested by build_loader
    tok = _CountingTok()
    loader, sampler = build_loader(StreamingBlockDataset(stream, tok, 16), 2, rank=1, world=W)
    n = sum(b["input_ids"].shape[0] for b in loader)
    assert sampler is None and n > 0 and tok.calls == 100


def test_sft_packing_shards_the_raw_stream():
    """reference: sft_llama2.py:99-138 streams question/answer pairs into ConstantLengthDataset; per-rank tokeniser
    calls must be about 1/W of the stream."""
    data = [{"question": f"q{i}", "response_j": "a" * (10 + i % 5)} for i in range(300)]
    W = 3
    calls = []
    for r in range(W):
        tok = _CountingTok()
        ds = ConstantLengthDataset(tok, data, infinite=False, seq_length=32, num_of_sequences=4, chars_per_token=1.0)
        loader, _ = build_loader(ds, 2, rank=r, world=W)
        assert sum(1 for _ in loader) > 0
        calls.append(tok.calls)
    assert calls == [100] * W


def test_streaming_hub_path_iterates_instead_of_indexing(tmp_path, monkeypatch):
    """--streaming with --dataset_name used to index an IterableDataset.  A fake `load_dataset` that
    returns plain generators (no indexing, no column_names) drives the same code."""
    sys.path.insert(0, str(ROOT))
    import run_clm

    class _Iter:
        def __init__(self, n):
            self.n = n

        def __iter__(self):
            return ({"text": f"doc {i} " + "y" * 50} for i in range(self.n))

    def fake_load(name, cfg=None, streaming=False, split=None):
        assert streaming
        return {"train": _Iter(200)}

    args = run_clm.DataTrainingArguments(dataset_name="fake/corpus", streaming=True, max_eval_samples=20)
    train, valid = run_clm._hub_dataset(args, ByteTokenizer(), 64, load_dataset=fake_load)
    vb = list(valid)
    assert 0 < len(vb) <= 20 and "doc 0 " in "".join(chr(c) for c in vb[0]["input_ids"].tolist())
    it = iter(train)
    first = next(it)
    assert "doc 20 " in "".join(chr(c) for c in first["input_ids"].tolist())       # the validation examples are skipped
    for _ in range(500):                                                           # infinite: restarts when exhausted
        next(it)


def test_run_clm_trains_from_a_large_text_file_with_bounded_memory(tmp_path):
    """A 200 MB corpus through the README-style command line on 2 gloo ranks: tokenised once by 4 processes into the
    cache, second launch hits the cache, peak RSS stays far below what holding the tokens as Python ints would need
    (200 M ints ≈ 1.6 GB of pointers alone), and the two ranks train on disjoint blocks."""
    corpus = tmp_path / "big.txt"
    line = ("the quick brown lion votes with one bit per parameter over nvlink " * 3 + "\n").encode()
    with open(corpus, "wb") as f:
        for _ in range(200 * (1 << 20) // len(line) + 1):
            f.write(line)
    assert corpus.stat().st_size >= 200 * (1 << 20)
    env = dict(os.environ, DLION_CACHE=str(tmp_path / "cache"), PYTHONPATH=str(ROOT), OMP_NUM_THREADS="1")
    from distutil import free_port
    base = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
            "--master-port", str(free_port()), "-m", "run_clm", "--model_type", "gpt2-tiny", "--train_file", str(corpus),
            "--validation_file", str(corpus), "--block_size", "64", "--per_device_train_batch_size", "2",
            "--gradient_accumulation_steps", "1", "--max_steps", "3", "--lion", "--async_grad", "--do_train", "--do_eval",
            "--preprocessing_num_workers", "4", "--max_eval_samples", "8", "--per_device_eval_batch_size", "2",
            "--logging_steps", "1", "--save_steps", "0", "--output_dir", str(tmp_path / "out"), "--overwrite_output_dir"]
    import time
    t0 = time.time()
    r = subprocess.run(base, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    first = time.time() - t0
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    peak_kb = resource.getrusage(resource.RUSAGE_CHILDREN).ru_maxrss
    assert peak_kb < 1_300_000, f"peak RSS of the largest child {peak_kb / 1e6:.2f} GB"
    caches = list((tmp_path / "cache").glob("*/index.json"))
    assert len(caches) == 1
    idx = json.loads(caches[0].read_text())
    assert idx["total_tokens"] == corpus.stat().st_size and len(idx["shards"]) >= 4
    stamp = caches[0].stat().st_mtime_ns
    # the two ranks' samplers over the cached blocks are disjoint (what the trainer's loaders iterate)
    ds = MemmapBlockDataset(caches[0].parent, 64)
    picks = []
    for rank in range(2):
        loader, sampler = build_loader(ds, 2, rank=rank, world=2, seed=42)
        picks.append(set(list(iter(sampler))[:1000]))
    assert picks[0] and picks[1] and not (picks[0] & picks[1])
    t0 = time.time()
    r = subprocess.run(base, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    # cache hit: the index and every shard are the files of the first launch, untouched (a wall-clock comparison of
    # the two launches would say the same but depends on how busy the machine is)
    assert caches[0].stat().st_mtime_ns == stamp
    assert all((caches[0].parent / sh["file"]).stat().st_mtime_ns <= stamp for sh in idx["shards"])
    second = time.time() - t0
    assert second < first + 30, (first, second)
