"""Tokenise once, in parallel, into memory-mapped token shards; read fixed-size blocks straight from the maps.

reference: run_clm.py:463-544 — ``raw_datasets.map(tokenize_function, batched=True, num_proc=preprocessing_num_workers,
load_from_cache_file=not overwrite_cache)`` followed by ``map(group_texts, ...)``: Arrow-backed, memory-mapped,
multi-process, cached.  ``datasets`` does that for hub corpora; for local text files (and for anything iterable) this
module provides the same properties without it:

* the corpus is cut into byte ranges on line boundaries; ``num_workers`` processes each stream their ranges line by
  line through the tokenizer and append ``uint16`` / ``uint32`` tokens to one ``.bin`` shard per range — no process
  ever holds more than one write buffer of tokens, whatever the corpus size (openwebtext is ~9 G tokens);
* shards + an ``index.json`` live under a cache directory keyed by (file identity, tokenizer identity, options), so a
  second launch — and every rank after rank 0 — opens the maps instead of tokenising again (``overwrite_cache``
  rebuilds);
* ``MemmapBlockDataset`` cuts the concatenation of all shards into ``block_size`` blocks (``group_texts`` semantics:
  one stream, remainder dropped, ``labels = input_ids``) by offset arithmetic over ``np.memmap`` views.
"""
from __future__ import annotations

import hashlib
import json
import multiprocessing as mp
import os
import tempfile
from bisect import bisect_right
from pathlib import Path
from typing import Dict, Iterable, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data import Dataset

_FLUSH_TOKENS = 1 << 22           # a worker's write buffer: 4 Mi tokens (8 - 16 MB)


def default_cache_dir() -> Path:
    return Path(os.environ.get("DLION_CACHE", os.path.join(tempfile.gettempdir(), "dlion_token_cache")))


def tokenizer_identity(tokenizer) -> str:
    name = getattr(tokenizer, "name_or_path", None) or type(tokenizer).__name__
    return f"{type(tokenizer).__module__}.{type(tokenizer).__name__}:{name}:{len(tokenizer)}"


def _token_dtype(vocab_size: int):
    return np.uint16 if vocab_size <= 65536 else np.uint32


def _encode(tokenizer, text: str) -> np.ndarray:
    fast = getattr(tokenizer, "encode_np", None)
    if fast is not None:
        return fast(text)
    return np.asarray(tokenizer(text)["input_ids"], dtype=np.int64)


def _line_ranges(path: str, n: int) -> List[Tuple[int, int]]:
    """``n`` byte ranges covering the file, each starting at the beginning of a line."""
    size = os.path.getsize(path)
    if size == 0:
        return []
    cuts = [0]
    with open(path, "rb") as f:
        for k in range(1, n):
            pos = size * k // n
            if pos <= cuts[-1]:
                continue
            f.seek(pos)
            f.readline()                      # move to the next line start
            pos = f.tell()
            if cuts[-1] < pos < size:
                cuts.append(pos)
    cuts.append(size)
    return [(a, b) for a, b in zip(cuts[:-1], cuts[1:]) if b > a]


def _tokenize_range(job) -> Tuple[str, int]:
    """Worker: stream one byte range of a text file through the tokenizer into one shard file."""
    path, start, end, out_path, tokenizer, dtype_name, keep_linebreaks = job
    dtype = np.dtype(dtype_name)
    total = 0
    buf: List[np.ndarray] = []
    buffered = 0
    with open(path, "rb") as f, open(out_path, "wb") as out:
        f.seek(start)
        while f.tell() < end:
            raw = f.readline()
            if not raw:
                break
            line = raw.decode("utf-8", errors="replace")
            if not keep_linebreaks:
                line = line.rstrip("\r\n")
            if not line.strip():
                continue                     # empty lines carry no tokens (as the reference's text loader drops them)
            ids = _encode(tokenizer, line)
            buf.append(ids.astype(dtype, copy=False))
            buffered += ids.size
            if buffered >= _FLUSH_TOKENS:
                np.concatenate(buf).tofile(out)
                total += buffered
                buf, buffered = [], 0
        if buf:
            np.concatenate(buf).tofile(out)
            total += buffered
    return out_path, total


def _tokenize_texts(job) -> Tuple[str, int]:
    """Worker: tokenise an in-memory list of texts (hub datasets handed over in slices) into one shard file."""
    texts, out_path, tokenizer, dtype_name = job
    dtype = np.dtype(dtype_name)
    total = 0
    with open(out_path, "wb") as out:
        buf, buffered = [], 0
        for t in texts:
            if not t:
                continue
            ids = _encode(tokenizer, t)
            buf.append(ids.astype(dtype, copy=False))
            buffered += ids.size
            if buffered >= _FLUSH_TOKENS:
                np.concatenate(buf).tofile(out)
                total += buffered
                buf, buffered = [], 0
        if buf:
            np.concatenate(buf).tofile(out)
            total += buffered
    return out_path, total


def _run_jobs(fn, jobs: list, num_workers: int) -> List[Tuple[str, int]]:
    if num_workers <= 1 or len(jobs) <= 1:
        return [fn(j) for j in jobs]
    ctx = mp.get_context("fork" if hasattr(os, "fork") else "spawn")
    with ctx.Pool(min(num_workers, len(jobs))) as pool:
        return pool.map(fn, jobs, chunksize=1)


def _write_index(d: Path, shards: List[Tuple[str, int]], dtype, meta: dict) -> dict:
    index = {"dtype": np.dtype(dtype).name, "shards": [{"file": os.path.basename(p), "tokens": int(n)} for p, n in shards if n > 0],
             "total_tokens": int(sum(n for _, n in shards)), **meta}
    tmp = d / "index.json.tmp"
    tmp.write_text(json.dumps(index, indent=1))
    os.replace(tmp, d / "index.json")            # atomic: a reader never sees a half-written index
    return index


def tokenize_text_file(path: str, tokenizer, *, num_workers: Optional[int] = None, cache_dir: Optional[str] = None,
                       overwrite_cache: bool = False, keep_linebreaks: bool = True) -> Path:
    """Tokenise ``path`` into cached shards (or find them); returns the cache directory holding ``index.json``."""
    path = os.path.abspath(path)
    st = os.stat(path)
    vocab = len(tokenizer)
    key = hashlib.sha1(json.dumps([path, st.st_size, int(st.st_mtime), tokenizer_identity(tokenizer), keep_linebreaks,
                                   "v1"]).encode()).hexdigest()[:20]
    d = Path(cache_dir) if cache_dir else default_cache_dir()
    d = d / f"{Path(path).stem}-{key}"
    if (d / "index.json").exists() and not overwrite_cache:
        return d
    d.mkdir(parents=True, exist_ok=True)
    workers = max(int(num_workers or 1), 1)
    ranges = _line_ranges(path, workers * 4 if workers > 1 else 1)
    dtype = _token_dtype(vocab)
    jobs = [(path, a, b, str(d / f"shard-{i:05d}.bin"), tokenizer, np.dtype(dtype).name, keep_linebreaks)
            for i, (a, b) in enumerate(ranges)]
    shards = _run_jobs(_tokenize_range, jobs, workers)
    _write_index(d, shards, dtype, {"source": path, "tokenizer": tokenizer_identity(tokenizer), "vocab_size": vocab})
    return d


def tokenize_text_iterable(texts: Iterable[str], tokenizer, name: str, *, num_workers: Optional[int] = None,
                           cache_dir: Optional[str] = None, overwrite_cache: bool = False, fingerprint: str = "",
                           slice_size: int = 20000) -> Path:
    """Same for any finite iterable of strings (a hub dataset column): slices of ``slice_size`` texts go to the
    worker pool, one shard per slice, in order."""
    key = hashlib.sha1(json.dumps([name, fingerprint, tokenizer_identity(tokenizer), "v1"]).encode()).hexdigest()[:20]
    d = (Path(cache_dir) if cache_dir else default_cache_dir()) / f"{name.replace('/', '_')}-{key}"
    if (d / "index.json").exists() and not overwrite_cache:
        return d
    d.mkdir(parents=True, exist_ok=True)
    workers = max(int(num_workers or 1), 1)
    dtype = _token_dtype(len(tokenizer))
    shards: List[Tuple[str, int]] = []
    pending: list = []
    cur: List[str] = []

    def flush_pending():
        nonlocal pending
        if pending:
            shards.extend(_run_jobs(_tokenize_texts, pending, workers))
            pending = []

    i = 0
    for t in texts:
        cur.append(t)
        if len(cur) >= slice_size:
            pending.append((cur, str(d / f"shard-{i:05d}.bin"), tokenizer, np.dtype(dtype).name))
            cur, i = [], i + 1
            if len(pending) >= workers * 2:      # bounded: at most 2 slices of text per worker in memory
                flush_pending()
    if cur:
        pending.append((cur, str(d / f"shard-{i:05d}.bin"), tokenizer, np.dtype(dtype).name))
    flush_pending()
    _write_index(d, shards, dtype, {"source": name, "tokenizer": tokenizer_identity(tokenizer), "vocab_size": len(tokenizer)})
    return d


class MemmapBlockDataset(Dataset):
    """Fixed-size blocks over the concatenation of the token shards of one cache directory."""

    def __init__(self, cache_dir, block_size: int, max_blocks: Optional[int] = None):
        self.dir = Path(cache_dir)
        self.index = json.loads((self.dir / "index.json").read_text())
        self.block_size = int(block_size)
        self.dtype = np.dtype(self.index["dtype"])
        self.files = [self.dir / s["file"] for s in self.index["shards"]]
        counts = [int(s["tokens"]) for s in self.index["shards"]]
        self.starts = [0]
        for c in counts:
            self.starts.append(self.starts[-1] + c)
        self.total_tokens = self.starts[-1]
        self.num_blocks = self.total_tokens // self.block_size
        if max_blocks is not None:
            self.num_blocks = min(self.num_blocks, int(max_blocks))
        self._maps: Dict[int, np.memmap] = {}

    def __len__(self):
        return self.num_blocks

    def __getstate__(self):                       # memmaps are re-opened lazily in dataloader workers
        st = self.__dict__.copy()
        st["_maps"] = {}
        return st

    def _map(self, k: int) -> np.memmap:
        m = self._maps.get(k)
        if m is None:
            m = self._maps[k] = np.memmap(self.files[k], dtype=self.dtype, mode="r")
        return m

    def tokens(self, start: int, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.int64)
        got = 0
        k = bisect_right(self.starts, start) - 1
        while got < n:
            off = start + got - self.starts[k]
            take = min(n - got, self.starts[k + 1] - self.starts[k] - off)
            out[got:got + take] = self._map(k)[off:off + take]
            got += take
            k += 1
        return out

    def __getitem__(self, i: int) -> Dict[str, torch.Tensor]:
        if not 0 <= i < self.num_blocks:
            raise IndexError(i)
        ids = torch.from_numpy(self.tokens(i * self.block_size, self.block_size))
        return {"input_ids": ids, "labels": ids.clone()}
