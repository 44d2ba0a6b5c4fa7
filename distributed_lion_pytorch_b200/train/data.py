"""Datasets, packing and loaders.

reference: run_clm.py:316-544 (load_dataset → tokenize → ``group_texts`` into fixed
``block_size`` blocks with ``labels = input_ids``), sft_llama2.py:99-138
(``ConstantLengthDataset`` packing of "Question/Answer" text), dpo_llama2.py:84-125
(prompt/chosen/rejected triples).  There is no network and no tokenizer cache on the box, so
every path has a synthetic twin that produces tensors of exactly the same shape.
"""
from __future__ import annotations

import itertools
import math
from typing import Callable, Dict, Iterable, Iterator, List, Optional, Sequence

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, IterableDataset, Sampler


# ----------------------------------------------------------------------------- synthetic
class SyntheticTokenDataset(Dataset):
    """Deterministic random token blocks: ``{"input_ids": int64[S], "labels": int64[S]}``.
    Sample ``i`` depends only on ``(seed, i)`` so every rank/worker can build its shard alone."""

    def __init__(self, vocab_size: int, seq_len: int, num_samples: int, seed: int = 0):
        self.vocab_size, self.seq_len, self.num_samples, self.seed = vocab_size, seq_len, num_samples, seed

    def __len__(self):
        return self.num_samples

    def __getitem__(self, i: int) -> Dict[str, torch.Tensor]:
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + int(i))
        ids = torch.randint(0, self.vocab_size, (self.seq_len,), generator=g, dtype=torch.int64)
        return {"input_ids": ids, "labels": ids.clone()}


class SyntheticPreferenceDataset(Dataset):
    """Token-level DPO triples with the shapes ``dpo_llama2`` produces after tokenisation:
    prompt ≤ max_prompt_length, prompt+answer ≤ max_length."""

    def __init__(self, vocab_size: int, max_prompt_length: int, max_length: int, num_samples: int, seed: int = 0):
        self.v, self.mp, self.ml, self.n, self.seed = vocab_size, max_prompt_length, max_length, num_samples, seed

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 7_000_003 + int(i))
        lp = int(torch.randint(self.mp // 2, self.mp + 1, (1,), generator=g))
        lc = int(torch.randint((self.ml - lp) // 2, self.ml - lp + 1, (1,), generator=g))
        lr = int(torch.randint((self.ml - lp) // 2, self.ml - lp + 1, (1,), generator=g))
        r = lambda n: torch.randint(3, self.v, (max(n, 1),), generator=g, dtype=torch.int64)
        return {"prompt_ids": r(lp), "chosen_ids": r(lc), "rejected_ids": r(lr)}


# ----------------------------------------------------------------------------- text → blocks
class ByteTokenizer:
    """Offline stand-in for AutoTokenizer: UTF-8 bytes + bos/eos/pad.  Used only when no HF
    tokenizer can be loaded (no cache, no network)."""
    vocab_size = 259
    bos_token_id, eos_token_id, pad_token_id = 256, 257, 258
    model_max_length = 1024
    padding_side = "right"

    def encode(self, text: str, add_special_tokens: bool = False) -> List[int]:
        ids = list(text.encode("utf-8"))
        return [self.bos_token_id] + ids if add_special_tokens else ids

    def __call__(self, text, **_):
        if isinstance(text, str):
            return {"input_ids": self.encode(text)}
        return {"input_ids": [self.encode(t) for t in text]}

    def encode_np(self, text: str) -> np.ndarray:
        """Tokens as a numpy array without a Python list in between (the bulk tokeniser's fast path)."""
        return np.frombuffer(text.encode("utf-8"), dtype=np.uint8)

    def decode(self, ids) -> str:
        return bytes(i for i in ids if i < 256).decode("utf-8", errors="replace")

    def __len__(self):
        return self.vocab_size


def load_tokenizer(name: Optional[str]):
    """HF tokenizer if it is cached locally, else the byte tokenizer."""
    if name:
        try:
            from transformers import AutoTokenizer
            return AutoTokenizer.from_pretrained(name, local_files_only=True)
        except Exception:
            pass
    return ByteTokenizer()


def group_texts(token_lists: Iterable[Sequence[int]], block_size: int) -> List[List[int]]:
    """Concatenate everything and cut into ``block_size`` blocks, dropping the remainder
    (reference: run_clm.py:509-522)."""
    buf: List[int] = []
    for t in token_lists:
        buf.extend(t)
    total = (len(buf) // block_size) * block_size
    return [buf[i:i + block_size] for i in range(0, total, block_size)]


class BlockDataset(Dataset):
    def __init__(self, blocks: List[List[int]]):
        self.blocks = blocks

    def __len__(self):
        return len(self.blocks)

    def __getitem__(self, i):
        ids = torch.tensor(self.blocks[i], dtype=torch.int64)
        return {"input_ids": ids, "labels": ids.clone()}


def text_file_dataset(path: str, tokenizer, block_size: int, keep_linebreaks: bool = True, *,
                      num_workers: Optional[int] = None, overwrite_cache: bool = False, cache_dir: Optional[str] = None,
                      max_blocks: Optional[int] = None):
    """A local text file as fixed-size token blocks (reference: run_clm.py:342-381 + :463-544).  The file is
    tokenised once by ``num_workers`` processes into memory-mapped shards under a cache directory
    (``train/token_cache.py``); later launches, and every rank after rank 0, just open the maps.  Neither the text
    nor the tokens are ever held in memory as a whole."""
    from .token_cache import MemmapBlockDataset, tokenize_text_file
    d = tokenize_text_file(path, tokenizer, num_workers=num_workers, cache_dir=cache_dir,
                           overwrite_cache=overwrite_cache, keep_linebreaks=keep_linebreaks)
    return MemmapBlockDataset(d, block_size, max_blocks=max_blocks)


class StreamingBlockDataset(IterableDataset):
    """A stream of raw text examples (``datasets`` streaming mode, or any iterable of dicts / strings) packed into
    fixed-size token blocks on the fly (reference: run_clm.py:316-341 with ``streaming=True``, then the same
    tokenize + ``group_texts`` maps applied lazily).

    Sharding happens on the RAW stream, before any tokenisation: with ``shard(rank, world)`` rank r keeps raw examples
    r, r+W, ... and inside a DataLoader with ``num_workers`` workers each worker keeps every n-th of those — so the
    tokeniser runs once per example across the whole job, not once per rank."""

    def __init__(self, source, tokenizer, block_size: int, text_key: Optional[str] = "text", skip: int = 0,
                 take: Optional[int] = None, max_blocks: Optional[int] = None, infinite: bool = False):
        self.source, self.tok, self.block_size, self.text_key = source, tokenizer, int(block_size), text_key
        self.skip, self.take, self.max_blocks = int(skip), take, max_blocks
        # training streams restart when exhausted: every rank must be able to reach --max_steps, and the shards of a
        # stream are not equally long
        self.infinite = bool(infinite)
        self._rank, self._world = 0, 1

    def shard(self, rank: int, world: int) -> "StreamingBlockDataset":
        self._rank, self._world = int(rank), int(world)
        return self

    def _raw(self) -> Iterator:
        it = iter(self.source)
        stop = None if self.take is None else self.skip + int(self.take)
        it = itertools.islice(it, self.skip, stop)
        info = torch.utils.data.get_worker_info()
        nw, wid = (info.num_workers, info.id) if info is not None else (1, 0)
        return itertools.islice(it, self._rank * nw + wid, None, self._world * nw)

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        from .token_cache import _encode
        buf = np.empty(0, dtype=np.int64)
        produced = 0
        while True:
            seen = 0
            for ex in self._raw():
                seen += 1
                text = ex if isinstance(ex, str) else ex[self.text_key]
                if not text:
                    continue
                buf = np.concatenate([buf, np.asarray(_encode(self.tok, text), dtype=np.int64)])
                while buf.size >= self.block_size:
                    ids = torch.from_numpy(buf[: self.block_size].copy())
                    buf = buf[self.block_size:]
                    yield {"input_ids": ids, "labels": ids.clone()}
                    produced += 1
                    if self.max_blocks is not None and produced >= self.max_blocks:
                        return
            if not self.infinite or seen == 0:
                return


def prepare_sample_text(example: dict) -> str:
    """reference: sft_llama2.py:93-96"""
    return f"Question: {example['question']}\n\nAnswer: {example['response_j']}"


def chars_token_ratio(dataset, tokenizer, nb_examples: int = 400) -> float:
    """reference: sft_llama2.py:62-75"""
    total_chars = total_tokens = 0
    for _, ex in zip(range(nb_examples), iter(dataset)):
        text = prepare_sample_text(ex)
        total_chars += len(text)
        total_tokens += len(tokenizer(text)["input_ids"])
    return total_chars / max(total_tokens, 1)


class ConstantLengthDataset(IterableDataset):
    """Pack formatted text samples into constant-length token blocks separated by EOS
    (stands in for TRL's ConstantLengthDataset, reference: sft_llama2.py:122-137)."""

    def __init__(self, tokenizer, dataset, formatting_func: Optional[Callable] = None, infinite: bool = False,
                 seq_length: int = 1024, num_of_sequences: int = 1024, chars_per_token: float = 3.6,
                 eos_token_id: Optional[int] = None):
        self.tokenizer, self.dataset, self.fmt = tokenizer, dataset, formatting_func or prepare_sample_text
        self.infinite, self.seq_length = infinite, seq_length
        self.max_buffer_size = int(seq_length * chars_per_token * num_of_sequences)
        self.eos = eos_token_id if eos_token_id is not None else getattr(tokenizer, "eos_token_id", 0)
        self._rank, self._world = 0, 1

    def shard(self, rank: int, world: int) -> "ConstantLengthDataset":
        """Rank ``rank`` of ``world`` formats, tokenises and packs only every world-th RAW example (accelerate's
        dataloader sharding, which the reference relies on at sft_llama2.py:171-181, dispatches whole batches
        instead; either way no example is tokenised more than once per job)."""
        self._rank, self._world = int(rank), int(world)
        return self

    def _raw(self) -> Iterator:
        info = torch.utils.data.get_worker_info()
        nw, wid = (info.num_workers, info.id) if info is not None else (1, 0)
        return itertools.islice(iter(self.dataset), self._rank * nw + wid, None, self._world * nw)

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        it = self._raw()
        more = True
        while more:
            buffer, buflen = [], 0
            while buflen < self.max_buffer_size:
                try:
                    text = self.fmt(next(it))
                    buffer.append(text)
                    buflen += len(text)
                except StopIteration:
                    if self.infinite:
                        it = self._raw()
                    else:
                        more = False
                        break
            all_ids: List[int] = []
            for text in buffer:
                all_ids.extend(self.tokenizer(text)["input_ids"] + [self.eos])
            for i in range(0, len(all_ids) - self.seq_length + 1, self.seq_length):
                ids = torch.tensor(all_ids[i:i + self.seq_length], dtype=torch.int64)
                yield {"input_ids": ids, "labels": ids.clone()}


# ----------------------------------------------------------------------------- loading
def default_data_collator(features: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
    out = {}
    for k in features[0]:
        v0 = features[0][k]
        if torch.is_tensor(v0):
            out[k] = torch.stack([f[k] for f in features])
        elif isinstance(v0, (list, tuple)) and v0 and isinstance(v0[0], int):
            out[k] = torch.tensor([f[k] for f in features], dtype=torch.int64)
        elif isinstance(v0, (int, float)):
            out[k] = torch.tensor([f[k] for f in features])
    return out


class ShardSampler(Sampler):
    """Epoch-seeded permutation, rank r takes indices r, r+W, … (what accelerate's dataloader
    sharding gives the reference); ``skip`` drops the batches a resumed run already consumed."""

    def __init__(self, n: int, rank: int, world: int, seed: int, shuffle: bool = True):
        self.n, self.rank, self.world, self.seed, self.shuffle = n, rank, world, seed, shuffle
        self.epoch = 0
        self.per_rank = n // world if n >= world else n

    def set_epoch(self, e: int):
        self.epoch = e

    def __len__(self):
        return self.per_rank

    def __iter__(self):
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            order = torch.randperm(self.n, generator=g).tolist()
        else:
            order = list(range(self.n))
        if self.n >= self.world:
            order = order[: self.per_rank * self.world][self.rank::self.world]
        return iter(order)


def build_loader(dataset, batch_size: int, *, rank: int = 0, world: int = 1, seed: int = 0, shuffle: bool = True,
                 collate_fn=None, num_workers: int = 0, pin_memory: bool = False, drop_last: bool = True):
    collate_fn = collate_fn or default_data_collator
    if isinstance(dataset, IterableDataset) and callable(getattr(dataset, "shard", None)):
        # shards its own RAW stream by rank (and by dataloader worker): tokenisation is not repeated on every rank
        dataset.shard(rank, world)
        return DataLoader(dataset, batch_size=batch_size, collate_fn=collate_fn, num_workers=num_workers,
                          pin_memory=pin_memory, drop_last=drop_last, persistent_workers=num_workers > 0), None
    if isinstance(dataset, IterableDataset):
        class _Sharded(IterableDataset):
            def __iter__(self_inner):
                for i, x in enumerate(dataset):
                    if i % world == rank:
                        yield x
        return DataLoader(_Sharded(), batch_size=batch_size, collate_fn=collate_fn, num_workers=0,
                          pin_memory=pin_memory, drop_last=drop_last), None
    sampler = ShardSampler(len(dataset), rank, world, seed, shuffle)
    return DataLoader(dataset, batch_size=batch_size, sampler=sampler, collate_fn=collate_fn,
                      num_workers=num_workers, pin_memory=pin_memory, drop_last=drop_last,
                      persistent_workers=num_workers > 0), sampler
