#!/usr/bin/env python
"""Evaluation forward (model + loss under no_grad) of GPT-2 small, 20 x 1024 tokens per batch, with the tcgen05 LM-head
kernel and with the cuBLAS block + CE-kernel path (DLION_LMHEAD_FUSED=0): tokens/s of `Trainer.evaluate`'s inner step.
    python bench/eval_loss_microbench.py [--iters 10]"""
import argparse
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion_pytorch_b200.models import build_model  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    torch.manual_seed(0)
    model = build_model("gpt2", dtype=torch.bfloat16, device="cuda").eval()
    ids = torch.randint(0, 50257, (20, 1024), device="cuda")
    out = {}
    for mode in ("1", "0"):
        os.environ["DLION_LMHEAD_FUSED"] = mode
        with torch.no_grad():
            for _ in range(3):
                loss = model(ids, labels=ids).loss
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.iters):
                loss = model(ids, labels=ids).loss
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        out["fused_lm_head" if mode == "1" else "cublas_block_plus_ce_kernel"] = {
            "ms_per_batch": round(ms, 3), "tokens_per_s": round(ids.numel() / ms * 1e3), "loss": float(loss)}
    os.environ.pop("DLION_LMHEAD_FUSED")
    out["speedup"] = round(out["cublas_block_plus_ce_kernel"]["ms_per_batch"] / out["fused_lm_head"]["ms_per_batch"], 3)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
