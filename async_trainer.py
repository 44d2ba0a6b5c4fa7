"""Public entry point kept from the reference: ``from async_trainer import AsyncTrainer,
AsyncSFTTrainer, AsyncDPOTrainer`` (reference: async_trainer.py:8,37,65; README.md:10-14).
The implementations live in ``distributed_lion_pytorch_b200.train``."""
from distributed_lion_pytorch_b200.train import (AsyncDPOTrainer, AsyncSFTTrainer, AsyncTrainer,  # noqa: F401
                                                 DPOTrainer, SFTTrainer, Trainer, TrainingArguments)

__all__ = ["AsyncTrainer", "AsyncSFTTrainer", "AsyncDPOTrainer", "Trainer", "SFTTrainer", "DPOTrainer",
           "TrainingArguments"]


def hf_async_trainer_class():
    """``transformers.Trainer`` subclass with the reference's no_sync ``training_step``
    (reference: async_trainer.py:8-34), for installations where HF's Trainer is usable
    (it needs ``accelerate``, which is absent on the B200 image — there the in-house
    ``AsyncTrainer`` above is the one to use).  Written against transformers >= 4.46, whose
    ``training_step`` takes ``num_items_in_batch``."""
    import contextlib

    from transformers import Trainer as HFTrainer

    class HFAsyncTrainer(HFTrainer):
        def training_step(self, model, inputs, num_items_in_batch=None):
            no_sync = getattr(model, "no_sync", None)
            ctx = no_sync() if callable(no_sync) else contextlib.nullcontext()      # works un-wrapped too (D8)
            with ctx:
                return super().training_step(model, inputs, num_items_in_batch)

    return HFAsyncTrainer
