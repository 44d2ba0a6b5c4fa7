"""Public entry point kept from the reference: ``from async_trainer import AsyncTrainer,
AsyncSFTTrainer, AsyncDPOTrainer`` (reference: async_trainer.py:8,37,65; README.md:10-14).
The implementations live in ``distributed_lion_pytorch_b200.train``."""
from distributed_lion_pytorch_b200.train import (AsyncDPOTrainer, AsyncSFTTrainer, AsyncTrainer,  # noqa: F401
                                                 DPOTrainer, SFTTrainer, Trainer, TrainingArguments)

__all__ = ["AsyncTrainer", "AsyncSFTTrainer", "AsyncDPOTrainer", "Trainer", "SFTTrainer", "DPOTrainer",
           "TrainingArguments"]
