"""Training loop, trainers and data utilities (see ``trainer.py`` for the design notes)."""
from .args import AsyncTrainingArguments, HfArgumentParser, TrainingArguments, parse_cli  # noqa: F401
from .data import (BlockDataset, ByteTokenizer, ConstantLengthDataset, SyntheticPreferenceDataset,  # noqa: F401
                   SyntheticTokenDataset, build_loader, default_data_collator, group_texts, load_tokenizer)
from .dpo import AsyncDPOTrainer, DPODataCollator, DPOTrainer, dpo_loss  # noqa: F401
from .schedules import (get_constant_schedule_with_warmup, get_cosine_schedule_with_warmup,  # noqa: F401
                        get_linear_schedule_with_warmup, get_scheduler)
from .sft import AsyncSFTTrainer, SFTTrainer  # noqa: F401
from .trainer import AsyncTrainer, Trainer, TrainerState, get_last_checkpoint, set_seed  # noqa: F401
