"""distributed_lion_pytorch_b200 — a Blackwell (sm_100a) native Distributed Lion framework.

Public surface (mirrors kyleliang919/distributed-lion-pytorch):

* ``Lion``                      — optimizer (reference: distributed_lion.py:138-200)
* ``AsyncTrainer`` & friends    — trainers that never all-reduce gradients
                                  (reference: async_trainer.py:8-90)
* ``models.GPT2LMHeadModel`` / ``models.LlamaForCausalLM`` — HF-named model families
* ``ops``                       — hand-written sm_100a kernels behind a C ABI (include/dlion.h)

The compute path is PyTorch + the in-tree CUDA library ``_lib/libdlion.so``; the
cross-GPU sign exchange of the optimizer runs inside one persistent kernel over
NVLink peer memory (``parallel.symm``), not through NCCL.
"""
from .version import __version__  # noqa: F401


def __getattr__(name):  # lazy: keep `import distributed_lion_pytorch_b200` cheap
    if name == "Lion":
        from .optim.lion import Lion
        return Lion
    if name in ("AsyncTrainer", "AsyncSFTTrainer", "AsyncDPOTrainer", "Trainer", "TrainingArguments"):
        from . import train
        return getattr(train, name)
    raise AttributeError(name)
