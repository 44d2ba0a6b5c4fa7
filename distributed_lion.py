"""Public entry point kept from the reference: ``from distributed_lion import Lion``
(reference: README.md:8-15, distributed_lion.py:138).  The implementation lives in
``distributed_lion_pytorch_b200.optim.lion``."""
from distributed_lion_pytorch_b200.optim.lion import Lion  # noqa: F401
from distributed_lion_pytorch_b200.optim.semantics import (  # noqa: F401
    pack_bits, unpack_bits, count_votes, vote_direction)

__all__ = ["Lion", "pack_bits", "unpack_bits", "count_votes", "vote_direction"]
