"""B200-native BERT question-answering fine-tuning step (drop-in for the hot path of
neuro-inc/ml-recipe-distributed-pytorch).  See DESIGN.md.

The directory name carries a hyphen (it is fixed by the project layout), so import it through
``__graft_entry__.load_package()`` which registers it as ``bqa_b200``.
"""
from . import _lib  # noqa: F401  (fails loudly if libbqa.so has not been built)
from . import engine  # noqa: F401,E402
