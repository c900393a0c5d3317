"""Drop-in for the reference's `models` package: exports BaseModel and AdaINModel
(reference src/models/__init__.py:1-2), resolved by name from `--model` (arguments.py:57)."""
from .adain_model import AdaINModel
from .base_model import BaseModel
