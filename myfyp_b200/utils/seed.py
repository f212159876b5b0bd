"""Reproducibility helper (the fork author's experiment script relies on one, exp_SAVE3.txt:1-40)."""

from __future__ import annotations

import random
from typing import Optional

import numpy as np


def set_seed(seed: Optional[int] = 666) -> None:
    if seed is None:
        return
    import torch

    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    from myfyp_b200.settings import Settings

    Settings.SEED = seed
