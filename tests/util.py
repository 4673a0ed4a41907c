import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def as_torch(arr, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(arr))
    if dtype == "bfloat16" or dtype is torch.bfloat16:
        return t.view(torch.bfloat16)
    return t


def golden_files(prefix):
    return sorted(f for f in os.listdir(GOLDEN) if f.startswith(prefix) and f.endswith(".npz"))
