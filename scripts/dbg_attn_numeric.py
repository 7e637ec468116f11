"""Debug: numeric check of lg_attention with many work items per CTA."""
import os, sys
import numpy as np
import torch
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
import test_kernels_gpu as T
from conftest import pkg
C = pkg("_cabi")
B, N = 16, 2048
rng = np.random.default_rng(11)
rnd = rng.integers(1024, N + 1, B).tolist()
for name, ql, kl in (("none", None, None), ("q1792", [1792] * B, None), ("k1792", None, [1792] * B),
                     ("q1793", [1793] * B, None), ("qk random", rnd, rnd), ("q2048", [2048] * B, None)):
    for rep in range(2):
        errs = T._attn_case(C, B, 4, N, N, ql, kl, seed=21, check=list(range(0, B, 3)))
        print(name, rep, " ".join(f"{eb:.3f}" for _, eb in errs), flush=True)
