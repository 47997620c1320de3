"""CPU: the prefix-sum formulation of the quad-tree that k_octree implements (tools/octree_parallel_model.py) equals the
sequential oracle (std::list semantics of the reference) on random candidate sets."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_parallel_formulation_equals_sequential_oracle(oracle_mod):
    from octree_parallel_model import distribute_parallel

    rng = np.random.default_rng(42)
    n_phase2 = 0
    for t in range(60):
        W, H = int(rng.integers(60, 700)), int(rng.integers(40, 400))
        if round(W / H) < 1:
            continue
        M, N = int(rng.integers(0, 1500)), int(rng.integers(1, 400))
        pos = rng.permutation(W * H)[:M]
        xyr = np.stack([pos % W, pos // W, rng.integers(7, 40, len(pos))], 1).astype(np.int32)
        ref = oracle_mod.distribute(xyr, 16, 16 + W, 16, 16 + H, N)
        got = distribute_parallel(xyr, 16, 16 + W, 16, 16 + H, N)
        assert np.array_equal(ref, got), (t, W, H, M, N)
        n_phase2 += int(len(ref) >= N and M > N)
    assert n_phase2 > 10
