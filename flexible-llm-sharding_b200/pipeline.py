"""Model-parallel layer pipeline across the GPUs of one process (reference utils.py:150-157,185-195,
main.py:71-73).  Implemented in `run_model_parallel`."""
from __future__ import annotations


def run_model_parallel(model, inputs, shards, num_gpu, rank, want_probs, answer_ids):
    raise NotImplementedError("model-parallel pipeline: not built yet")
