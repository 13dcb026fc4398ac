"""B200-native layer-sharded prompt scoring (drop-in for flexible-LLM-sharding's hot path).

Sub-modules (imported lazily so that `synthetic`/`planner` work without the CUDA library):
  _lib        ctypes binding of the C-ABI library `csrc/libfls_b200.so` (include/fls_b200.h)
  planner     shard plan / batching / prompt partitioning (reference utils.py:143-157, main.py:14-25)
  utils       DeviceManager, ShardedLlama (reference utils.py:24-305)
  activation_store  where token rows live between shards: HBM / pinned host / `.npy` files (utils.py:159-213)
  pipeline    model-parallel layer pipeline with GPU-to-GPU hand-off (utils.py:150-157,185-195)
  main        CLI + run_sharded_llama (reference main.py)
  prepare_weights  per-layer safetensors splitter (reference prepare_weights.py)
  synthetic   random-init checkpoints / tokenizer / prompts for tests and bench
"""
__version__ = "0.1.0"
