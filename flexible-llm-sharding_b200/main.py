"""CLI - same flags, pickle formats and side effects as the reference's `main.py`
(reference main.py:28-98), driving the B200 path.

    python main.py --model_path ./llama2_layers --prompt_pickle prompts.pkl --output_file scores.pkl \
                   [--layer_num_per_shard 1] [--storage_location cpu|gpu|disk] [--num_batch 1] \
                   [--data_parallel True] [--max_activation_in_cpu 100] [--disk_folder ./temp] [--num_gen_token 1]

Input pickle: list of `(prefix: str, suffixes: tuple[str, ...])`.  Outputs: `output_file` = list of
`np.ndarray [n_suffixes, num_gen_token, vocab]` float16, and `<prompt_pickle>_updated.pkl` with the
generated tokens appended to every suffix (main.py:86-98).

Multi-GPU: inside one process the reference's threading model is kept (one host thread per visible
GPU, main.py:67-73).  Under `torchrun` (WORLD_SIZE set) each rank drives its own GPU instead:
`--data_parallel` partitions the prompts (np.array_split order); without it the ranks form the reference's
default layer pipeline (shard j on rank j mod G) with the activations handed rank to rank over NVLink
(pipeline_dist.py).  Rank 0 gathers the scores either way.
"""
from __future__ import annotations

import argparse
import copy
import os
import pickle
import sys
from concurrent.futures import ThreadPoolExecutor
from functools import partial
from time import perf_counter

import numpy as np

from . import planner


def run_sharded_llama(device, prompt_list, device_manager, args):
    """reference main.py:14-25."""
    from .utils import ShardedLlama
    # The reference builds a new ShardedLlama per call (main.py:15); here the per-device model (CUDA context,
    # pinned/HBM-resident weights) is kept on the DeviceManager and reused for later batches and tokens.
    models = getattr(device_manager, "models", None)
    model = models.get(device) if models is not None else None
    if model is None:
        model = ShardedLlama(args, device_manager, device=device)
        if models is not None:
            models[device] = model
    if not args.data_parallel and device != device_manager.devices[0]:  # don't need actual prompts
        prompt_list = [None] * len(prompt_list)
    outputs = []
    t0, s0 = perf_counter(), model.ctx.stats()
    for lo, hi in planner.batch_ranges(len(prompt_list), args.num_batch):
        outputs += model(prompt_list[lo:hi])
    # the reference's end-of-call line (utils.py:304), with the rates this build is judged by; stderr, so that
    # stdout stays free for callers that print results there
    dt, s1 = perf_counter() - t0, model.ctx.stats()
    gb, ms = (s1.h2d_bytes - s0.h2d_bytes) / 1e9, s1.h2d_ms - s0.h2d_ms
    what = (f"scored {len(outputs)} prompts in {dt:.2f}s ({len(outputs) / max(dt, 1e-9):.1f} prompts/s)" if outputs
            else f"ran its pipeline stages in {dt:.2f}s")
    sys.stderr.write(f"{device} {what}; streamed {gb:.1f} GB of weights"
                     + (f" at {gb / (ms * 1e-3):.1f} GB/s" if ms > 0 else "")
                     + f"; activations: {model.last_timing.get('activation_tier', '-')}, prefix K/V: "
                     f"{model.last_timing.get('prefix_kv', '-')}\n")        # one write: threads do not interleave
    return outputs


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    parser.add_argument("--model_path", type=str, default="./")
    parser.add_argument("--prompt_pickle", type=str, required=True, help="Path to the input prompt pickle file")
    parser.add_argument("--output_file", type=str, required=True, help="Path to the LLM output scores file")
    parser.add_argument("--num_batch", type=int, default=1)
    parser.add_argument("--layer_num_per_shard", type=int, default=1)
    parser.add_argument("--storage_location", type=str, default="cpu",
                        help="'gpu': use vRAM to store intermediate activations, 'cpu': use RAM, 'disk': use disk")
    parser.add_argument("--max_activation_in_cpu", type=int, default=100)
    # type=bool as in the reference (main.py:40): any non-empty string, including "False", is True
    parser.add_argument("--data_parallel", type=bool, default=False, required=False,
                        help="if True, multiple GPUs run with data parallelism; if False, with model parallelism.")
    parser.add_argument("--disk_folder", type=str, default="./temp",
                        help="folder path for writing files of intermediate activations in 'disk' mode")
    parser.add_argument("--num_gen_token", type=int, default=1, help="how many new tokens to be generated")
    # not a reference flag: rows per micro-batch (a prompt's prefix + all its option rows must fit one);
    # also settable through the FLS_MAX_ROWS environment variable
    parser.add_argument("--max_rows", type=int, default=None,
                        help="token rows per micro-batch (default 16384); raise it for prompts with more rows")
    # not a reference flag either (FLS_PACK_ORDER): how prompts are assigned to micro-batches; the scores do not change
    parser.add_argument("--pack_order", choices=["length", "input"], default=None,
                        help="length (default): balanced, longest-first micro-batches; input: in the order given")
    return parser


def score_all(args, input_prompts, devices, device_manager=None):
    """One pass over all prompts on all devices of this process (main.py:66-76)."""
    from .utils import DeviceManager
    if device_manager is None:
        device_manager = DeviceManager(args, devices)

    def run(device, prompt_list):
        try:
            return run_sharded_llama(device, prompt_list, device_manager=device_manager, args=args)
        except BaseException as e:      # a dead worker must not leave its pipeline peers waiting (reference: they spin forever)
            with device_manager.condition:
                device_manager.failed = e
                device_manager.condition.notify_all()
            raise

    if len(devices) > 1:
        if args.data_parallel:
            parts = [input_prompts[lo:hi] for lo, hi in planner.partition_prompts(len(input_prompts), len(devices))]
            with ThreadPoolExecutor(max_workers=len(devices)) as ex:
                outputs = list(ex.map(run, devices, parts))
        else:
            with ThreadPoolExecutor(max_workers=len(devices)) as ex:
                outputs = list(ex.map(run, devices, [list(input_prompts) for _ in devices]))
        return sum(outputs, [])
    return run(devices[0], input_prompts)


def setup_staging(args):
    """Ranks that score the same model from a DISK directory would each hold their own pinned copy of every layer
    (8 x 26 GB for 13B on 8 GPUs).  Only files on tmpfs can be page-locked in place and shared, so with --data_parallel
    under torchrun the layer files are staged once per box in a tmpfs directory (`fls_set_stage_dir`) when it has
    room: FLS_STAGE_DIR names it (then it is the user's to remove; "" switches staging off), otherwise
    /dev/shm/fls_stage_<uid>_<port> is used and removed by local rank 0 at the end of main().  Returns the
    directory this call created, else None."""
    import shutil
    from . import _lib
    from .utils import _on_tmpfs
    d, created = os.environ.get("FLS_STAGE_DIR"), None
    if d is None and args.data_parallel and int(os.environ.get("WORLD_SIZE", "1")) > 1 \
            and not _on_tmpfs(args.model_path) and os.path.isdir("/dev/shm"):
        try:
            need = sum(os.path.getsize(os.path.join(args.model_path, f)) for f in os.listdir(args.model_path)
                       if f.endswith(".safetensors"))
            free = shutil.disk_usage("/dev/shm").free
        except OSError:
            need, free = 1, 0
        if need * 1.05 < free:
            d = created = f"/dev/shm/fls_stage_{os.getuid()}_{os.environ.get('MASTER_PORT', '0')}"
    if d:
        os.makedirs(d, exist_ok=True)
        _lib.load().fls_set_stage_dir(d.encode())
    return created


def distributed_device_manager(args):
    """DeviceManager of one torchrun rank (one process per GPU of one node).  --data_parallel: the process sees
    only its own GPU, as a one-device run.  Model parallel: `devices` lists every rank's GPU in rank order, so
    the reference's shard ownership (utils.py:150-153: rank = index of the own device) holds across processes."""
    import torch.distributed as dist
    from .utils import DeviceManager
    local = int(os.environ.get("LOCAL_RANK", dist.get_rank()))
    if args.data_parallel:
        return DeviceManager(args, [f"cuda:{local}"])
    if local != dist.get_rank():
        raise SystemExit("the model-parallel pipeline spans the GPUs of ONE node (LOCAL_RANK must equal RANK)")
    dm = DeviceManager(args, [f"cuda:{i}" for i in range(dist.get_world_size())])
    dm.distributed = True
    return dm


def score_all_distributed(args, input_prompts, device_manager=None):
    """torchrun mode.  --data_parallel: rank r scores its np.array_split part on its own GPU.  Otherwise the ranks
    form the layer pipeline (pipeline_dist.py): every rank gets the whole list (main.py:72-73) and the owner of
    `lm_head` returns the scores.  Rank 0 gathers either way."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    local = int(os.environ.get("LOCAL_RANK", rank))
    device = f"cuda:{local}"
    dm = device_manager if device_manager is not None else distributed_device_manager(args)
    if args.data_parallel:
        lo, hi = planner.partition_prompts(len(input_prompts), world)[rank]
    else:
        lo, hi = 0, len(input_prompts)
    mine = run_sharded_llama(device, input_prompts[lo:hi], dm, args)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(mine, gathered, dst=0)
    if rank == 0:
        return sum(gathered, [])
    return None


def main(argv=None):
    import torch
    args = build_parser().parse_args(argv)
    print(args)
    if args.storage_location == "disk":
        os.makedirs(args.disk_folder, exist_ok=True)
    with open(args.prompt_pickle, "rb") as f:
        original_input_prompts = pickle.load(f)
    input_prompts = copy.deepcopy(original_input_prompts)

    distributed = int(os.environ.get("WORLD_SIZE", "1")) > 1
    if distributed:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("gloo")       # host-side object exchange only; activations travel over NVLink
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", dist.get_rank())))
        is_root = dist.get_rank() == 0
    else:
        is_root = True
        num_device = torch.cuda.device_count()
        devices = [f"cuda:{i}" for i in range(num_device)]
        if not devices:
            raise SystemExit("no CUDA device visible: this implementation has no CPU path")

    from .utils import DeviceManager
    stage_dir = None
    if distributed:
        stage_dir = setup_staging(args)
        device_manager = distributed_device_manager(args)
    else:
        device_manager = DeviceManager(args, devices)          # created once, like main.py:61
    output_scores, tokenizer = [], None
    # Prefix K/V cache: the reference recomputes everything for every new token (main.py:65); the
    # prefixes do not change between tokens, so every pass but the last keeps their per-layer K/V in HBM when they
    # fit and later passes run the option rows only.  Same scores bit for bit; FLS_PREFIX_KV_CACHE=off disables it.
    kv_user = getattr(args, "prefix_kv_cache", os.environ.get("FLS_PREFIX_KV_CACHE", "auto"))
    use_kv = kv_user != "off" and args.num_gen_token > 1
    for i_new in range(args.num_gen_token):
        if use_kv:
            args.prefix_kv_cache = "last" if i_new == args.num_gen_token - 1 else "on"
        prompts_now = list(input_prompts)
        outputs = (score_all_distributed(args, prompts_now, device_manager) if distributed
                   else score_all(args, prompts_now, devices, device_manager))
        if is_root:
            if i_new == 0:
                output_scores = outputs
                from .utils import load_tokenizer
                tokenizer = load_tokenizer(args.model_path)
                tokenizer.pad_token = tokenizer.eos_token
            else:
                output_scores = [np.concatenate((old, new), axis=1) for old, new in zip(output_scores, outputs)]
            # append the greedy token to every suffix (main.py:84-90)
            for idx in range(len(input_prompts)):
                prefix, suffix = original_input_prompts[idx]
                new_tokens = np.argmax(output_scores[idx], axis=-1)
                input_prompts[idx] = (prefix, tuple(s + tokenizer.decode(t) for s, t in zip(suffix, new_tokens)))
        if distributed and args.num_gen_token > 1:
            import torch.distributed as dist
            box = [input_prompts if is_root else None]
            dist.broadcast_object_list(box, src=0)
            input_prompts = box[0]
    args.prefix_kv_cache = kv_user
    for model in getattr(device_manager, "models", {}).values():
        model.drop_prefix_kv_cache()
    if stage_dir is not None:                 # every rank is done with the staged files before they go
        import shutil
        import torch.distributed as dist
        for model in getattr(device_manager, "models", {}).values():
            model.close()
        dist.barrier()
        if int(os.environ.get("LOCAL_RANK", "0")) == 0:
            shutil.rmtree(stage_dir, ignore_errors=True)
    if is_root:
        with open(args.prompt_pickle.replace(".pkl", "_updated.pkl"), "wb") as f:
            pickle.dump(input_prompts, f)
        with open(args.output_file, "wb") as f:
            pickle.dump(output_scores, f)
    return output_scores if is_root else None


if __name__ == "__main__":
    main()
