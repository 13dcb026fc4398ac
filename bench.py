#!/usr/bin/env python
"""Benchmark of the layer-sharded prompt-scoring path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One *step* = one full scoring pass: every shard of the model streamed from pinned host memory
through the double-buffered HBM slots, all prompts pushed through it, full-vocabulary fp16
probabilities produced.  Default workload = BASELINE.json configs[1]: Llama-2-7B architecture,
random-init fp16, layer_num_per_shard=1, storage_location=gpu, 1024 synthetic prompts x 512 token
rows (prefix 384 + 4 options x 32) per GPU.  With N>1 every rank scores its own 1024 prompts on its
own GPU (the `--data_parallel` prompt partitioning: independent units, no data-path collective), so
scaling is weak and `value` is the aggregate prompts/s.

Prints ONE JSON line (rank 0).  `value` = prompts/s with token ids resident in HBM and results left
in HBM; `e2e` = the same through the public API `ShardedLlama.__call__` from Python strings to numpy
probabilities on the host.  `roofline` is for the tcgen05 GEMM kernel (all four decoder GEMMs),
timed with CUDA events around every launch inside the timed region.  `cpu_baseline` / `--impl
reference` time the CPU oracle (the restated reference loop, one prompt at a time) on the host cores.
"""
from __future__ import annotations

import argparse
import copy
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from flexible_llm_sharding_b200 import synthetic  # noqa: E402

WORKLOADS = {
    # name: (shape, n_prompts, prefix_tokens, n_options, option_tokens)
    "7b-1024x512": ("7b", 1024, 384, 4, 32),
    "13b-1024x512": ("13b", 1024, 384, 4, 32),
    "70b-1024x512": ("70b", 1024, 384, 4, 32),
    "tiny-32x256": ("tiny", 32, 192, 4, 16),
}


# DRAM bytes (read + write) of one launch of each decoder GEMM at 16384 rows, 7B shapes, from the `ncu --set full`
# capture summarised in profiles/r01_ncu_full_prof_gemm2.csv (dram__bytes_read.sum + dram__bytes_write.sum).
# Static evidence, not measured by this run: quoted only for the workload the capture was taken on.
NCU_GEMM_TRAFFIC = {"7b": {"gemm_qkv": 0.716098e9 + 4.211310e6, "gemm_o": 0.465435e9 + 3.910457e6,
                           "gemm_gate_up": 1.080926e9 + 2.213196e6, "gemm_down": 2.097680e9 + 1.500891e6}}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(tensor_burst=d["bf16_tflops"], tensor_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    hbm=d["hbm_gbs"], source="measured (MEASURED_PEAKS.json)")
    return dict(tensor_burst=1590.0, tensor_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "reasons": sorted(reasons), "samples": len(sm)}


def model_dir_for(shape_name: str, rank: int, world: int, device, alias_layers: int = 0) -> str:
    """Random-init layer files in prepare_weights.py's layout (written once per box, rank 0)."""
    base = os.environ.get("FLS_BENCH_DIR") or ("/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir())
    d = os.path.join(base, f"fls_bench_{shape_name}" + (f"_alias{alias_layers}" if alias_layers else ""))
    done = os.path.join(d, ".complete")
    if rank == 0 and not os.path.exists(done):
        shutil.rmtree(d, ignore_errors=True)
        shape = synthetic.LlamaShape.named(shape_name)
        synthetic.write_layer_files(d, shape, seed=0, init_std=0.02,
                                    device=str(device) if torch.cuda.is_available() else "cpu",
                                    alias_decoder_layers=alias_layers)
        open(done, "w").close()
    if world > 1:
        torch.distributed.barrier()
    return d


def make_args(model_path, **kw):
    d = dict(model_path=model_path, num_batch=1, layer_num_per_shard=1, storage_location="gpu",
             max_activation_in_cpu=100, data_parallel=False, disk_folder="./temp", num_gen_token=1,
             max_rows=16384, max_head_rows=4096,
             resident_weights="never")     # a bench step streams every shard from pinned host memory
    d.update(kw)
    return types.SimpleNamespace(**d)


def measure_h2d_gbs(device) -> float:
    n = 1 << 30
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=device)
    best = 0.0
    for _ in range(4):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); d.copy_(h, non_blocking=True); b.record(); b.synchronize()
        best = max(best, n / (a.elapsed_time(b) * 1e-3) / 1e9)
    return best


def cpu_oracle_rate(model_path: str, prompts, n_layers_full: int, sample_prompts: int):
    """prompts/s of the restated reference loop on the host cores, from a bounded sample:
    `sample_prompts` prompts through 1 and through 2 decoder layers; the difference is the cost of
    one layer, the 1-layer run carries tokenisation + embedding + norm + lm_head."""
    from oracle import llama_oracle as O
    torch.set_num_threads(os.cpu_count())
    sample = copy.deepcopy(prompts[:sample_prompts])
    ts = []
    for nl in (1, 2):
        orc = O.OracleShardedLlama(model_path, dtype=torch.float32, max_decoder_layers=nl)
        t0 = time.perf_counter()
        orc(copy.deepcopy(sample))
        ts.append(time.perf_counter() - t0)
    per_layer = max(ts[1] - ts[0], 1e-9)
    full = ts[0] + (n_layers_full - 1) * per_layer
    return sample_prompts / full, dict(t_1layer_s=ts[0], t_2layer_s=ts[1], extrapolated_full_pass_s=full)


def bench_model_parallel(a, shape_name, shape, n_prompts, prefix_tokens, n_options, option_tokens, alias,
                         workload_name, pk):
    """Layer pipeline over all visible GPUs of ONE process (the reference's default multi-GPU mode,
    main.py:71-73): shard j on GPU j mod G, activations handed GPU-to-GPU over NVLink.  Steps are
    whole passes through main.score_all (public API, strings in / numpy out), timed on the host."""
    import __graft_entry__ as ge
    ge.build()
    from flexible_llm_sharding_b200.main import score_all
    G = torch.cuda.device_count()
    devices = [f"cuda:{i}" for i in range(G)]
    md = model_dir_for(shape_name, 0, 1, torch.device("cuda:0"), alias)
    prompts = synthetic.make_prompts(n_prompts, shape.vocab_size, prefix_tokens, n_options, option_tokens, seed=1234)
    args = make_args(md, layer_num_per_shard=a.layer_num_per_shard, storage_location="gpu", max_rows=a.max_rows)
    from flexible_llm_sharding_b200.utils import DeviceManager
    dm = DeviceManager(args, devices)        # created once per process, as main.py:61 does; models live on it
    for _ in range(max(1, a.warmup)):
        out = score_all(args, copy.deepcopy(prompts), devices, dm)
    for d in devices:
        torch.cuda.synchronize(d)
    for d in devices:                       # per-launch CUDA events: how busy each rank's GPU really is
        dm.models[d].ctx.call("fls_reset_stats")
        dm.models[d].ctx.call("fls_set_profiling", 1)
    with ClockSampler(0) as clocks:
        t0 = time.perf_counter()
        for _ in range(a.steps):
            out = score_all(args, copy.deepcopy(prompts), devices, dm)
        for d in devices:
            torch.cuda.synchronize(d)
        dt = time.perf_counter() - t0
    per_rank = {d: {k: round(v, 3) if isinstance(v, float) else v for k, v in dm.models[d].last_timing.items()}
                for d in devices}
    for d in devices:
        prof = dm.models[d].ctx.profile()
        dm.models[d].ctx.call("fls_set_profiling", 0)
        per_rank[d]["kernel_s_per_pass"] = round(sum(v["ms"] for v in prof.values()) / 1e3 / a.steps, 3)
    assert len(out) == n_prompts
    v = n_prompts * a.steps / dt
    rows_per_prompt = prefix_tokens + n_options * option_tokens
    print(json.dumps({
        "metric": "prompts/sec (layer-sharded prompt scoring)", "value": v, "unit": "prompts/s", "n_gpus": G,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f16 (fp32 accumulate)", "data": "synthetic",
        "config": {"workload": workload_name + (f" [{alias} distinct decoder layers, others aliased]" if alias else ""),
                   "parallelism": f"mp{G}: layer pipeline, shard j on GPU j mod {G}, NVLink peer hand-off",
                   "timing": "host wall clock around main.score_all (strings in / numpy out: tokenisation, all "
                             "ranks' passes, D2H of the scores; the DeviceManager and its per-GPU models are "
                             "created once before the warm-up pass, as in main.py)",
                   "last_pass_per_rank": per_rank},
        "prefill_tok_per_s": v * rows_per_prompt, "clocks": clocks.summary(),
        "e2e": {"value": v, "unit": "prompts/s"},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="7b-1024x512", choices=sorted(WORKLOADS))
    ap.add_argument("--prompts", type=int, default=0, help="override the prompt count (not the headline config)")
    ap.add_argument("--layer_num_per_shard", type=int, default=1)
    ap.add_argument("--storage_location", default="gpu")
    ap.add_argument("--max_rows", type=int, default=16384)
    ap.add_argument("--cpu_sample_prompts", type=int, default=8)
    ap.add_argument("--no_cpu_baseline", action="store_true")
    ap.add_argument("--no_e2e", action="store_true")
    ap.add_argument("--alias_layers", type=int, default=-1,
                    help="write only N distinct decoder layers and symlink the rest (default: 4 for 70b, else 0)")
    ap.add_argument("--parallel", default="dp", choices=["dp", "mp"],
                    help="mp: ONE process drives all visible GPUs as a layer pipeline (not under torchrun)")
    a = ap.parse_args()
    if os.environ.get("FLS_WATCHDOG_S"):      # debugging aid: dump every thread's stack and exit if the run stalls
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["FLS_WATCHDOG_S"]), exit=True)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    shape_name, n_prompts, prefix_tokens, n_options, option_tokens = WORKLOADS[a.workload]
    if a.prompts:
        n_prompts = a.prompts
    shape = synthetic.LlamaShape.named(shape_name)
    alias = a.alias_layers if a.alias_layers >= 0 else (4 if shape_name == "70b" else 0)
    rows_per_prompt = prefix_tokens + n_options * option_tokens
    workload_name = (f"Llama-2-{shape_name} arch random-init fp16, layer_num_per_shard={a.layer_num_per_shard}, "
                     f"storage_location={a.storage_location}, {n_prompts} prompts x {rows_per_prompt} token rows "
                     f"(prefix {prefix_tokens} + {n_options} options x {option_tokens}) per GPU")
    pk = peaks()

    # ------------------------------------------------------------------ reference arm (CPU oracle)
    if a.impl == "reference":
        if rank != 0:
            return
        gen_dev = torch.device("cuda:0") if torch.cuda.is_available() else "cpu"    # weight generation only
        md = model_dir_for(shape_name, 0, 1, gen_dev, alias)
        prompts = synthetic.make_prompts(max(a.cpu_sample_prompts, 1), shape.vocab_size, prefix_tokens, n_options,
                                         option_tokens, seed=1234)
        vals = []
        for _ in range(max(1, min(a.warmup, 1)) + max(1, a.steps)):
            v, detail = cpu_oracle_rate(md, prompts, shape.num_hidden_layers, a.cpu_sample_prompts)
            vals.append((v, detail))
        v = float(np.mean([x[0] for x in vals[1:]])) if len(vals) > 1 else vals[0][0]
        sample = (f"{a.cpu_sample_prompts} prompts x {rows_per_prompt} rows through 1 and 2 decoder layers + embed/norm/"
                  f"lm_head, fp32, extrapolated linearly to {shape.num_hidden_layers} layers")
        print(json.dumps({
            "impl": "reference", "metric": "prompts/sec (layer-sharded prompt scoring)", "value": v,
            "unit": "prompts/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * n_prompts / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (CPU oracle of the fp16 reference loop)", "data": "synthetic",
            "config": {"workload": workload_name},
            "cpu_baseline": {"value": v, "unit": "prompts/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": sample, **vals[-1][1]},
            "e2e": {"value": v, "unit": "prompts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "prefill_tok_per_s": v * rows_per_prompt,
        }))
        return

    # ------------------------------------------------------------------ B200 arm
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    device = torch.device(f"cuda:{local}")
    if world > 1:
        # stdout carries exactly one JSON line: keep NCCL's version banner out of it
        os.environ["NCCL_DEBUG"] = os.environ.get("FLS_NCCL_DEBUG", "WARN")
        torch.distributed.init_process_group("nccl", device_id=device)
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if world > 1:
        torch.distributed.barrier()
    from flexible_llm_sharding_b200 import planner
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama

    if a.parallel == "mp":
        return bench_model_parallel(a, shape_name, shape, n_prompts, prefix_tokens, n_options, option_tokens, alias,
                                    workload_name, pk)
    md = model_dir_for(shape_name, rank, world, device, alias)
    prompts = synthetic.make_prompts(n_prompts, shape.vocab_size, prefix_tokens, n_options, option_tokens,
                                     seed=1234 + rank)
    args = make_args(md, layer_num_per_shard=a.layer_num_per_shard, storage_location=a.storage_location,
                     max_rows=a.max_rows)
    model = ShardedLlama(args, DeviceManager(args, [str(device)]), device=str(device))
    tok = planner.tokenize_prompts(model.tokenizer, prompts)
    packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, model.prefix_causal)
    link_gbs = measure_h2d_gbs(device)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def step_device():
        model.run(None, want_probs=True, packed=packed, return_device=True)

    for _ in range(a.warmup):
        step_device()
    barrier()
    model.ctx.call("fls_reset_stats")
    model.ctx.call("fls_set_profiling", 1)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        ev0.record()
        for _ in range(a.steps):
            step_device()
        ev1.record()
        barrier()
    dt = ev0.elapsed_time(ev1) * 1e-3
    prof = model.ctx.profile()
    stats = model.ctx.stats()
    model.ctx.call("fls_set_profiling", 0)
    launches = int(stats.kernel_launches)
    if world > 1:
        t = torch.tensor([dt], device=device, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        dt = float(t.item())
    value = world * n_prompts * a.steps / dt

    # ---- end-to-end through the public API (strings in, numpy probabilities out)
    e2e = None
    if not a.no_e2e:
        model(copy.deepcopy(prompts))           # warm the path (tokenizer, pinned output buffers)
        barrier()
        t0 = time.perf_counter()
        h2d = d2h = 0
        s0 = model.ctx.stats().h2d_bytes
        for _ in range(a.steps):
            out = model(copy.deepcopy(prompts))
            h2d += model.last_timing["h2d_input_bytes"]
            d2h += model.last_timing["d2h_output_bytes"]
        barrier()
        et = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([et], device=device, dtype=torch.float64)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            et = float(t.item())
        w_h2d = (model.ctx.stats().h2d_bytes - s0)
        e2e = {"value": world * n_prompts * a.steps / et, "unit": "prompts/s",
               "h2d_bytes_per_step": int((h2d + w_h2d) / a.steps), "d2h_bytes_per_step": int(d2h / a.steps),
               "tokenize_s_per_step": model.last_timing["tokenize_s"],
               "note": "h2d = token ids/positions/tile descriptors + the weight stream (every shard, every step)"}
        assert len(out) == n_prompts and out[0].shape == (n_options, 1, shape.vocab_size)

    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    # ---- roofline of the dominant kernel: the tcgen05 GEMM (4 decoder GEMMs per layer)
    gemm_cls = ["gemm_qkv", "gemm_o", "gemm_gate_up", "gemm_down"]
    g_ms = sum(prof[c]["ms"] for c in gemm_cls)
    g_fl = sum(prof[c]["work"] for c in gemm_cls)
    g_n = sum(prof[c]["launches"] for c in gemm_cls)
    achieved = g_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0
    peak = pk["tensor_sustained"]
    per_class = {}
    for c, v in prof.items():
        if v["launches"]:
            rate = v["work"] / (v["ms"] * 1e-3) if v["ms"] > 0 else 0.0
            per_class[c] = {"ms_per_step": v["ms"] / a.steps, "launches_per_step": v["launches"] / a.steps,
                            "share_of_step": v["ms"] * 1e-3 / dt,
                            ("tflops" if c.startswith("gemm") else "gbs"): rate / (1e12 if c.startswith("gemm") else 1e9)}
    traffic, traffic_source = None, None
    ncu = NCU_GEMM_TRAFFIC.get(shape_name)
    if ncu and a.max_rows == 16384 and g_n:
        # launch-weighted mean over the four GEMM classes, like `achieved` / `avg_launch_ms`
        traffic = sum(ncu[c] * prof[c]["launches"] for c in gemm_cls) / g_n
        traffic_source = "profiles/r01_ncu_full_prof_gemm2.csv (ncu --set full, one launch per class, bytes/launch)"
    roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak if peak else None, "traffic": traffic, "traffic_source": traffic_source,
                "kernel": "gemm_tcgen05_kernel (qkv, o+residual, gate/up+SwiGLU, down+residual)",
                "avg_launch_ms": g_ms / max(g_n, 1), "flops_per_launch": g_fl / max(g_n, 1),
                "peak_source": pk["source"] + ", sustained figure (kernel timed inside a long step)",
                "per_class": per_class}

    # ---- whole-pass roofline: per shard max(GEMM flops / tensor peak, weight bytes / host link)
    P_layer = 2 * shape.hidden_size * (shape.num_attention_heads + shape.num_key_value_heads) * shape.head_dim \
        + 3 * shape.hidden_size * shape.intermediate_size
    rows = n_prompts * rows_per_prompt
    t_layer = max(2.0 * P_layer * rows / (peak * 1e12), 2.0 * P_layer / (link_gbs * 1e9))
    t_head = max(2 * 2.0 * shape.vocab_size * shape.hidden_size * n_prompts * n_options / (peak * 1e12),
                 2.0 * shape.vocab_size * shape.hidden_size / (link_gbs * 1e9))
    t_embed = 2.0 * shape.vocab_size * shape.hidden_size / (link_gbs * 1e9)
    ideal = shape.num_hidden_layers * t_layer + t_head + t_embed
    path_roofline = {"ideal_s_per_step": ideal, "measured_s_per_step": dt / a.steps, "frac": ideal / (dt / a.steps),
                     "h2d_link_gbs_measured": link_gbs,
                     "weight_stream_gbs": stats.h2d_bytes / (stats.h2d_ms * 1e-3) / 1e9 if stats.h2d_ms > 0 else None,
                     "weight_bytes_per_step": int(stats.h2d_bytes / a.steps),
                     "definition": "sum over shards of max(2*P*rows/tensor_peak, weight_bytes/host_link)"}

    cpu = None
    if not a.no_cpu_baseline:
        v, detail = cpu_oracle_rate(md, prompts, shape.num_hidden_layers, a.cpu_sample_prompts)
        cpu = {"value": v, "unit": "prompts/s", "cores": os.cpu_count(), "kind": "port",
               "sample": (f"{a.cpu_sample_prompts} prompts x {rows_per_prompt} rows through 1 and 2 decoder layers + "
                          f"embed/norm/lm_head, fp32 oracle, extrapolated linearly to {shape.num_hidden_layers} layers"),
               **detail}

    line = {
        "metric": "prompts/sec (layer-sharded prompt scoring)", "value": value, "unit": "prompts/s",
        "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 (fp32 accumulate)",
        "data": "synthetic",
        "config": {"workload": workload_name + (f" [{alias} distinct decoder layers, others aliased]" if alias else ""),
                   "micro_batch_rows": a.max_rows,
                   "cache": "inputs larger than L2 (activations %.1f GB, weights %.1f GB per step)" % (
                       rows * shape.hidden_size * 2 / 1e9, stats.h2d_bytes / a.steps / 1e9),
                   "parallelism": f"dp{world} (independent prompt partitions, no collective)"},
        "prefill_tok_per_s": value * rows_per_prompt,
        "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": launches,
        "roofline": roofline, "path_roofline": path_roofline, "cpu_baseline": cpu,
    }
    print(json.dumps(line))
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
