#!/usr/bin/env python
"""Benchmark of the layer-sharded prompt-scoring path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One *step* = one full scoring pass: every shard of the model streamed from page-locked host memory through the
double-buffered HBM slots, all prompts pushed through it, full-vocabulary fp16 probabilities produced.

The JSON line (rank 0, last line of stdout):
  value / e2e / roofline / path_roofline   BASELINE.json configs[1]: Llama-2-7B architecture, random-init fp16,
        layer_num_per_shard=1, storage_location=gpu, 1024 synthetic prompts x 512 token rows per GPU, K timed
        steps after W warm-up steps.  With N>1 every rank scores its own 1024 prompts (`--data_parallel`: independent
        units, no data-path collective; weak scaling) and `value` is the aggregate.
  headline (N=1)   the configuration the metric is QUOTED on - Llama-2-70B, 1024 x 512, one GPU, 138 GB of weights
        streamed per pass - measured in the same process: 1 warm-up + `--headline_steps` (2) timed passes with inputs
        resident in HBM plus one pass end to end through the public API.  A 70B pass takes about a minute, so the
        driver's `--steps 20 --warmup 5` cannot be spent on it (BASELINE.md section 6 has the time budget); the block
        carries its own roofline / path_roofline / e2e / weight-stream figures, the number of distinct decoder layers
        and the host bytes really page-locked.
  dp_13b_k4 (N>1)  BASELINE.json configs[2]: Llama-2-13B, layer_num_per_shard=4, --data_parallel, 1024 prompts per rank.
  mp (N>1)         BASELINE.json configs[3]: the same 70B workload as ONE job over the N ranks - the reference's
        default model-parallel layer pipeline (shard j on rank j mod N, rows handed rank to rank over NVLink):
        strong scaling, with a tiny-model check that the N-rank pipeline returns the 1-GPU scores bit for bit.
  cpu_baseline / parity_check (N=1)   the CPU oracle (restated reference loop, fp32) timed on a bounded sample on
        the host cores, and its scores compared with the CUDA path's on the same sample; gpu_eager_baseline = the same
        loop in fp16 on the GPU through torch eager + cuBLAS (the reference's own arithmetic path on this box).
`--impl reference` times the CPU oracle alone (the reference has no CPU path and cannot be installed here).
"""
from __future__ import annotations

import argparse
import copy
import csv
import gc
import glob
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
T_START = time.perf_counter()

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
GEMM_CLASSES = ["gemm_qkv", "gemm_o", "gemm_gate_up", "gemm_down"]
METRIC = "prompts/sec (layer-sharded prompt scoring)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(tensor_burst=d["bf16_tflops"], tensor_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    hbm=d["hbm_gbs"], source="measured (MEASURED_PEAKS.json)")
    return dict(tensor_burst=1590.0, tensor_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


def ncu_gemm_traffic(shape_name: str):
    """DRAM bytes (read + write) of one launch of each decoder GEMM from the newest committed `ncu --set full`
    summary for this model shape: profiles/r*_ncu_full_gemm_<shape>.csv, one row per launch in the order
    qkv, o, gate/up, down (tools/gemm_probe.py --ncu launches them in that order), columns found BY NAME and
    scaled by the unit row.  Static evidence of that capture, not measured by this run."""
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", f"r*_ncu_full_gemm_{shape_name}.csv")))
    if not files:
        return None, None
    unit_scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    with open(files[-1], newline="") as f:
        rows = list(csv.reader(f))
    if len(rows) < 6:
        return None, None
    head, units, data = rows[0], rows[1], [r for r in rows[2:] if r and "gemm" in r[0]]
    try:
        ir, iw = head.index("dram__bytes_read.sum"), head.index("dram__bytes_write.sum")
    except ValueError:
        return None, None
    if len(data) < 4 or units[ir] not in unit_scale or units[iw] not in unit_scale:
        return None, None
    out = {}
    for cls, r in zip(GEMM_CLASSES, data[:4]):
        out[cls] = float(r[ir]) * unit_scale[units[ir]] + float(r[iw]) * unit_scale[units[iw]]
    return out, os.path.relpath(files[-1], ROOT)


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


# ------------------------------------------------------------------------------------------ synthetic model files
def bench_base_dir() -> str:
    return os.environ.get("FLS_BENCH_DIR") or ("/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir())


def mem_available_bytes() -> int:
    try:
        with open("/proc/meminfo") as f:
            return next(int(l.split()[1]) for l in f if l.startswith("MemAvailable")) * 1024
    except (OSError, StopIteration, ValueError):
        return 0


def model_bytes(shape) -> int:
    return sum(2 * int(np.prod(s)) for ln in shape.layer_names() for s in shape.layer_tensor_shapes(ln).values())


def choose_alias(shape, requested: int, reserve_bytes: int = 24 << 30) -> int:
    """0 = every decoder layer distinct.  The files go to tmpfs and are page-locked in place (no second copy), so
    a model needs its own size in host RAM once; if even that does not fit, only as many distinct decoder layers
    as do are written and the rest are symlinks to them (stated in the results)."""
    if requested >= 0:
        return requested
    d = os.path.join(bench_base_dir(), f"fls_bench_70b_L{shape.num_hidden_layers}")
    if os.path.exists(os.path.join(d, ".complete")):
        return 0                                          # already written in full on this box
    need, avail = model_bytes(shape), mem_available_bytes()
    if avail == 0 or need + reserve_bytes <= avail:
        return 0
    per_layer = sum(2 * int(np.prod(s)) for s in shape.layer_tensor_shapes("model.layers.0").values())
    fixed = need - shape.num_hidden_layers * per_layer
    return int(max(1, min(shape.num_hidden_layers - 1, (avail - reserve_bytes - fixed) // per_layer)))


def model_dir_for(shape_name: str, rank: int, world: int, device, alias_layers: int = 0, shape=None) -> str:
    """Random-init layer files in prepare_weights.py's layout, written once per box: every rank writes the layers
    whose index is congruent to its rank (generated on its GPU, streamed to tmpfs by writer threads)."""
    shape = shape or synthetic.LlamaShape.named(shape_name)
    d = os.path.join(bench_base_dir(), f"fls_bench_{shape_name}_L{shape.num_hidden_layers}"
                     + (f"_alias{alias_layers}" if alias_layers else ""))
    done = os.path.join(d, ".complete")
    have = os.path.exists(done)
    if world > 1:
        flag = torch.tensor([int(have)], device=device)
        torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN)
        have = bool(flag.item())
    if not have:
        if rank == 0:
            shutil.rmtree(d, ignore_errors=True)
            os.makedirs(d, exist_ok=True)
        if world > 1:
            torch.distributed.barrier()
        names = shape.layer_names()
        mine = {ln for i, ln in enumerate(names) if i % world == rank}
        synthetic.write_layer_files(d, shape, seed=0, init_std=0.02,
                                    device=str(device) if torch.cuda.is_available() else "cpu",
                                    alias_decoder_layers=alias_layers, only_layers=mine, fast=True)
        if world > 1:
            torch.distributed.barrier()
        if rank == 0:
            open(done, "w").close()
    if world > 1:
        torch.distributed.barrier()
    return d


def truncated_model_view(model_dir: str, n_layers: int) -> str:
    """A directory that presents the first `n_layers` decoder layers of `model_dir` as a whole model (symlinks + a
    config.json with num_hidden_layers = n_layers): what the bounded CPU sample runs through."""
    d = tempfile.mkdtemp(prefix="fls_bench_view_")
    with open(os.path.join(model_dir, "config.json")) as f:
        cfg = json.load(f)
    cfg["num_hidden_layers"] = n_layers
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump(cfg, f)
    keep = {"model.embed_tokens", "model.norm", "lm_head"} | {f"model.layers.{i}" for i in range(n_layers)}
    for fn in os.listdir(model_dir):
        if fn.endswith(".safetensors"):
            if fn[:-len(".safetensors")] in keep:
                os.symlink(os.path.realpath(os.path.join(model_dir, fn)), os.path.join(d, fn))
        elif fn not in ("config.json", ".complete"):
            os.symlink(os.path.realpath(os.path.join(model_dir, fn)), os.path.join(d, fn))
    return d


def make_args(model_path, **kw):
    d = dict(model_path=model_path, num_batch=1, layer_num_per_shard=1, storage_location="gpu",
             max_activation_in_cpu=100, data_parallel=False, disk_folder="./temp", num_gen_token=1,
             max_rows=16384, max_head_rows=4096,
             resident_weights="never")     # a bench step streams every shard from page-locked host memory
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


def release_host_images():
    gc.collect()
    from flexible_llm_sharding_b200 import _lib
    _lib.load().fls_trim_host_cache(0)


# ------------------------------------------------------------------------------------------ CPU / eager baselines
def cpu_oracle_sample(model_path: str, prompts, n_layers_full: int, repeats: int = 3):
    """prompts/s of the restated reference loop on the host cores from a bounded sample: the sample prompts
    through 1 and through 2 decoder layers (median of `repeats` runs each); the difference is the cost of one
    layer, the 1-layer run carries tokenisation + embedding + norm + lm_head.  Returns (rate, detail, scores of
    the 2-layer run) - rate None when the two timings are not ordered (a disturbed sample is not extrapolated)."""
    from oracle import llama_oracle as O
    torch.set_num_threads(os.cpu_count())
    ts, scores = {}, None
    for nl in (1, 2):
        view = truncated_model_view(model_path, nl)
        try:
            runs = []
            for _ in range(repeats):
                orc = O.OracleShardedLlama(view, dtype=torch.float32)
                t0 = time.perf_counter()
                out = orc(copy.deepcopy(prompts))
                runs.append(time.perf_counter() - t0)
            ts[nl] = float(np.median(runs))
            if nl == 2:
                scores = out
        finally:
            shutil.rmtree(view, ignore_errors=True)
    detail = dict(t_1layer_s=ts[1], t_2layer_s=ts[2], repeats=repeats)
    if ts[2] <= ts[1]:
        detail["rejected"] = "t(2 layers) <= t(1 layer): the host was disturbed during the sample"
        return None, detail, scores
    full = ts[1] + (n_layers_full - 1) * (ts[2] - ts[1])
    detail["extrapolated_full_pass_s"] = full
    return len(prompts) / full, detail, scores


def parity_against(scores_ref, scores_got, answer_token: int = 5, p_min: float = 1e-4, tol: float = 5e-2):
    ref = np.stack([s[:, 0] for s in scores_ref]).astype(np.float64)
    got = np.stack([s[:, 0] for s in scores_got]).astype(np.float64)
    m = ref >= p_min
    dlog = float(np.abs(np.log(np.maximum(got[m], 1e-30)) - np.log(ref[m])).max())
    srt = np.sort(np.log(np.maximum(ref, 1e-30)), axis=-1)
    rankable = (srt[..., -1] - srt[..., -2]) > 2 * tol
    top1 = bool((got.argmax(-1)[rankable] == ref.argmax(-1)[rankable]).all())
    rl, gl = np.log(np.maximum(ref[..., answer_token], 1e-30)), np.log(np.maximum(got[..., answer_token], 1e-30))
    same, rankable_prompts = True, 0
    for p in range(ref.shape[0]):
        order = np.argsort(-rl[p])
        if (-np.diff(rl[p][order]) > 2 * tol).all():
            rankable_prompts += 1
            same &= bool((np.argsort(-gl[p]) == order).all())
    return dict(max_dlogp=dlog, tolerance=tol, within_tolerance=dlog <= tol, top1_identical=top1,
                rankings_identical=same, prompts_with_rankable_margins=rankable_prompts, tokens_compared=int(m.sum()))


def gpu_eager_baseline(model_path: str, prompts, n_prompts_full: int, device):
    """The restated reference loop in fp16 on the GPU (torch eager + cuBLAS, one prompt at a time, every layer
    file loaded per pass like utils.py:121-131): the reference's own arithmetic path on this box.  Two sample
    sizes separate the per-pass cost (weight loading) from the per-prompt cost; the rate is for the full prompt
    count."""
    from oracle import llama_oracle as O
    ts = {}
    for n in (max(1, len(prompts) // 4), len(prompts)):
        orc = O.OracleShardedLlama(model_path, dtype=torch.float16, device=str(device))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        orc(copy.deepcopy(prompts[:n]))
        torch.cuda.synchronize()
        ts[n] = time.perf_counter() - t0
    n1, n2 = sorted(ts)
    per_prompt = max((ts[n2] - ts[n1]) / max(n2 - n1, 1), 1e-9)
    per_pass = max(ts[n2] - n2 * per_prompt, 0.0)
    full = per_pass + n_prompts_full * per_prompt
    return {"value": n_prompts_full / full, "unit": "prompts/s",
            "kind": "restated reference loop, torch eager fp16 + cuBLAS on this GPU, one prompt at a time, layer files "
                    "loaded per pass",
            "sample": f"{n1} and {n2} prompts through all layers; per-pass and per-prompt cost separated, "
                      f"extrapolated to {n_prompts_full} prompts",
            "per_prompt_s": per_prompt, "per_pass_s": per_pass, "t_samples_s": {str(k): v for k, v in ts.items()}}


# ------------------------------------------------------------------------------------------ timing + rooflines
def timed_device_passes(model, packed, steps, warmup, world, device, local):
    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        model.run(None, want_probs=True, packed=packed, return_device=True)
    barrier()
    model.ctx.call("fls_reset_stats")
    model.ctx.call("fls_set_profiling", 1)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        ev0.record()
        for _ in range(steps):
            model.run(None, want_probs=True, packed=packed, return_device=True)
        ev1.record()
        barrier()
    dt = ev0.elapsed_time(ev1) * 1e-3
    prof = model.ctx.profile()
    stats = model.ctx.stats()
    model.ctx.call("fls_set_profiling", 0)
    if world > 1:
        t = torch.tensor([dt], device=device, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        dt = float(t.item())
    return dt, prof, stats, clocks.summary()


def timed_e2e(model, prompts, steps, world, device, warm=True):
    """`steps` passes end to end through the public API (strings in, numpy probabilities out)."""
    if warm:
        model(copy.deepcopy(prompts))           # tokenizer, pinned output buffers
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    h2d = d2h = 0
    s0 = model.ctx.stats().h2d_bytes
    for _ in range(steps):
        out = model(copy.deepcopy(prompts))
        h2d += model.last_timing["h2d_input_bytes"]
        d2h += model.last_timing["d2h_output_bytes"]
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    et = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([et], device=device, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        et = float(t.item())
    w_h2d = model.ctx.stats().h2d_bytes - s0
    return et, out, dict(h2d_bytes_per_step=int((h2d + w_h2d) / steps), d2h_bytes_per_step=int(d2h / steps),
                         tokenize_s_per_step=model.last_timing["tokenize_s"],
                         note="h2d = token ids/positions/tile descriptors + the weight stream (every shard, every step)")


def gemm_roofline(prof, dt, steps, pk, shape_name, max_rows):
    g_ms = sum(prof[c]["ms"] for c in GEMM_CLASSES)
    g_fl = sum(prof[c]["work"] for c in GEMM_CLASSES)
    g_n = sum(prof[c]["launches"] for c in GEMM_CLASSES)
    achieved = g_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0
    per_class = {}
    for c, v in prof.items():
        if v["launches"]:
            rate = v["work"] / (v["ms"] * 1e-3) if v["ms"] > 0 else 0.0
            per_class[c] = {"ms_per_step": v["ms"] / steps, "launches_per_step": v["launches"] / steps,
                            "share_of_step": v["ms"] * 1e-3 / dt,
                            ("tflops" if c.startswith("gemm") else "gbs"): rate / (1e12 if c.startswith("gemm") else 1e9)}
    traffic, traffic_source = None, None
    ncu, src = ncu_gemm_traffic(shape_name)
    if ncu and max_rows == 16384 and g_n:
        traffic = sum(ncu[c] * prof[c]["launches"] for c in GEMM_CLASSES) / g_n       # launch-weighted, like `achieved`
        traffic_source = f"{src} (ncu --set full, one launch per class at 16384 rows, dram read + write bytes)"
    peak = pk["tensor_sustained"]
    return {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
            "frac": achieved / peak if peak else None, "frac_of_burst_peak": achieved / pk["tensor_burst"],
            "peak_burst": pk["tensor_burst"], "traffic": traffic, "traffic_source": traffic_source,
            "kernel": "gemm_tcgen05_kernel (qkv+RoPE, o+residual, gate/up+SwiGLU, down+residual)",
            "avg_launch_ms": g_ms / max(g_n, 1), "flops_per_launch": g_fl / max(g_n, 1),
            "peak_source": pk["source"] + ": `peak` is the sustained figure (kernel timed inside a long step), "
                           "`peak_burst` the best-of-10 one",
            "per_class": per_class}


def path_roofline(shape, n_prompts, n_options, rows_per_prompt, link_gbs, pk, dt_step, stats, steps, n_ranks=1):
    """Sum over shards of max(GEMM FLOPs / tensor peak, weight bytes / host link); with the layer pipeline over
    n_ranks GPUs every rank computes and streams 1/n_ranks of the shards at the same time."""
    P_layer = 2 * shape.hidden_size * (shape.num_attention_heads + shape.num_key_value_heads) * shape.head_dim \
        + 3 * shape.hidden_size * shape.intermediate_size
    rows = n_prompts * rows_per_prompt
    out = {"definition": "sum over shards of max(2*P*rows/tensor_peak, weight_bytes/host_link)"
                         + (f", divided by the {n_ranks} pipeline ranks" if n_ranks > 1 else "")}
    for name, peak in (("", pk["tensor_sustained"]), ("_burst", pk["tensor_burst"])):
        t_layer = max(2.0 * P_layer * rows / (peak * 1e12), 2.0 * P_layer / (link_gbs * 1e9))
        t_head = max(2 * 2.0 * shape.vocab_size * shape.hidden_size * n_prompts * n_options / (peak * 1e12),
                     2.0 * shape.vocab_size * shape.hidden_size / (link_gbs * 1e9))
        t_embed = 2.0 * shape.vocab_size * shape.hidden_size / (link_gbs * 1e9)
        ideal = (shape.num_hidden_layers * t_layer + t_head + t_embed) / n_ranks
        out["ideal_s_per_step" + name] = ideal
        out["frac" + name] = ideal / dt_step
    out.update(measured_s_per_step=dt_step, h2d_link_gbs_measured=link_gbs,
               weight_stream_gbs=stats.h2d_bytes / (stats.h2d_ms * 1e-3) / 1e9 if stats.h2d_ms > 0 else None,
               weight_bytes_per_step=int(stats.h2d_bytes / steps),
               peaks="frac vs the sustained tensor peak, frac_burst vs the burst peak")
    return out


# ------------------------------------------------------------------------------------------ 70B headline (N = 1)
def headline_block(a, pk, device, local, link_gbs):
    from flexible_llm_sharding_b200 import planner
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    shape_name, n_prompts, prefix_tokens, n_options, option_tokens = WORKLOADS["70b-1024x512"]
    if a.headline_prompts:
        n_prompts = a.headline_prompts
    shape = synthetic.LlamaShape.named(shape_name)
    rows_per_prompt = prefix_tokens + n_options * option_tokens
    t_setup = time.perf_counter()
    alias = choose_alias(shape, a.alias_layers)
    md = model_dir_for(shape_name, 0, 1, device, alias)
    t_files = time.perf_counter() - t_setup
    prompts = synthetic.make_prompts(n_prompts, shape.vocab_size, prefix_tokens, n_options, option_tokens, seed=1234)
    args = make_args(md, layer_num_per_shard=1, storage_location="gpu", max_rows=a.max_rows, pinned_limit_bytes=0)
    model = ShardedLlama(args, DeviceManager(args, [str(device)]), device=str(device))
    tok = planner.tokenize_prompts(model.tokenizer, prompts)
    packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, model.prefix_causal)
    t_warm = time.perf_counter()
    model.run(None, want_probs=True, packed=packed, return_device=True)      # warm-up pass: files page-locked
    torch.cuda.synchronize()
    t_warm = time.perf_counter() - t_warm
    dt, prof, stats, clocks = timed_device_passes(model, packed, a.headline_steps, 0, 1, device, local)
    dt_step = dt / a.headline_steps
    value = n_prompts / dt_step
    # The end-to-end pass is a fourth minute-long pass: it is dropped (and said so) when the run is already long, so
    # that the driver's per-run limit is never what ends this program.
    spent = time.perf_counter() - T_START
    if spent > float(os.environ.get("FLS_BENCH_E2E_CUTOFF_S", "650")):
        et, e2e = None, {"skipped": f"run already {spent:.0f}s long"}
    else:
        et, out, e2e = timed_e2e(model, prompts, 1, 1, device, warm=False)
        assert len(out) == n_prompts and out[0].shape == (n_options, 1, shape.vocab_size)
    blk = {
        "workload": f"Llama-2-70b arch random-init fp16, layer_num_per_shard=1, storage_location=gpu, {n_prompts} "
                    f"prompts x {rows_per_prompt} token rows (prefix {prefix_tokens} + {n_options} options x "
                    f"{option_tokens}), 1 GPU, every shard streamed from page-locked host memory every pass",
        "metric": METRIC, "value": value, "unit": "prompts/s", "prefill_tok_per_s": value * rows_per_prompt,
        "steps": a.headline_steps, "warmup": 1, "ms_per_step": 1e3 * dt_step,
        "e2e": {"value": n_prompts / et if et else None, "unit": "prompts/s", "steps": 1, **e2e},
        "roofline": gemm_roofline(prof, dt, a.headline_steps, pk, shape_name, a.max_rows),
        "path_roofline": path_roofline(shape, n_prompts, n_options, rows_per_prompt, link_gbs, pk, dt_step, stats,
                                       a.headline_steps),
        "clocks": clocks, "gpu_launches": int(stats.kernel_launches),
        "distinct_decoder_layers": alias or shape.num_hidden_layers,
        "aliased_decoder_layers": (shape.num_hidden_layers - alias) if alias else 0,
        "host_bytes_page_locked": int(stats.pinned_bytes), "of_which_file_pages_locked_in_place": int(stats.mapped_bytes),
        "model_bytes": model_bytes(shape),
        "setup_s": {"write_layer_files": round(t_files, 1), "warmup_pass_incl_page_locking": round(t_warm, 1)},
    }
    model.ctx.close()
    del model
    release_host_images()
    return blk


# ------------------------------------------------------------------------------------------ 13B, k = 4, data parallel (N > 1)
def c3_block(a, pk, device, local, rank, world, link_gbs):
    """configs[2]: Llama-2-13B, layer_num_per_shard = 4, --data_parallel: every rank scores its own 1024 prompts.
    The layer files lie on tmpfs and every rank page-locks the SAME pages, so the box holds one copy of the model."""
    from flexible_llm_sharding_b200 import planner
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    shape_name, n_prompts, prefix_tokens, n_options, option_tokens = WORKLOADS["13b-1024x512"]
    shape = synthetic.LlamaShape.named(shape_name)
    rows_per_prompt = prefix_tokens + n_options * option_tokens
    md = model_dir_for(shape_name, rank, world, device, 0)
    prompts = synthetic.make_prompts(n_prompts, shape.vocab_size, prefix_tokens, n_options, option_tokens,
                                     seed=4321 + rank)
    args = make_args(md, layer_num_per_shard=4, storage_location="gpu", max_rows=a.max_rows, pinned_limit_bytes=0,
                     data_parallel=True)
    model = ShardedLlama(args, DeviceManager(args, [str(device)]), device=str(device))
    tok = planner.tokenize_prompts(model.tokenizer, prompts)
    packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, model.prefix_causal)
    dt, prof, stats, clocks = timed_device_passes(model, packed, a.headline_steps, 1, world, device, local)
    dt_step = dt / a.headline_steps
    value = world * n_prompts / dt_step
    with open("/proc/meminfo") as f:
        shmem_kb = next((int(l.split()[1]) for l in f if l.startswith("Shmem:")), 0)
    blk = {
        "workload": f"Llama-2-13b arch random-init fp16, layer_num_per_shard=4, storage_location=gpu, {n_prompts} prompts "
                    f"x {rows_per_prompt} token rows per GPU, --data_parallel over {world} GPUs (weak scaling)",
        "metric": METRIC, "value": value, "unit": "prompts/s", "prefill_tok_per_s": value * rows_per_prompt,
        "steps": a.headline_steps, "warmup": 1, "ms_per_step": 1e3 * dt_step, "scaling": "weak",
        "roofline": gemm_roofline(prof, dt, a.headline_steps, pk, shape_name, a.max_rows),
        "path_roofline": path_roofline(shape, n_prompts, n_options, rows_per_prompt, link_gbs, pk, dt_step, stats,
                                       a.headline_steps),
        "clocks_rank0": clocks, "model_bytes": model_bytes(shape),
        "host_bytes_page_locked_per_rank": int(stats.pinned_bytes),
        "of_which_file_pages_locked_in_place_per_rank": int(stats.mapped_bytes),
        "box_shmem_bytes": shmem_kb * 1024,
        "note": "every rank locks the same tmpfs pages: the box holds one copy of the model (box_shmem_bytes counts all "
                "benchmark models lying in /dev/shm), not one per rank",
    }
    model.close()
    del model
    release_host_images()
    if world > 1:
        torch.distributed.barrier()
    return blk if rank == 0 else None


# ------------------------------------------------------------------------------------------ 70B layer pipeline (N > 1)
def mp_block(a, pk, device, local, rank, world, link_gbs):
    """configs[3]: one 70B scoring job over all ranks as the reference's default model-parallel layer pipeline
    (main.py:71-73), one process per GPU, rows handed rank to rank over NVLink (pipeline_dist.py)."""
    import torch.distributed as dist
    from flexible_llm_sharding_b200 import main as M
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    shape_name, n_prompts, prefix_tokens, n_options, option_tokens = WORKLOADS["70b-1024x512"]
    if a.headline_prompts:
        n_prompts = a.headline_prompts
    shape = synthetic.LlamaShape.named(shape_name)
    rows_per_prompt = prefix_tokens + n_options * option_tokens
    dev_name = f"cuda:{local}"

    # 1. correctness first, on a model small enough for one GPU to score alone: N-rank pipeline == 1 GPU, bit for bit
    tiny = synthetic.LlamaShape.named("tiny-gqa")
    td = model_dir_for("tiny-gqa", rank, world, device, 0)
    tprompts = synthetic.make_prompts(24, tiny.vocab_size, 40, 3, 10, seed=99, ragged=True)
    targs = make_args(td, layer_num_per_shard=1, max_rows=256)
    tdm = M.distributed_device_manager(targs)
    got = M.run_sharded_llama(dev_name, copy.deepcopy(tprompts), tdm, targs)
    gathered = [None] * world
    dist.all_gather_object(gathered, got)
    got = sum(gathered, [])
    identical = None
    if rank == 0:
        solo = ShardedLlama(targs, DeviceManager(targs, [str(device)]), device=str(device))
        want = solo(copy.deepcopy(tprompts))
        identical = len(got) == len(want) and all(np.array_equal(x, y) for x, y in zip(got, want))
        solo.ctx.close()
    for m in tdm.models.values():
        m.close()
    tdm.models.clear()
    dist.barrier()

    # 2. the 70B job
    t_setup = time.perf_counter()
    alias = choose_alias(shape, a.alias_layers)
    flag = torch.tensor([alias], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    alias = int(flag.item())
    md = model_dir_for(shape_name, rank, world, device, alias)
    t_files = time.perf_counter() - t_setup
    prompts = synthetic.make_prompts(n_prompts, shape.vocab_size, prefix_tokens, n_options, option_tokens, seed=1234)
    args = make_args(md, layer_num_per_shard=1, storage_location="gpu", max_rows=a.max_rows, pinned_limit_bytes=0)
    dm = M.distributed_device_manager(args)

    def one_pass():
        return M.run_sharded_llama(dev_name, copy.deepcopy(prompts), dm, args)

    one_pass()                                                   # warm-up: files page-locked, rings exported
    model = dm.models[dev_name]
    torch.cuda.synchronize()
    dist.barrier()
    model.ctx.call("fls_reset_stats")
    model.ctx.call("fls_set_profiling", 1)
    with ClockSampler(local) as clocks:
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(a.headline_steps):
            out = one_pass()
        torch.cuda.synchronize()
        dist.barrier()
        dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt_step = float(t.item()) / a.headline_steps
    prof = model.ctx.profile()
    stats = model.ctx.stats()
    model.ctx.call("fls_set_profiling", 0)
    kernel_s = sum(v["ms"] for v in prof.values()) / 1e3 / a.headline_steps
    mine = dict(rank=rank, kernel_s_per_pass=round(kernel_s, 3), busy_frac=round(kernel_s / dt_step, 4),
                p2p_bytes_per_pass=int(stats.p2p_bytes / a.headline_steps),
                weight_bytes_per_pass=int(stats.h2d_bytes / a.headline_steps),
                weight_stream_gbs=round(stats.h2d_bytes / (stats.h2d_ms * 1e-3) / 1e9, 2) if stats.h2d_ms > 0 else None,
                ring_depth=model.last_timing.get("ring_depth"), ring_bytes=model.last_timing.get("ring_bytes"),
                host_bytes_page_locked=int(stats.pinned_bytes),
                launches_per_pass=int(stats.kernel_launches / a.headline_steps), scored=len(out))
    per_rank = [None] * world
    dist.all_gather_object(per_rank, mine)
    value = n_prompts / dt_step
    blk = None
    if rank == 0:
        assert sum(r["scored"] for r in per_rank) == n_prompts
        blk = {
            "workload": f"Llama-2-70b arch random-init fp16, layer_num_per_shard=1, storage_location=gpu, {n_prompts} "
                        f"prompts x {rows_per_prompt} token rows, ONE job over {world} GPUs",
            "parallelism": f"mp{world}: layer pipeline, shard j on rank j mod {world}, one process per GPU, rows handed "
                           "rank to rank over NVLink through IPC-mapped rings with device-side ready/free counters",
            "scaling": "strong", "metric": METRIC, "value": value, "unit": "prompts/s",
            "prefill_tok_per_s": value * rows_per_prompt, "steps": a.headline_steps, "warmup": 1,
            "ms_per_step": 1e3 * dt_step,
            "timing": "wall clock between barriers around whole passes through main.run_sharded_llama on every rank "
                      "(strings in / numpy probabilities out on the lm_head owner: tokenisation, broadcast of the token "
                      "ids, all ranks' shards, D2H of the scores), max over ranks",
            "e2e": {"value": value, "unit": "prompts/s", "note": "the timed region already is end to end"},
            "path_roofline": path_roofline(shape, n_prompts, n_options, rows_per_prompt, link_gbs, pk, dt_step, stats,
                                           a.headline_steps, n_ranks=world),
            "per_rank": per_rank, "p2p_bytes_per_pass": sum(r["p2p_bytes_per_pass"] for r in per_rank),
            "clocks_rank0": clocks.summary(),
            "tiny_model_pipeline_bit_identical_to_1gpu": identical,
            "distinct_decoder_layers": alias or shape.num_hidden_layers,
            "setup_s": {"write_layer_files": round(t_files, 1)},
        }
    for m in dm.models.values():
        m.close()
    dm.models.clear()
    release_host_images()
    dist.barrier()
    return blk


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
    ap.add_argument("--eager_sample_prompts", type=int, default=64)
    ap.add_argument("--no_cpu_baseline", action="store_true")
    ap.add_argument("--no_eager_baseline", action="store_true")
    ap.add_argument("--no_e2e", action="store_true")
    ap.add_argument("--no_headline", action="store_true", help="skip the 70B headline block (N=1) / mp block (N>1)")
    ap.add_argument("--headline_only", action="store_true", help="the main workload gets 1+1 passes and no baselines")
    ap.add_argument("--headline_steps", type=int, default=2)
    ap.add_argument("--headline_prompts", type=int, default=0, help="override the 70B prompt count (debugging)")
    ap.add_argument("--alias_layers", type=int, default=-1,
                    help="70B: write only N distinct decoder layers and symlink the rest (default: all distinct "
                         "when the host's RAM holds the model, else as many as fit)")
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
        alias = choose_alias(shape, a.alias_layers) if shape_name == "70b" else max(a.alias_layers, 0)
        md = model_dir_for(shape_name, 0, 1, gen_dev, alias)
        prompts = synthetic.make_prompts(max(a.cpu_sample_prompts, 1), shape.vocab_size, prefix_tokens, n_options,
                                         option_tokens, seed=1234)
        vals, detail = [], {}
        for _ in range(max(1, min(a.warmup, 1)) + max(1, min(a.steps, 3))):
            v, d, _ = cpu_oracle_sample(md, prompts, shape.num_hidden_layers, repeats=1)
            if v is not None:
                vals.append(v)
                detail = d
        v = float(np.median(vals[1:] if len(vals) > 1 else vals)) if vals else 0.0
        sample = (f"{a.cpu_sample_prompts} prompts x {rows_per_prompt} rows through 1 and 2 decoder layers + embed/norm/"
                  f"lm_head, fp32, extrapolated linearly to {shape.num_hidden_layers} layers; median of the "
                  f"{max(1, len(vals) - 1)} samples after the first")
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v,
            "unit": "prompts/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * n_prompts / v if v else None, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (CPU oracle of the fp16 reference loop)", "data": "synthetic",
            "config": {"workload": workload_name},
            "cpu_baseline": {"value": v, "unit": "prompts/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": sample, **detail},
            "e2e": {"value": v, "unit": "prompts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "prefill_tok_per_s": v * rows_per_prompt,
        }), flush=True)
        return

    # ------------------------------------------------------------------ B200 arm
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    device = torch.device(f"cuda:{local}")
    if world > 1:
        # NCCL_DEBUG / NCCL_DEBUG_FILE are left exactly as the caller set them (the driver reads NCCL's own lines to
        # count ranks); whatever NCCL prints comes first, the JSON line is the LAST line of stdout
        torch.distributed.init_process_group("nccl", device_id=device)
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if world > 1:
        torch.distributed.barrier()
    from flexible_llm_sharding_b200 import planner
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama

    alias = choose_alias(shape, a.alias_layers) if shape_name == "70b" else max(a.alias_layers, 0)
    md = model_dir_for(shape_name, rank, world, device, alias)
    prompts = synthetic.make_prompts(n_prompts, shape.vocab_size, prefix_tokens, n_options, option_tokens,
                                     seed=1234 + rank)
    args = make_args(md, layer_num_per_shard=a.layer_num_per_shard, storage_location=a.storage_location,
                     max_rows=a.max_rows, pinned_limit_bytes=0)
    model = ShardedLlama(args, DeviceManager(args, [str(device)]), device=str(device))
    tok = planner.tokenize_prompts(model.tokenizer, prompts)
    packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, model.prefix_causal)
    link_gbs = measure_h2d_gbs(device)

    steps, warmup = (1, 1) if a.headline_only else (a.steps, a.warmup)
    dt, prof, stats, clocks = timed_device_passes(model, packed, steps, warmup, world, device, local)
    launches = int(stats.kernel_launches)
    value = world * n_prompts * steps / dt

    e2e = None
    if not a.no_e2e and not a.headline_only:
        et, out, extra = timed_e2e(model, prompts, steps, world, device)
        e2e = {"value": world * n_prompts * steps / et, "unit": "prompts/s", **extra}
        assert len(out) == n_prompts and out[0].shape == (n_options, 1, shape.vocab_size)
    pinned_main, mapped_main = int(stats.pinned_bytes), int(stats.mapped_bytes)

    # bounded CPU sample + parity of the CUDA path against it + the torch-eager GPU arm (1-GPU runs only)
    cpu = parity = eager = None
    if world == 1 and not a.headline_only:
        sample = prompts[:a.cpu_sample_prompts]
        if not a.no_cpu_baseline:
            v, detail, ref_scores = cpu_oracle_sample(md, sample, shape.num_hidden_layers)
            cpu = {"value": v, "unit": "prompts/s", "cores": os.cpu_count(), "kind": "port",
                   "sample": (f"{len(sample)} prompts x {rows_per_prompt} rows through 1 and 2 decoder layers + embed/"
                              f"norm/lm_head, fp32 oracle, median of 3 runs each, extrapolated linearly to "
                              f"{shape.num_hidden_layers} layers"), **detail}
            view = truncated_model_view(md, 2)
            try:
                vargs = make_args(view, max_rows=a.max_rows)
                vm = ShardedLlama(vargs, DeviceManager(vargs, [str(device)]), device=str(device))
                parity = parity_against(ref_scores, vm(copy.deepcopy(sample)))
                parity["sample"] = (f"the CPU sample's {len(sample)} prompts through embed + 2 decoder layers + norm + "
                                    f"lm_head at {shape_name} shapes: CUDA path vs the fp32 oracle")
                vm.ctx.close()
                del vm
            finally:
                shutil.rmtree(view, ignore_errors=True)
        if not a.no_eager_baseline:
            eager = gpu_eager_baseline(md, prompts[:a.eager_sample_prompts], n_prompts, device)
    model.ctx.close()
    del model
    release_host_images()

    headline = mp = c3 = None
    if not a.no_headline:
        if world == 1:
            headline = headline_block(a, pk, device, local, link_gbs)
        else:
            c3 = c3_block(a, pk, device, local, rank, world, link_gbs)
            mp = mp_block(a, pk, device, local, rank, world, link_gbs)

    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    line = {
        "metric": METRIC, "value": value, "unit": "prompts/s",
        "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * dt / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 (fp32 accumulate)",
        "data": "synthetic",
        "config": {"workload": workload_name + (f" [{alias} distinct decoder layers, others aliased]" if alias else ""),
                   "micro_batch_rows": a.max_rows,
                   "cache": "inputs larger than L2 (activations %.1f GB, weights %.1f GB per step)" % (
                       n_prompts * rows_per_prompt * shape.hidden_size * 2 / 1e9, stats.h2d_bytes / steps / 1e9),
                   "parallelism": f"dp{world} (independent prompt partitions, no collective)",
                   "host_bytes_page_locked": pinned_main, "of_which_file_pages_locked_in_place": mapped_main},
        "prefill_tok_per_s": value * rows_per_prompt,
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        "roofline": gemm_roofline(prof, dt, steps, pk, shape_name, a.max_rows),
        "path_roofline": path_roofline(shape, n_prompts, n_options, rows_per_prompt, link_gbs, pk, dt / steps, stats,
                                       steps),
        "cpu_baseline": cpu, "parity_check": parity, "gpu_eager_baseline": eager,
    }
    if headline is not None:
        line["headline"] = headline
    if c3 is not None:
        line["dp_13b_k4"] = c3
    if mp is not None:
        line["mp"] = mp
    print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
