"""Activation store: where the packed token rows live between shards.

Replaces `store_intermediate_activation` / `fetch_intermediate_activation` (reference
utils.py:159-213).  The reference keeps one `(prefix, suffix)` pair per prompt in a dict and
round-trips it through `.cpu()` / `.to(device)` / `np.save` once per prompt per shard.  Here the
rows of all prompts form one `[total_rows, H]` fp16 matrix and a tier decides where it lives:

  gpu   one HBM tensor; fetch/store are views (no copies).  With 180 GB per B200 this holds
        e.g. 1024 x 2048-token prompts of Llama-2-70B (34 GB).
  cpu   a pinned host tensor; each micro-batch is copied in on an H2D stream while the previous
        one computes and copied back on a D2H stream (three rotating HBM staging buffers).
  disk  per-prompt `.npy` files with the reference's names and shapes
        (`prefix{rank}-{idx:05d}.npy` [1,len_p,H], `suffix…` [S,len_s,H]; utils.py:171-177).
        Files are written and read by a small thread pool straight from / into pinned host buffers
        (three for each direction), a micro-batch ahead of the GPU, so the file system works while
        the GPU computes.

`storage_location` names the tier the user allows; unless `force_tier` is set the store upgrades
`cpu`/`disk` to `gpu` when the matrix fits the HBM budget ("HBM-resident where capacity allows").
The scored rows after `model.norm` (S per prompt) are small and always stay in HBM.
"""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

DISK_IO_THREADS = 8      # per-prompt file reads / writes in flight


class ActivationStore:
    def __init__(self, location: str, packed, hidden: int, device: torch.device, compute_stream,
                 disk_folder: str = "./temp", rank_tag: str = "", hbm_budget_bytes: int | None = None,
                 force_tier: bool = False, max_rows: int = 0, pinned_pool: dict | None = None):
        assert location in ("gpu", "cpu", "disk"), location
        self.packed = packed
        self.H = hidden
        self.device = device
        self.compute_stream = compute_stream
        self.disk_folder = disk_folder
        self.rank_tag = rank_tag
        total_bytes = packed.total_rows * hidden * 2
        tier = location
        if tier != "gpu" and not force_tier:
            if hbm_budget_bytes is None:
                free, _ = torch.cuda.mem_get_info(device)     # plus what torch's allocator holds but is not using
                free += torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
                hbm_budget_bytes = int(free * 0.8)
            if total_bytes <= hbm_budget_bytes:
                tier = "gpu"
        self.tier = tier
        self.requested = location
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.heads = torch.empty((packed.total_head_rows, hidden), dtype=torch.float16, device=device)
        if tier == "gpu":
            self.x_all = torch.empty((packed.total_rows, hidden), dtype=torch.float16, device=device)
        else:
            cap = max(mb.n_rows for mb in packed.micro_batches)
            self.staging = [torch.empty((cap, hidden), dtype=torch.float16, device=device) for _ in range(3)]
            # high priority: these copies gate compute, the (low-priority) weight stream does not
            self.h2d_stream = torch.cuda.Stream(device, priority=-1)
            self.d2h_stream = torch.cuda.Stream(device, priority=-1)
            self.h2d_done = [None] * 3
            self.d2h_done = [None] * 3
            self.loaded = {}     # micro-batch index -> staging slot holding it
            self.seq = 0         # staging slots rotate over ALL fetches, so a shard's first micro-batch can be
            #                      brought in while the previous shard's last one still computes
            self.saved = {}      # micro-batch index -> event: its rows' latest version is in host memory
            if tier == "cpu":
                self.x_host = self._pinned(pinned_pool, "x_host", packed.total_rows, hidden)
            else:
                os.makedirs(disk_folder, exist_ok=True)
                self.host_w = [self._pinned(pinned_pool, f"disk_w{j}", cap, hidden) for j in range(3)]
                self.host_r = [self._pinned(pinned_pool, f"disk_r{j}", cap, hidden) for j in range(3)]
                self.jobs = ThreadPoolExecutor(max_workers=4, thread_name_prefix="fls-disk-job")
                self.io = ThreadPoolExecutor(max_workers=DISK_IO_THREADS, thread_name_prefix="fls-disk-io")
                self.w_job = [None] * 3   # writer job that last used host_w[j]
                self.r_job = {}           # micro-batch index -> reader job filling its staging slot
                self.written = {}         # micro-batch index -> writer job of its latest files

    @staticmethod
    def _pinned(pool, name, rows, hidden):
        """Page-locked [rows, hidden] fp16 buffer; page-locking a GB costs tenths of a second, so the owner
        (ShardedLlama) keeps the largest one per name and hands it back on the next call."""
        need = rows * hidden
        if pool is not None and name in pool and pool[name].numel() >= need:
            return pool[name][:need].view(rows, hidden)
        buf = torch.empty(need, dtype=torch.float16).pin_memory()
        if pool is not None:
            pool[name] = buf
        return buf.view(rows, hidden)

    # ------------------------------------------------------------------ token-stage rows
    def _take_slot(self) -> int:
        s = self.seq % 3
        self.seq += 1
        return s

    def prefetch(self, i: int, has_input: bool):
        """Start bringing micro-batch i into HBM (no-op for the gpu tier).  `i == len(micro_batches)` names the
        first micro-batch of the NEXT shard: the caller passes it only when another token-stage shard follows."""
        n = len(self.packed.micro_batches)
        if self.tier == "gpu" or i > n:
            return
        if i == n:
            i = 0
            if not has_input or n < 2 or i not in self.saved:
                return                                     # its rows are not in host memory yet
        if i in self.loaded:
            return
        mb = self.packed.micro_batches[i]
        s = self._take_slot()
        buf = self.staging[s][:mb.n_rows]
        if has_input and self.tier == "disk":
            ev = torch.cuda.Event()
            self.r_job[i] = self.jobs.submit(self._disk_load, mb, s, self.d2h_done[s], self.written.get(i),
                                             self.h2d_done[s], ev)
            self.h2d_done[s] = ev                          # recorded by the job; fetch() waits for the job first
            self.h2d_bytes += mb.n_rows * self.H * 2
        elif has_input:
            with torch.cuda.stream(self.h2d_stream):
                if self.d2h_done[s] is not None:           # previous tenant fully written back
                    self.h2d_stream.wait_event(self.d2h_done[s])
                if i in self.saved:                        # and these rows' own write-back has landed
                    self.h2d_stream.wait_event(self.saved[i])
                buf.copy_(self.x_host[mb.row_lo:mb.row_lo + mb.n_rows], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.h2d_stream)
                self.h2d_done[s] = ev
            self.h2d_bytes += mb.n_rows * self.H * 2
        else:
            if self.d2h_done[s] is not None:
                self.compute_stream.wait_event(self.d2h_done[s])
            self.h2d_done[s] = None
        self.loaded[i] = s

    def fetch(self, i: int, has_input: bool) -> torch.Tensor:
        """Device tensor [n_rows, H] holding micro-batch i's rows (uninitialised if not has_input)."""
        mb = self.packed.micro_batches[i]
        if self.tier == "gpu":
            return self.x_all[mb.row_lo:mb.row_lo + mb.n_rows]
        self.prefetch(i, has_input)
        s = self.loaded[i]
        if self.tier == "disk" and i in self.r_job:
            self.r_job.pop(i).result()                     # files read and the H2D enqueued (event recorded)
        if self.h2d_done[s] is not None:
            self.compute_stream.wait_event(self.h2d_done[s])
        return self.staging[s][:mb.n_rows]

    def store(self, i: int, x: torch.Tensor, keep_tokens: bool):
        """Persist micro-batch i after a shard (`keep_tokens` False once model.norm has run)."""
        if self.tier == "gpu":
            return
        s = self.loaded.pop(i)
        mb = self.packed.micro_batches[i]
        ev = torch.cuda.Event()
        ev.record(self.compute_stream)
        if keep_tokens and self.tier == "disk" and self.w_job[s] is not None:
            self.w_job[s].result()                         # host_w[s] is free again (bounds the host's run-ahead)
        with torch.cuda.stream(self.d2h_stream):
            self.d2h_stream.wait_event(ev)
            if keep_tokens:
                if self.tier == "cpu":
                    self.x_host[mb.row_lo:mb.row_lo + mb.n_rows].copy_(x, non_blocking=True)
                else:
                    self.host_w[s][:mb.n_rows].copy_(x, non_blocking=True)
                self.d2h_bytes += mb.n_rows * self.H * 2
            done = torch.cuda.Event()
            done.record(self.d2h_stream)
            self.d2h_done[s] = done
            self.saved[i] = done
        if keep_tokens and self.tier == "disk":
            self.w_job[s] = self.written[i] = self.jobs.submit(self._disk_save, mb, s, done)

    def end_shard(self):
        """Shard boundary.  Whatever is still in `loaded` was prefetched for the next shard and stays."""

    # ------------------------------------------------------------------ scored rows
    def heads_view(self, i: int) -> torch.Tensor:
        mb = self.packed.micro_batches[i]
        return self.heads[mb.head_lo:mb.head_lo + mb.n_head]

    # ------------------------------------------------------------------ disk tier
    def _names(self, prompt_idx: int):
        f = self.disk_folder
        return (f"{f}/prefix{self.rank_tag}-{prompt_idx:05d}.npy", f"{f}/suffix{self.rank_tag}-{prompt_idx:05d}.npy")

    def _prompt_slices(self, mb):
        """(prefix file, suffix file, first row in the micro-batch, len_p, S, len_s) per prompt."""
        for p in range(mb.prompt_lo, mb.prompt_hi):
            len_p, S, len_s = self.packed.shapes[p]
            pf, sf = self._names(self.packed.prompt_index(p))     # the reference's name: the caller's prompt index
            yield pf, sf, int(self.packed.prompt_row[p]) - mb.row_lo, len_p, S, len_s

    def _disk_save(self, mb, s: int, d2h_done):
        """Writer job: once the rows are in host_w[s], write the reference's two files per prompt."""
        d2h_done.synchronize()
        arr = self.host_w[s][:mb.n_rows].numpy()

        def save_one(item):
            pf, sf, b, len_p, S, len_s = item
            np.save(pf, arr[b:b + len_p].reshape(1, len_p, self.H))
            np.save(sf, arr[b + len_p:b + len_p + S * len_s].reshape(S, len_s, self.H))

        list(self.io.map(save_one, self._prompt_slices(mb)))

    @staticmethod
    def _read_npy_into(path: str, dest: np.ndarray, shape: tuple):
        """Read a `.npy` file's payload straight into `dest` (rows of a pinned buffer), checking its header."""
        with open(path, "rb", buffering=0) as fp:
            major, _ = np.lib.format.read_magic(fp)
            hdr = (np.lib.format.read_array_header_1_0 if major == 1 else np.lib.format.read_array_header_2_0)(fp)
            if tuple(hdr[0]) != tuple(shape) or hdr[1] or hdr[2] != np.float16:
                raise ValueError(f"{path}: expected float16 C-order {shape}, found {hdr}")
            view = memoryview(dest.reshape(-1).view(np.uint8))
            got = 0
            while got < len(view):
                n = fp.readinto(view[got:])
                if not n:
                    raise IOError(f"{path}: truncated ({got} of {len(view)} bytes)")
                got += n

    def _disk_load(self, mb, s: int, slot_written_back, files_written, host_free, h2d_ev):
        """Reader job: files -> host_r[s] -> staging[s] on the H2D stream; records `h2d_ev` when enqueued."""
        torch.cuda.set_device(self.device)
        if files_written is not None:
            files_written.result()
        if host_free is not None:
            host_free.synchronize()                        # the previous copy out of host_r[s] has finished
        host = self.host_r[s][:mb.n_rows]
        arr = host.numpy()

        def load_one(item):
            pf, sf, b, len_p, S, len_s = item
            self._read_npy_into(pf, arr[b:b + len_p], (1, len_p, self.H))
            self._read_npy_into(sf, arr[b + len_p:b + len_p + S * len_s], (S, len_s, self.H))

        list(self.io.map(load_one, self._prompt_slices(mb)))
        with torch.cuda.stream(self.h2d_stream):
            if slot_written_back is not None:              # staging[s]'s previous tenant has left
                self.h2d_stream.wait_event(slot_written_back)
            self.staging[s][:mb.n_rows].copy_(host, non_blocking=True)
            h2d_ev.record(self.h2d_stream)

    keep_disk_files = False      # tests set this to look at the files a pass leaves behind

    def cleanup(self, abort: bool = False):
        """End of the pass: wait for the disk tier's jobs, stop its threads, remove its files.  `abort` (the pass
        failed) drops queued jobs and swallows their errors so the original exception surfaces."""
        if self.tier != "disk" or self.jobs is None:
            return
        pending = list(self.r_job.values()) + list(self.written.values())
        self.r_job.clear()
        self.written.clear()
        try:
            for f in pending:
                if abort:
                    f.cancel()
                    try:
                        f.result()
                    except BaseException:
                        pass
                else:
                    f.result()
        finally:
            self.jobs.shutdown(wait=True, cancel_futures=abort)
            self.io.shutdown(wait=True, cancel_futures=abort)
            self.jobs = self.io = None
            if not self.keep_disk_files:
                for p in range(len(self.packed.shapes)):
                    for fn in self._names(p):
                        if os.path.exists(fn):
                            os.remove(fn)
