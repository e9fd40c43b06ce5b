#!/usr/bin/env python3
"""bench.py — the hot path of quorum on B200: one JSON line per run.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg4] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic input (quorum_b200/synth.py, fixed seeds):
  cfg2 (default, BASELINE.json configs[1], the configuration the metric is quoted on): streaming
       hide_intermediate_think, 65,536 concurrent streams x 256 chunks x 64 B per GPU, tags split across chunk cuts;
  cfg3: 1 M responses x 4 KB, 5 thinking_tags, 4 per request, multi-byte separator (non-stream concatenate);
  cfg4: 8 backends x 32 KB reasoning-heavy responses, hide_final_think, 16,384 requests.
`value` is GB/s of INPUT bytes with the batch already resident in HBM; `e2e` is the same metric through the public
host-buffer call (pinned host -> device copies, kernels, device -> host copies inside the timed region).
Multi-GPU: units (streams / requests) are independent, every rank runs its own batch, no collective on the data
path (weak scaling); torch.distributed is used for the barrier and the max-over-ranks time only.
`--impl reference` times the reference's own CPU code (oracle/_ref = unmodified copy of the reference module when
staged, else the C port oracle/quorum_oracle.c) on all host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "cfg2": dict(kind="stream", units=65536, chunks=256, chunk_bytes=64, n_base=512,
                 name="cfg2: streaming hide_intermediate_think, 65536 streams x 256 chunks x 64 B, tags split across chunk cuts"),
    "cfg3": dict(kind="resp", units=1000000, unit_bytes=4096, per_request=4, n_base=4096,
                 name="cfg3: 1M responses x 4 KB, 5 thinking_tags, 4 per request, multi-byte separator"),
    "cfg4": dict(kind="resp", units=131072, unit_bytes=32768, per_request=8, n_base=256,
                 name="cfg4: 8 backends x 32 KB reasoning-heavy (70% inside <think>), hide_final_think, 16384 requests"),
    "cfg5": dict(kind="ragged", total_bytes=1250000000, pool_bytes=24 << 20, per_request=4,
                 name="cfg5: adversarial mix (nested/unclosed/stray/crossed tags, fold characters, multi-byte white space), "
                      "response lengths log-uniform 64 B - 64 KB, 1.25 GB per GPU (10 GB over 8), 4 per request"),
}
TAGS4 = ["think", "reason", "reasoning", "thought"]


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed regions (device and end to end) run."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# workload construction
# ---------------------------------------------------------------------------------------------------------------

class RaggedBase:
    """Pool of ragged units (list of bytes) with the bits of the ndarray interface the CPU legs use."""

    def __init__(self, units):
        self.units = units
        self.shape = (len(units), int(np.mean([len(u) for u in units])))

    def __getitem__(self, idx):
        return [self.units[int(i) % len(self.units)] for i in idx]


def build_base(wl, name):
    from quorum_b200 import synth
    if name == "cfg5":
        return RaggedBase(synth.cfg5_units(wl["pool_bytes"], seed=1005))
    if name == "cfg2":
        return synth.cfg2_base(wl["n_base"], wl["chunks"], wl["chunk_bytes"])
    if name == "cfg3":
        return synth.cfg3_base(wl["n_base"], wl["unit_bytes"])
    return synth.cfg4_base(wl["n_base"], wl["unit_bytes"])


def unit_index(wl, n_units, rank):
    from quorum_b200 import synth
    if wl["kind"] == "ragged":
        return np.arange(n_units)
    return synth.gather_units(np.empty((wl["n_base"], 1)), n_units, seed=1000 + rank)


def _raw(r):
    return r if isinstance(r, (bytes, bytearray)) else r.tobytes()


# ---------------------------------------------------------------------------------------------------------------
# CPU legs (the only place bench.py touches oracle/)
# ---------------------------------------------------------------------------------------------------------------

def _ref_worker_init():
    global _REFMOD
    import logging
    logging.disable(logging.CRITICAL)
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "src"))
    import quorum.oai_proxy as q
    _REFMOD = q


def _ref_stream_shard(args):
    rows, chunk_bytes, tags = args
    q = _REFMOD
    t0 = time.perf_counter()
    nbytes = 0
    for row in rows:
        text = row.tobytes()
        f = q.ThinkingTagFilter(tags)
        for p in range(0, len(text), chunk_bytes):
            f.feed(text[p:p + chunk_bytes].decode())
        f.flush()
        nbytes += len(text)
    return nbytes, time.perf_counter() - t0


def _ref_resp_shard(args):
    rows, per_request, tags, sep = args
    q = _REFMOD
    t0 = time.perf_counter()
    nbytes = 0
    texts = [_raw(r).decode() for r in rows]
    t0 = time.perf_counter()
    for i in range(0, len(texts), per_request):
        processed = [q.strip_thinking_tags(c, tags, hide_intermediate=True) for c in texts[i:i + per_request]]
        sep.join(processed)
    for r in rows:
        nbytes += len(_raw(r))
    return nbytes, time.perf_counter() - t0


def cpu_reference_available() -> bool:
    return os.path.isfile(os.path.join(ROOT, "oracle", "_ref", "src", "quorum", "oai_proxy.py"))


def time_cpu(wl, name, base, cores, sample_units, repeats=1):
    """Returns (GB/s, kind, sample description).  Whole sample wall clock over `cores` worker processes/threads."""
    from quorum_b200 import synth
    idx = unit_index(wl, sample_units, rank=0)
    rows = base[idx]
    shards = [rows[i::cores] for i in range(cores)]
    tags = synth.TAGS5
    if cpu_reference_available():
        import multiprocessing as mp
        ctx = mp.get_context("fork")
        with ctx.Pool(cores, initializer=_ref_worker_init) as pool:
            pool.map(_ref_stream_shard if wl["kind"] == "stream" else _ref_resp_shard,
                     [(s[:2], wl.get("chunk_bytes"), tags) if wl["kind"] == "stream" else
                      (s[:wl["per_request"]], wl["per_request"], tags, synth.SEP_MULTIBYTE) for s in shards])  # warm-up
            best = None
            for _ in range(repeats):
                t0 = time.perf_counter()
                if wl["kind"] == "stream":
                    res = pool.map(_ref_stream_shard, [(s, wl["chunk_bytes"], tags) for s in shards])
                else:
                    res = pool.map(_ref_resp_shard, [(s, wl["per_request"], tags, synth.SEP_MULTIBYTE) for s in shards])
                dt = time.perf_counter() - t0
                nbytes = sum(r[0] for r in res)
                best = dt if best is None else min(best, dt)
        kind = "reference"
        how = f"unmodified reference module (CPython {sys.version.split()[0]}), multiprocessing.Pool({cores})"
    else:
        from concurrent.futures import ThreadPoolExecutor
        from oracle import c_oracle
        ctags = c_oracle.Tags(tags)

        def run(shard):
            if wl["kind"] == "ragged":
                flat, io = synth.pack([_raw(u) for u in shard])
                ro = np.arange(0, len(shard) + 1, wl["per_request"], dtype=np.int64)
                if ro[-1] != len(shard):
                    ro = np.append(ro, len(shard))
                c_oracle.strip_concat_batch(ctags, flat, io, ro, synth.SEP_MULTIBYTE.encode(), c_oracle.FLAG_HIDE)
                return flat.size
            flat = np.ascontiguousarray(shard).reshape(-1)
            if wl["kind"] == "stream":
                co = np.arange(0, flat.size + 1, wl["chunk_bytes"], dtype=np.int64)
                so = np.arange(0, shard.shape[0] * wl["chunks"] + 1, wl["chunks"], dtype=np.int64)
                c_oracle.stream_feed_batch(ctags, flat, co, so, None, c_oracle.FLAG_FLUSH)
            else:
                io = np.arange(0, flat.size + 1, wl["unit_bytes"], dtype=np.int64)
                ro = np.arange(0, shard.shape[0] + 1, wl["per_request"], dtype=np.int64)
                if ro[-1] != shard.shape[0]:
                    ro = np.append(ro, shard.shape[0])
                c_oracle.strip_concat_batch(ctags, flat, io, ro, synth.SEP_MULTIBYTE.encode(), c_oracle.FLAG_HIDE)
            return flat.size

        with ThreadPoolExecutor(cores) as ex:
            list(ex.map(run, [s[:2] for s in shards]))
            best = None
            for _ in range(repeats):
                t0 = time.perf_counter()
                nbytes = sum(ex.map(run, shards))
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
        kind = "port"
        how = f"C restatement oracle/quorum_oracle.c, {cores} threads (ctypes releases the GIL)"
    unit_b = base.shape[1]
    return nbytes / best / 1e9, kind, f"{sample_units} units x ~{unit_b} B of the same workload; {how}"


def run_reference_arm(args, wl, name):
    """--impl reference: the reference's CPU implementation on all host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    base = build_base(wl, name)
    per_core = 96 if wl["kind"] == "stream" else (2048 if wl.get("unit_bytes", 9000) <= 4096 else 256)
    if wl["kind"] == "ragged":
        per_core = min(per_core, max(4, len(base.units) // cores))
    if not cpu_reference_available():
        per_core *= 16
    sample_units = per_core * cores
    vals = []
    for i in range(args.warmup + args.steps):
        v, kind, sample = time_cpu(wl, name, base, cores, sample_units)
        if i >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals))
    unit_b = base.shape[1]
    line = {
        "impl": "reference", "metric": "filtered GB/s (input bytes)", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sample_units * unit_b / (value * 1e9) * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic (quorum_b200/synth.py, fixed seeds)",
        "config": {"workload": wl["name"], "sample": sample},
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))  # cfg2 = BASELINE.json configs[1]
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--units", type=int, default=0, help="override the number of units per GPU (debug)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    name = args.workload
    wl = dict(WORKLOADS[name])
    if args.units:
        wl["units"] = args.units

    if args.impl == "reference":
        run_reference_arm(args, wl, name)
        return

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: quorum_b200 has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    import quorum_b200
    from quorum_b200 import batch as qbatch
    from quorum_b200 import synth

    peak, peak_src = read_peaks()
    base = build_base(wl, name)
    unit_b = base.shape[1]
    ragged_off = None
    if wl["kind"] == "ragged":
        # the pool of adversarial units, rotated by rank and repeated until the per-GPU byte count is reached
        units = base.units[rank % len(base.units):] + base.units[:rank % len(base.units)]
        pool, poff = synth.pack(units)
        reps = max(1, int(wl["total_bytes"] // max(1, pool.size)))
        if args.units:
            reps = 1
        data = torch.from_numpy(pool.copy()).to(dev).repeat(reps).contiguous()
        lens = torch.from_numpy(np.diff(poff)).to(dev).repeat(reps)
        ragged_off = torch.zeros(lens.numel() + 1, dtype=torch.int64, device=dev)
        torch.cumsum(lens, 0, out=ragged_off[1:])
        n_units = int(lens.numel())
    else:
        n_units = wl["units"]
        idx = unit_index(wl, n_units, rank)
        base_dev = torch.from_numpy(base).to(dev)
        idx_dev = torch.from_numpy(idx).to(dev)
        data = base_dev.index_select(0, idx_dev).reshape(-1).contiguous()      # uint8[n_units * unit_b], device resident
        del base_dev
    total_in = data.numel()
    table = qbatch.tag_table(synth.TAGS5)

    if wl["kind"] == "stream":
        cb, ch = wl["chunk_bytes"], wl["chunks"]
        chunk_off = torch.arange(0, total_in + 1, cb, dtype=torch.int64, device=dev)
        stream_off = torch.arange(0, n_units * ch + 1, ch, dtype=torch.int64, device=dev)
        n_items = n_units * ch
        out = torch.empty(total_in + 32 * n_units + 16, dtype=torch.uint8, device=dev)
        oco = torch.empty(n_items + 1, dtype=torch.int64, device=dev)
        state_out = torch.empty((n_units, 32), dtype=torch.uint8, device=dev)
        scratch = torch.empty(qbatch.lib().qb200_stream_feed_scratch_bytes(n_items, n_units, out.numel()), dtype=torch.uint8, device=dev)

        def step():
            qbatch.stream_feed_batch(table, data, chunk_off, stream_off, None, flush=True, state_out=state_out,
                                     out_bytes=out, out_chunk_off=oco, scratch=scratch)

        def out_total():
            return int(oco[-1].item())
        meta_bytes = 16 * n_items + 32 * n_units     # chunk_off read + out_chunk_off write per chunk, state write per stream
        items_name = "chunks"
    else:
        pr = wl["per_request"]
        if ragged_off is not None:
            n_units -= n_units % pr
            in_off = ragged_off[: n_units + 1].contiguous()
            total_in = int(in_off[-1].item())
            data = data[:total_in]
        else:
            in_off = torch.arange(0, total_in + 1, unit_b, dtype=torch.int64, device=dev)
        n_req = n_units // pr
        req_off = torch.arange(0, n_units + 1, pr, dtype=torch.int64, device=dev)
        sep_b = synth.SEP_MULTIBYTE.encode()
        sep = torch.frombuffer(bytearray(sep_b), dtype=torch.uint8).to(dev)
        n_items = n_units
        out = torch.empty(total_in + len(sep_b) * n_units + 16, dtype=torch.uint8, device=dev)
        out_off = torch.empty(n_req + 1, dtype=torch.int64, device=dev)
        scratch = torch.empty(qbatch.lib().qb200_strip_concat_scratch_bytes(n_units, n_req, out.numel()), dtype=torch.uint8, device=dev)

        def step():
            qbatch.strip_concat_batch(table, data, in_off, req_off, sep, hide=True, streaming=False, out_bytes=out,
                                      out_off=out_off, scratch=scratch)

        def out_total():
            return int(out_off[-1].item())
        meta_bytes = 16 * n_units + 8 * n_req
        items_name = "responses"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value) ----
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    total_out = out_total()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    l0 = quorum_b200.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    launches = quorum_b200.launch_count() - l0
    barrier()
    ms = ev0.elapsed_time(ev1) / args.steps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt)
        launches = int(lt.item())
    value = world * total_in / (ms * 1e-3) / 1e9

    # ---- per-kernel split of a call (outside the timed region): the library records an event after each launch ----
    kernels = None
    if rank == 0:
        from quorum_b200 import _lib as qlib
        ksteps = max(1, min(args.steps, 10))
        qlib.profile_begin(torch.cuda.current_stream().cuda_stream)
        for _ in range(ksteps):
            step()
        marks = qlib.profile_end()
        agg = {}
        for kname, kms in marks:
            a = agg.setdefault(kname, [0.0, 0])
            a[0] += kms
            a[1] += 1
        tot_ms = sum(a[0] for a in agg.values()) or 1.0
        kernels = [{"name": k, "ms_per_step": a[0] / ksteps, "launches_per_step": a[1] / ksteps, "share": a[0] / tot_ms}
                   for k, a in agg.items()]
    barrier()

    # ---- end to end through the host-buffer call ----
    e2e = None
    if not args.no_e2e:
        h_data = torch.empty(total_in, dtype=torch.uint8).pin_memory()
        h_data.copy_(data)
        if wl["kind"] == "stream":
            h_co, h_so = chunk_off.cpu().pin_memory(), stream_off.cpu().pin_memory()
            sess = qbatch.HostStreamFeeder(table, dev, total_in, n_items, n_units)

            def e2e_step():
                return sess.run(h_data, h_co, h_so, flush=True)
        else:
            h_io, h_ro = in_off.cpu().pin_memory(), req_off.cpu().pin_memory()
            sess = qbatch.HostStripConcat(table, dev, total_in, n_units, n_req, sep_b)

            def e2e_step():
                return sess.run(h_data, h_io, h_ro, hide=True, streaming=False)
        e2e_steps = max(3, min(args.steps, 10))
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            h2d, d2h = e2e_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": world * total_in / dt / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": dt * 1e3}
        del h_data, sess
    clocks = sampler.stop()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    alg_bytes = total_in + total_out + meta_bytes
    achieved = alg_bytes / (ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(alg_bytes),
                "kernel": "whole call (all kernels of one qb200_*_batch call), CUDA events on the launching stream"}
    if kernels:
        # the two kernels that touch the text: what each must move, against the time the library's own events give it
        moved = {"tokenize_kernel": total_in, "stream_finish_kernel": 2 * total_out, "strip_copy_kernel": 2 * total_out}
        for k in kernels:
            if k["name"] in moved and k["ms_per_step"] > 0:
                k["bytes"] = int(moved[k["name"]])
                k["GBps"] = moved[k["name"]] / (k["ms_per_step"] * 1e-3) / 1e9
                k["frac_of_peak"] = k["GBps"] / peak
        roofline["kernels"] = kernels
        roofline["kernels_source"] = "qb200_profile_begin/end: CUDA events after each launch on the launching stream, separate pass after the timed region"
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(prof):
        try:
            with open(prof) as f:
                tj = json.load(f)
            if name in tj:
                roofline["traffic"] = tj[name].get("dram_bytes_per_launch")
                roofline["traffic_source"] = tj[name].get("source")
        except Exception:
            pass

    cpu = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        per_core = 64 if wl["kind"] == "stream" else (1024 if unit_b <= 4096 else 128)
        if wl["kind"] == "ragged":
            per_core = min(per_core, max(4, len(base.units) // cores))
        if not cpu_reference_available():
            per_core *= 16
        v, kind, sample = time_cpu(wl, name, base, cores, per_core * cores)
        cpu = {"value": v, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample}

    line = {
        "metric": "filtered GB/s (input bytes)", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic (quorum_b200/synth.py, fixed seeds)",
        "config": {"workload": wl["name"], "units_per_gpu": n_units, "unit_bytes": unit_b,
                   "input_bytes_per_gpu": int(total_in), "output_bytes_per_gpu": total_out,
                   "thinking_tags": synth.TAGS5, "l2": "inputs larger than L2 (no flush needed)",
                   "sharding": f"{world} rank(s), independent units, no collective"},
        items_name + "_per_s": world * n_items / (ms * 1e-3),
        "units_per_s": world * n_units / (ms * 1e-3),
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
