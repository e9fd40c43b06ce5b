#!/usr/bin/env python3
"""bench.py — the hot path of quorum on B200: one JSON line per run.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg4|cfg5] [--impl reference]

BASELINE.json's metric is "filtered+concatenated GB/s and responses/sec at 1/2/4/8 B200; % of HBM roofline" and names no
configuration, so the headline is the largest single-GPU configuration that exercises filter AND concatenate:
  cfg3 (headline, BASELINE.json configs[2]): 1 M responses x 4 KB, 5 thinking_tags, 4 per request, multi-byte separator,
       non-stream concatenate (oai_proxy.py:1194-1199 + 1286) through qb200_strip_concat_batch;
and the same line carries a `configs` map with every other workload measured the same way:
  cfg2 (configs[1]): streaming hide_intermediate_think, 65,536 streams x 256 chunks x 64 B, tags split across chunk cuts
       (oai_proxy.py:278-371 through qb200_stream_feed_batch);
  cfg4 (configs[3]): 8 backends x 32 KB reasoning-heavy responses (70 % inside <think>), hide_final_think, 16,384 requests;
  cfg5 (configs[4], one GPU's share): adversarial mix, response lengths log-uniform 64 B - 64 KB, 1.25 GB per GPU;
plus `sharded`: configs[4] as BASELINE.json states it - ONE 10 GB batch partitioned by bytes over the N ranks
(quorum_b200.sharding.partition_by_bytes, strong scaling, no collective on the data path).
A "step" is one pass of the hot path over one batch of synthetic input (quorum_b200/synth.py, fixed seeds).  `value` is
GB/s of INPUT bytes with the batch already resident in HBM, timed over exactly K steps (a burst); `sustained` is the same
call repeated back to back for at least one second.  `e2e` is the same metric through the public host-buffer call
(pinned host -> device copies, kernels, device -> host copies inside the timed region), with the raw copy rate of the same
byte counts beside it (`copy_ceiling`: what the PCIe link allows, whatever the kernels do).
Multi-GPU (--gpus N, one rank per GPU under torchrun): the headline is one replica of the workload per rank (weak
scaling: units are independent, no collective; torch.distributed only for the barrier and the max-over-ranks time);
`sharded` is the strong-scaling number.
`--impl reference` times the reference's own CPU code (oracle/_ref = unmodified copy of the reference module when staged,
else the C port oracle/quorum_oracle.c) on all host cores, on a bounded sample of the same workloads.
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

HEADLINE = "cfg3"
WORKLOADS = {
    "cfg2": dict(kind="stream", units=65536, chunks=256, chunk_bytes=64, n_base=512,
                 name="cfg2: streaming hide_intermediate_think, 65536 streams x 256 chunks x 64 B, tags split across chunk cuts"),
    "cfg3": dict(kind="resp", units=1000000, unit_bytes=4096, per_request=4, n_base=4096,
                 name="cfg3: 1M responses x 4 KB, 5 thinking_tags, 4 per request, multi-byte separator, non-stream concatenate"),
    "cfg4": dict(kind="resp", units=131072, unit_bytes=32768, per_request=8, n_base=256,
                 name="cfg4: 8 backends x 32 KB reasoning-heavy (70% inside <think>), hide_final_think, 16384 requests"),
    "cfg5": dict(kind="ragged", total_bytes=1250000000, pool_bytes=24 << 20, per_request=4,
                 name="cfg5: adversarial mix (nested/unclosed/stray/crossed tags, fold characters, multi-byte white space), "
                      "response lengths log-uniform 64 B - 64 KB, 1.25 GB per GPU (10 GB over 8), 4 per request"),
}
SHARDED_TOTAL = 10_000_000_000
NOMINAL_HBM_GBS = 8000.0


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def git_head():
    try:
        return subprocess.check_output(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], stderr=subprocess.DEVNULL, text=True).strip()
    except Exception:
        return None


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
        busy = [x for x in sm if x > 0]
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
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


_BASES = {}


def build_base(wl, name):
    from quorum_b200 import synth
    if name in _BASES:
        return _BASES[name]
    if name == "cfg5":
        b = RaggedBase(synth.cfg5_units(wl["pool_bytes"], seed=1005))
    elif name == "cfg2":
        b = synth.cfg2_base(wl["n_base"], wl["chunks"], wl["chunk_bytes"])
    elif name == "cfg3":
        b = synth.cfg3_base(wl["n_base"], wl["unit_bytes"])
    else:
        b = synth.cfg4_base(wl["n_base"], wl["unit_bytes"])
    _BASES[name] = b
    return b


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


def time_cpu(wl, name, base, cores, sample_units, repeats=1, force_port=False):
    """Returns (GB/s, kind, sample description).  Whole sample wall clock over `cores` worker processes/threads."""
    from quorum_b200 import synth
    idx = unit_index(wl, sample_units, rank=0)
    rows = base[idx]
    shards = [rows[i::cores] for i in range(cores)]
    tags = synth.TAGS5
    if cpu_reference_available() and not force_port:
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
                wall = time.perf_counter() - t0
                nbytes = sum(r[0] for r in res)
                # the workers run side by side, one shard each: the parallel compute time is the slowest worker's own
                # clock around its loop (shipping the shard to the worker - pickling, pipes - is not the reference's work)
                dt = max(r[1] for r in res)
                best = dt if best is None else min(best, dt)
        kind = "reference"
        how = (f"unmodified reference module (CPython {sys.version.split()[0]}), multiprocessing.Pool({cores}), slowest worker's "
               f"compute time (wall clock incl. shipping the shards: {wall:.2f} s)")
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


def _sample_units(wl, base, cores, scale=1.0):
    per_core = 96 if wl["kind"] == "stream" else (2048 if wl.get("unit_bytes", 9000) <= 4096 else 256)
    if wl["kind"] == "ragged":
        per_core = min(per_core, max(4, len(base.units) // cores))
    if not cpu_reference_available():
        per_core *= 16
    per_core = max(wl.get("per_request", 1), int(per_core * scale))
    per_core -= per_core % wl.get("per_request", 1)
    return max(per_core, wl.get("per_request", 1)) * cores


def cpu_baseline_record(wl, name, base, full: bool):
    """The reference's CPU code on this box's host cores, bounded sample.  full: also one core, and the C port on all cores
    (SURVEY.md section 8 d5)."""
    cores = os.cpu_count() or 1
    v, kind, sample = time_cpu(wl, name, base, cores, _sample_units(wl, base, cores, 2.0 if full else 0.5))
    rec = {"value": v, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample}
    if full:
        v1, k1, s1 = time_cpu(wl, name, base, 1, max(wl.get("per_request", 1), _sample_units(wl, base, 1, 4.0)))
        rec["one_core"] = {"value": v1, "unit": "GB/s", "cores": 1, "kind": k1, "sample": s1}
        try:
            vc, kc, sc = time_cpu(wl, name, base, cores, _sample_units(wl, base, cores, 8.0), force_port=True)
            rec["c_port"] = {"value": vc, "unit": "GB/s", "cores": cores, "kind": kc, "sample": sc}
        except Exception as e:  # the C port is test infrastructure; its absence must not void the line
            rec["c_port"] = {"unavailable": str(e)[:200]}
    return rec


def run_reference_arm(args, names):
    """--impl reference: the reference's CPU implementation on all host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    name = names[0]
    wl = dict(WORKLOADS[name])
    base = build_base(wl, name)
    sample_units = _sample_units(wl, base, cores, 1.0)
    vals = []
    for i in range(args.warmup + args.steps):
        v, kind, sample = time_cpu(wl, name, base, cores, sample_units)
        if i >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals))
    unit_b = base.shape[1]
    configs = {}
    for other in names[1:]:
        wo = dict(WORKLOADS[other])
        bo = build_base(wo, other)
        vo, ko, so = time_cpu(wo, other, bo, cores, _sample_units(wo, bo, cores, 0.5))
        configs[other] = {"workload": wo["name"], "value": vo, "unit": "GB/s", "cpu_baseline": {"value": vo, "unit": "GB/s", "cores": cores, "kind": ko, "sample": so}}
    line = {
        "impl": "reference", "metric": "filtered+concatenated GB/s (input bytes)", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sample_units * unit_b / (value * 1e9) * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic (quorum_b200/synth.py, fixed seeds)",
        "config": {"workload": wl["name"], "sample": sample},
        "responses_per_s" if wl["kind"] != "stream" else "streams_per_s": value * 1e9 / unit_b,
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "configs": configs, "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------

class Ctx:
    pass


def bind_to_gpu_numa(torch, local_rank):
    """Keep this rank's threads (and so its pinned allocations, first touch) on the NUMA node of its GPU."""
    try:
        p = torch.cuda.get_device_properties(local_rank)
        path = f"/sys/bus/pci/devices/{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0/local_cpulist"
        with open(path) as f:
            txt = f.read().strip()
        cpus = set()
        for part in txt.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return txt
    except Exception:
        pass
    return None


def make_workload(ctx, name, wl, units_override=0, shard=None):
    """Device-resident batch + the step closure.  shard = (rank, world) builds this rank's part of the 10 GB cfg5 batch."""
    torch, dev, rank = ctx.torch, ctx.dev, ctx.rank
    from quorum_b200 import batch as qbatch
    from quorum_b200 import sharding, synth
    W = Ctx()
    W.name, W.wl = name, wl
    base = build_base(wl, name)
    W.base = base
    unit_b = base.shape[1]
    table = qbatch.tag_table(synth.TAGS5)
    W.table = table
    ragged_off = None
    if wl["kind"] == "ragged":
        pool, poff = synth.pack(base.units)
        plen = np.diff(poff)
        pr = wl["per_request"]
        if shard is None:
            # the pool of adversarial units, rotated by rank and repeated until the per-GPU byte count is reached
            k = rank % len(base.units)
            units = base.units[k:] + base.units[:k]
            pool, poff = synth.pack(units)
            reps = max(1, int(wl["total_bytes"] // max(1, pool.size)))
            if units_override:
                reps = 1
            data = torch.from_numpy(pool.copy()).to(dev).repeat(reps).contiguous()
            lens = torch.from_numpy(np.diff(poff)).to(dev).repeat(reps)
            W.shard_info = None
        else:
            # ONE batch of SHARDED_TOTAL bytes = the pool repeated; requests of `pr` responses; this rank's byte-balanced range
            r_, world = shard
            reps = int(np.ceil(SHARDED_TOTAL / pool.size))
            n_units_all = len(plen) * reps
            n_units_all -= n_units_all % pr
            lens_all = np.tile(plen, reps)[:n_units_all]
            req_bytes = lens_all.reshape(-1, pr).sum(1)
            parts = sharding.partition_by_bytes(req_bytes, world)
            lo, hi = parts[r_]
            off_all = np.concatenate([[0], np.cumsum(lens_all)])
            b0, b1 = int(off_all[lo * pr]), int(off_all[hi * pr])
            start = b0 % pool.size
            k = (start + (b1 - b0) + pool.size - 1) // pool.size
            data = torch.from_numpy(pool.copy()).to(dev).repeat(max(k, 1))[start:start + (b1 - b0)].clone()   # fresh, 16-byte aligned
            lens = torch.from_numpy(lens_all[lo * pr:hi * pr].copy()).to(dev)
            W.shard_info = {"requests": [int(lo), int(hi)], "of_requests": int(len(req_bytes)), "bytes": b1 - b0,
                            "total_bytes": int(off_all[-1]), "parts_bytes": [int(off_all[h * pr] - off_all[l * pr]) for l, h in parts]}
        ragged_off = torch.zeros(lens.numel() + 1, dtype=torch.int64, device=dev)
        torch.cumsum(lens, 0, out=ragged_off[1:])
        n_units = int(lens.numel())
    else:
        n_units = units_override or wl["units"]
        idx = unit_index(wl, n_units, rank)
        base_dev = torch.from_numpy(base).to(dev)
        idx_dev = torch.from_numpy(idx).to(dev)
        data = base_dev.index_select(0, idx_dev).reshape(-1).contiguous()      # uint8[n_units * unit_b], device resident
        del base_dev
    total_in = data.numel()
    W.unit_b = unit_b

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

        W.out_total = lambda: int(oco[-1].item())
        W.meta_bytes = 16 * n_items + 32 * n_units     # chunk_off read + out_chunk_off write per chunk, state write per stream
        W.items_name, W.n_items = "chunks", n_items
        # host-side form of the batch: int32 length of every chunk, chunk index and byte offset of every stream's start
        # (HostStreamFeeder.run_compact: 4 bytes per chunk over PCIe each way instead of 8)
        W.host = lambda: ((chunk_off[1:] - chunk_off[:-1]).to(torch.int32).cpu().pin_memory(), stream_off.cpu().pin_memory(),
                          chunk_off[stream_off].cpu().pin_memory())
        W.session = lambda: qbatch.HostStreamFeeder(table, dev, total_in, n_items, n_units)
        W.e2e_call = lambda sess, h_data, h_len, h_so, h_sbo: sess.run_compact(h_data, h_len, h_so, h_sbo, flush=True)
        W.keep = (chunk_off, stream_off, out, oco, state_out, scratch)
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
        out = torch.empty(total_in + len(sep_b) * n_units + 16, dtype=torch.uint8, device=dev)
        out_off = torch.empty(n_req + 1, dtype=torch.int64, device=dev)
        scratch = torch.empty(qbatch.lib().qb200_strip_concat_scratch_bytes(n_units, n_req, out.numel()), dtype=torch.uint8, device=dev)

        def step():
            qbatch.strip_concat_batch(table, data, in_off, req_off, sep, hide=True, streaming=False, out_bytes=out,
                                      out_off=out_off, scratch=scratch)

        W.out_total = lambda: int(out_off[-1].item())
        W.meta_bytes = 16 * n_units + 8 * n_req
        W.items_name, W.n_items = "responses", n_units
        W.host = lambda: (in_off.cpu().pin_memory(), req_off.cpu().pin_memory())
        W.session = lambda: qbatch.HostStripConcat(table, dev, total_in, n_units, n_req, sep_b)
        W.e2e_call = lambda sess, h_data, h_a, h_b: sess.run(h_data, h_a, h_b, hide=True, streaming=False)
        W.keep = (in_off, req_off, sep, out, out_off, scratch)
    W.data, W.step, W.total_in, W.n_units = data, step, int(total_in), n_units
    return W


def measure(ctx, W, args, want_e2e=True, want_kernels=True, sustained_s=1.0):
    """One workload: burst of exactly K steps, >= sustained_s of back-to-back calls, per-kernel split, end to end."""
    torch, dist, dev, world = ctx.torch, ctx.dist, ctx.dev, ctx.world
    import quorum_b200

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    step = W.step
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    total_out = W.out_total()
    # ---- burst: exactly K steps ----
    barrier()
    l0 = quorum_b200.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    launches = quorum_b200.launch_count() - l0
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1) / args.steps)
    if world > 1:
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt)
        launches = int(lt.item())
    # ---- sustained: the same call back to back for at least sustained_s seconds of device time ----
    sustained = None
    if sustained_s > 0:
        n_sus = max(args.steps, int(np.ceil(sustained_s * 1e3 / max(ms, 1e-3) * 1.05)))
        barrier()
        ev0.record()
        for _ in range(n_sus):
            step()
        ev1.record()
        torch.cuda.synchronize()
        barrier()
        tot_ms = ev0.elapsed_time(ev1)
        ms_sus = max_over_ranks(tot_ms / n_sus)
        sustained = {"steps": n_sus, "timed_region_s": tot_ms * 1e-3, "ms_per_step": ms_sus,
                     "value": world * W.total_in / (ms_sus * 1e-3) / 1e9, "unit": "GB/s"}
    # ---- per-kernel split of a call (outside the timed regions): the library records an event after each launch ----
    kernels = None
    if want_kernels and ctx.rank == 0:
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
    # ---- end to end through the host-buffer call, and the raw copies of the same byte counts ----
    e2e = None
    if want_e2e:
        try:
            h_data = torch.empty(W.total_in, dtype=torch.uint8).pin_memory()
            h_data.copy_(W.data)
            h_meta = W.host()
            sess = W.session()
            e2e_steps = max(3, min(args.steps, 5))
            for _ in range(2):
                W.e2e_call(sess, h_data, *h_meta)
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                h2d, d2h = W.e2e_call(sess, h_data, *h_meta)
            torch.cuda.synchronize()
            dt = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
            e2e = {"value": world * W.total_in / dt / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(h2d),
                   "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": dt * 1e3}
            # the platform ceiling: the same bytes up and down on two streams, no kernels (all ranks at once)
            s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
            up_n, dn_n = min(int(h2d), W.total_in), min(int(d2h), sess.h_out.numel(), sess.d_out.numel())
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                with torch.cuda.stream(s_up):
                    sess.d_data[:up_n].copy_(h_data[:up_n], non_blocking=True)
                with torch.cuda.stream(s_dn):
                    sess.h_out[:dn_n].copy_(sess.d_out[:dn_n], non_blocking=True)
            torch.cuda.synchronize()
            dtc = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
            e2e["copy_ceiling"] = {"value": world * W.total_in / dtc / 1e9, "unit": "GB/s", "ms_per_step": dtc * 1e3,
                                   "what": "pinned H2D of the input bytes and D2H of the output bytes on two streams, no kernels, all ranks at once"}
            e2e["frac_of_copy_ceiling"] = e2e["value"] / e2e["copy_ceiling"]["value"]
            del h_data, sess
        except Exception as ex:  # host memory for the pinned copies can be short on a shared box
            e2e = {"unavailable": str(ex)[:200]}
    alg_bytes = W.total_in + total_out + W.meta_bytes
    achieved = alg_bytes / (ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": ctx.peak, "unit": "GB/s", "frac": achieved / ctx.peak,
                "frac_of_nominal_8000": achieved / NOMINAL_HBM_GBS, "traffic": None, "peak_source": ctx.peak_src,
                "algorithmic_bytes_per_launch": int(alg_bytes),
                "kernel": "whole call (all kernels of one qb200_*_batch call), CUDA events on the launching stream"}
    if sustained:
        roofline["frac_sustained"] = alg_bytes / (sustained["ms_per_step"] * 1e-3) / 1e9 / ctx.peak
    if kernels:
        roofline["kernels"] = kernels
        roofline["kernels_source"] = ("qb200_profile_begin/end: CUDA events after each launch on the launching stream, separate pass "
                                      "after the timed region")
        # the two kernels that touch the text, each against its OWN algorithmic bytes: the tokenizer reads every input byte
        # once (+ one count byte per KiB written); the copy kernel reads every kept byte and writes it once (+ 16 B of
        # offsets per item).  The whole-call fraction above is lower because the kept bytes are read by both.
        own = {"tokenize_kernel": W.total_in + W.total_in // 1024,
               "strip_copy_kernel": 2 * total_out + 16 * W.n_items, "stream_finish_kernel": 2 * total_out + 16 * W.n_units}
        per = []
        for k in kernels:
            if k["name"] in own and k["ms_per_step"] > 0:
                a = own[k["name"]] / (k["ms_per_step"] * 1e-3) / 1e9
                per.append({"name": k["name"], "algorithmic_bytes_per_launch": int(own[k["name"]]), "achieved": a, "frac": a / ctx.peak})
        if per:
            roofline["dominant_kernels"] = per
    tj = ctx.traffic.get(W.name)
    # the ncu capture is of the workload at its configs size: a batch of another size (the sharded 10 GB cfg5) has none
    if tj and tj.get("dram_bytes_per_launch") and not 0.5 < tj["dram_bytes_per_launch"] / alg_bytes < 4.0:
        tj = None
    if tj:
        roofline["traffic"] = tj.get("dram_bytes_per_launch")
        roofline["traffic_ratio"] = (tj.get("dram_bytes_per_launch") or 0) / alg_bytes if tj.get("dram_bytes_per_launch") else None
        roofline["traffic_source"] = tj.get("source")
        roofline["traffic_commit"] = tj.get("commit")
    rec = {
        "workload": W.wl["name"], "value": world * W.total_in / (ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms,
        W.items_name + "_per_s": world * W.n_items / (ms * 1e-3), "units_per_s": world * W.n_units / (ms * 1e-3),
        "units_per_gpu": W.n_units, "unit_bytes": W.unit_b, "input_bytes_per_gpu": W.total_in, "output_bytes_per_gpu": total_out,
        "timed_region_s": ms * args.steps * 1e-3, "sustained": sustained, "roofline": roofline, "e2e": e2e,
        "gpu_launches": int(launches),
    }
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=HEADLINE, choices=sorted(WORKLOADS), help="the headline workload of the line")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--units", type=int, default=0, help="override the number of units per GPU (debug)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--only-headline", action="store_true", help="skip the `configs` map and the sharded cfg5 batch")
    ap.add_argument("--all-configs", action="store_true", help="with --gpus N > 1: also fill the `configs` map (replica per rank)")
    ap.add_argument("--no-sustained", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    name = args.workload
    others = [n for n in ("cfg3", "cfg2", "cfg4", "cfg5") if n != name]

    if args.impl == "reference":
        run_reference_arm(args, [name] + ([] if args.only_headline else others))
        return

    import torch
    import torch.distributed as dist

    ctx = Ctx()
    ctx.torch, ctx.dist = torch, dist
    ctx.rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    ctx.world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: quorum_b200 has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    ctx.dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa(torch, local_rank) if ctx.world > 1 else None
    if ctx.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=ctx.dev)
    from quorum_b200 import synth
    ctx.peak, ctx.peak_src = read_peaks()
    ctx.traffic = {}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(prof):
        try:
            with open(prof) as f:
                ctx.traffic = json.load(f)
        except Exception:
            ctx.traffic = {}
    sus = 0.0 if args.no_sustained else 1.0
    sampler = ClockSampler(local_rank)
    sampler.start()

    def run_one(nm, shard=None, e2e=True):
        wl = dict(WORKLOADS[nm])
        W = make_workload(ctx, nm, wl, args.units if nm == name else 0, shard)
        rec = measure(ctx, W, args, want_e2e=e2e and not args.no_e2e, want_kernels=True, sustained_s=sus)
        if shard is not None:
            rec["shard"] = W.shard_info
        W.keep = None
        del W
        torch.cuda.empty_cache()
        return rec

    head = run_one(name)
    configs, sharded = {}, None
    if not args.only_headline:
        if ctx.world == 1 or args.all_configs:
            for nm in others:
                configs[nm] = run_one(nm)
        # BASELINE.json configs[4] as stated: one 10 GB batch, partitioned by bytes over the ranks (strong scaling)
        sharded = run_one("cfg5", shard=(ctx.rank, ctx.world))
        total = sharded["shard"]["total_bytes"]
        for key in ("value",):
            sharded[key] = total / (sharded["ms_per_step"] * 1e-3) / 1e9
        if sharded.get("sustained"):
            sharded["sustained"]["value"] = total / (sharded["sustained"]["ms_per_step"] * 1e-3) / 1e9
        if sharded.get("e2e") and "value" in sharded["e2e"]:
            sharded["e2e"]["value"] = total / (sharded["e2e"]["ms_per_step"] * 1e-3) / 1e9
            if "copy_ceiling" in sharded["e2e"]:
                sharded["e2e"]["copy_ceiling"]["value"] = total / (sharded["e2e"]["copy_ceiling"]["ms_per_step"] * 1e-3) / 1e9
                sharded["e2e"]["frac_of_copy_ceiling"] = sharded["e2e"]["value"] / sharded["e2e"]["copy_ceiling"]["value"]
        sharded["scaling"] = "strong"
        sharded["workload"] = (f"cfg5 as BASELINE.json configs[4]: ONE adversarial batch of {total} B partitioned by bytes over "
                               f"{ctx.world} rank(s) (sharding.partition_by_bytes), no collective, host gathers")
        sharded["note"] = ("value / e2e = the whole batch's bytes over the max-over-ranks time; per-GPU fields describe rank 0's shard; "
                           "roofline.frac is rank 0's shard against one GPU's peak")
        for k in ("responses_per_s", "units_per_s"):
            sharded.pop(k, None)
    clocks = sampler.stop()

    if ctx.rank != 0:
        if ctx.world > 1:
            dist.destroy_process_group()
        return

    cpu = None
    if ctx.world == 1 and not args.no_cpu:
        wl = dict(WORKLOADS[name])
        cpu = cpu_baseline_record(wl, name, build_base(wl, name), full=True)
        for nm, rec in configs.items():
            wo = dict(WORKLOADS[nm])
            rec["cpu_baseline"] = cpu_baseline_record(wo, nm, build_base(wo, nm), full=False)

    line = {
        "metric": "filtered+concatenated GB/s (input bytes)", "value": head["value"], "unit": "GB/s", "n_gpus": ctx.world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic (quorum_b200/synth.py, fixed seeds)",
        "config": {"workload": head["workload"], "units_per_gpu": head["units_per_gpu"], "unit_bytes": head["unit_bytes"],
                   "input_bytes_per_gpu": head["input_bytes_per_gpu"], "output_bytes_per_gpu": head["output_bytes_per_gpu"],
                   "thinking_tags": synth.TAGS5, "l2": "inputs larger than L2 (no flush needed)",
                   "sharding": f"{ctx.world} rank(s), one replica of the workload per rank, independent units, no collective",
                   "numa_cpus": numa, "head": git_head()},
        "responses_per_s": head.get("responses_per_s"), "chunks_per_s": head.get("chunks_per_s"), "units_per_s": head["units_per_s"],
        "timed_region_s": head["timed_region_s"], "sustained": head["sustained"],
        "roofline": head["roofline"], "cpu_baseline": cpu, "e2e": head["e2e"], "gpu_launches": head["gpu_launches"],
        "clocks": clocks, "configs": configs, "sharded": sharded,
    }
    if line["chunks_per_s"] is None:
        line.pop("chunks_per_s")
    if line["responses_per_s"] is None:
        line.pop("responses_per_s")
    print(json.dumps(line), flush=True)
    if ctx.world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
