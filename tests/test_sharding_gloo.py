"""World-size-2 check of the multi-GPU host logic on CPU (gloo): contiguous byte-balanced shards of independent units,
no data-path collective, host-side gather.  The per-rank compute here is the oracle (this test has no GPU); what is
under test is quorum_b200.sharding and the gather, i.e. that sharded == unsharded."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, tmpdir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import c_oracle
    from quorum_b200 import sharding, synth

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tags = synth.TAGS5
    units = synth.cfg5_units(2 << 20, seed=5)                       # same bytes on every rank (seeded)
    data, in_off = synth.pack(units)
    req_off = np.arange(0, len(units) + 1, 3, dtype=np.int64)
    if req_off[-1] != len(units):
        req_off = np.append(req_off, len(units))
    n_req = len(req_off) - 1
    req_bytes = [int(in_off[req_off[q + 1]] - in_off[req_off[q]]) for q in range(n_req)]
    lo, hi = sharding.partition_by_bytes(req_bytes, world)[rank]
    b0, b1, off_l, req_l = sharding.shard_offsets(in_off, req_off, lo, hi)
    out, out_off, _ = c_oracle.strip_concat_batch(c_oracle.Tags(tags), data[b0:b1] if b1 > b0 else np.zeros(1, np.uint8),
                                                  np.ascontiguousarray(off_l), np.ascontiguousarray(req_l), b"\n--\n",
                                                  c_oracle.FLAG_HIDE)
    # host gather: lengths first, then the bytes (object gather keeps the test small)
    gathered = [None] * world
    dist.all_gather_object(gathered, (lo, hi, bytes(out), out_off.tolist()))
    if rank == 0:
        full_out, full_off, _ = c_oracle.strip_concat_batch(c_oracle.Tags(tags), data, in_off, req_off, b"\n--\n",
                                                            c_oracle.FLAG_HIDE)
        pieces, offs, base = [], [0], 0
        for (l, h, ob, oo) in sorted(gathered):
            pieces.append(ob)
            offs.extend(base + x for x in oo[1:])
            base += oo[-1]
        ok = b"".join(pieces) == bytes(full_out) and offs == full_off.tolist()
        covered = sorted((l, h) for l, h, _, _ in gathered)
        ok = ok and covered[0][0] == 0 and covered[-1][1] == n_req and all(a[1] == b[0] for a, b in zip(covered, covered[1:]))
        with open(os.path.join(tmpdir, "result"), "w") as f:
            f.write("ok" if ok else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_matches_unsharded(tmp_path):
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "result").read_text() == "ok"
