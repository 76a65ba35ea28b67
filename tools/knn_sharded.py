#!/usr/bin/env python
"""BASELINE.json configs[4]: brute-force Hamming kNN-2 (k = 2, ratio test), nq x nt 256-bit descriptors, QUERY rows
sharded over the GPUs of one node, one NCCL all-gather of 8-byte result rows — through the C ABI
(orbfe_comm_create / orbfe_knn2_sharded).  Strong scaling: the problem is fixed, every rank holds the same inputs
(i.i.d. rows, seeds 1 and 2, 10 % planted near-duplicates, ratio 0.7, TH_LOW 50).

  python tools/knn_sharded.py [--nq 1000000 --nt 1000000 --engine auto|popc|mma|fp4 --steps 3 --check]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/knn_sharded.py ...

Rank 0 prints one JSON line.  --check: sharded result == single-GPU result of rank 0 on all queries == CPU oracle on a
4096-query subsample, plus the planted property on all queries; exit code 1 on any mismatch."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(nq, nt, engine="auto", steps=3, warmup=1, check=False, rank=None, world=None, local_rank=None, oracle_sub=4096):
    import torch
    import torch.distributed as dist
    import orb_slam2_annotate_b200 as orbfe
    from orb_slam2_annotate_b200 import synth

    rank = int(os.environ.get("RANK", "0")) if rank is None else rank
    world = int(os.environ.get("WORLD_SIZE", "1")) if world is None else world
    local_rank = int(os.environ.get("LOCAL_RANK", "0")) if local_rank is None else local_rank
    torch.cuda.set_device(local_rank)
    own_group = False
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        own_group = True

    train = synth.descriptors(nt, 1)
    query, rows, src, nflip = synth.planted_descriptors_fast(train, nq, 2)
    dq, dt = torch.from_numpy(query).cuda(), torch.from_numpy(train).cuda()
    comm = orbfe.Comm.from_torch(local_rank)
    mt = orbfe.ORBmatcher(0.7, device=local_rank)
    mt.set_knn_engine(engine)
    per = orbfe.shard_rows(nq, world)
    out = torch.zeros((world * per, 8), dtype=torch.uint8, device="cuda")
    ms = torch.cuda.ExternalStream(mt.stream, device=local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        mt.knn2_sharded(comm, dq.data_ptr(), nq, dt.data_ptr(), nt, out.data_ptr(), 50, 0.7)

    for _ in range(warmup):
        step()
    mt.sync()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ms)
    for _ in range(steps):
        step()
    e1.record(ms)
    mt.sync()
    barrier()
    ms_step = e0.elapsed_time(e1) / steps
    # the gather alone (same payload, nothing to wait for): what the collective adds to a step
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gather_us = 0.0
    if world > 1:
        chunks = list(out.view(world, per * 8).unbind(0))
        dist.all_gather(chunks, chunks[rank].clone())
        torch.cuda.synchronize()
        g0.record()
        for _ in range(10):
            dist.all_gather(chunks, chunks[rank])
        g1.record()
        torch.cuda.synchronize()
        gather_us = g0.elapsed_time(g1) * 100.0
        step(); mt.sync()                      # the timing loop above reused `out`: recompute before checking
    t = torch.tensor([ms_step, gather_us], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step, gather_us = t.tolist()

    parity = "not checked"
    ok = True
    if check:
        idx, d1, d2 = orbfe.unpack_results(out.cpu().numpy(), nq)
        problems = []
        # planted property on ALL queries: a planted row's best match is its source at exactly nflip bits
        if not np.array_equal(d1[rows], nflip.astype(np.int32)):
            problems.append("planted d1")
        acc = idx[rows] >= 0
        if not np.array_equal(idx[rows][acc], src[acc].astype(np.int32)):
            problems.append("planted idx")
        if nt >= 100000 and (acc.mean() < 0.99 or (idx >= 0).sum() != acc.sum()):
            problems.append("acceptance")
        if rank == 0:
            one = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
            mt.knn2_device(dq.data_ptr(), nq, dt.data_ptr(), nt, one[0].data_ptr(), one[1].data_ptr(), one[2].data_ptr(), 50, 0.7)
            mt.sync()
            o = one.cpu().numpy()
            if not (np.array_equal(o[0], idx) and np.array_equal(o[1], d1) and np.array_equal(o[2], d2)):
                problems.append("sharded != single GPU")
            from oracle import oracle_py as O
            sub = np.random.default_rng(3).choice(nq, min(oracle_sub, nq), replace=False)
            oi, o1, o2 = O.knn2(query[sub], train, 50, 0.7)
            if not (np.array_equal(oi, idx[sub]) and np.array_equal(o1, d1[sub]) and np.array_equal(o2, d2[sub])):
                problems.append("!= oracle on the subsample")
        flag = torch.tensor([len(problems)], device="cuda")
        if world > 1:
            dist.all_reduce(flag)
        ok = int(flag.item()) == 0
        parity = "ok" if ok else "MISMATCH " + ",".join(problems)

    line = None
    if rank == 0:
        gcmp = nq * nt / (ms_step * 1e-3) / 1e9
        line = {"metric": "Hamming Gcmp/s (brute-force kNN-2, 256-bit rows, queries sharded, NCCL result gather)",
                "value": gcmp, "unit": "Gcmp/s", "n_gpus": world, "scaling": "strong", "nq": nq, "nt": nt, "engine": engine,
                "ms_per_step": ms_step, "steps": steps, "gather_us": gather_us, "gather_bytes_per_rank": per * 8,
                "gather_bytes_total": world * per * 8, "parity": parity}
    comm.close()
    if own_group:
        dist.destroy_process_group()
    return line, ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nq", type=int, default=1_000_000)
    ap.add_argument("--nt", type=int, default=1_000_000)
    ap.add_argument("--engine", default="auto", choices=["auto", "popc", "mma", "fp4"])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    line, ok = run(a.nq, a.nt, a.engine, a.steps, a.warmup, a.check)
    if line is not None:
        print(json.dumps(line), flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
