#!/usr/bin/env python
"""How fast does one GPU write into a peer's memory over NVLink?  Every rank pushes a 2 GB buffer into its
right neighbour's symmetric-memory buffer (all ranks at once, as in the exchange) with 16-byte register
stores and with bulk async stores of several sizes; one JSON line per variant (GB/s per GPU, max time over
ranks, clocks).  N = 1 measures the same kernels on local memory.

    tools/run_n.sh 2 tools/nvlink_probe.py
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import pyspark_tf_gke_b200 as pkg  # noqa: E402
from pyspark_tf_gke_b200 import _lib  # noqa: E402
from tools.clocks import ClockSampler  # noqa: E402


def all_to_all(eng, world, rank, local, handle, buf, src, nbytes):
    """Every rank writes an equal share of its buffer into EVERY other rank's buffer at once (the exchange's traffic
    pattern): one probe kernel per peer, each on its own stream."""
    share = nbytes // world // 4096 * 4096
    streams = [torch.cuda.Stream() for _ in range(world - 1)]
    peers = [(rank + 1 + i) % world for i in range(world - 1)]
    for name, mode, chunk, ctas in (("all-to-all, st.v2.s64, 16 CTAs per peer", 0, 0, 16), ("all-to-all, st.v2.s64, 64 CTAs per peer", 0, 0, 64),
                                    ("all-to-all, st.v2.s64, 148 CTAs per peer", 0, 0, 148), ("all-to-all, bulk store 4096 B, 64 CTAs per peer", 1, 4096, 64)):
        def run():
            for st, p in zip(streams, peers):
                st.wait_stream(torch.cuda.current_stream())
                dst = int(handle.buffer_ptrs[p]) + rank * share
                _lib.check(eng.lib.etl_peer_write_probe(eng._h, src.data_ptr() + p * share, dst, share, mode, chunk, ctas, st.cuda_stream))
            for st in streams:
                torch.cuda.current_stream().wait_stream(st)
        run()
        torch.cuda.synchronize()
        dist.barrier()
        sampler = ClockSampler(local)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            run()
        e1.record()
        torch.cuda.synchronize()
        clocks = sampler.stop()
        t = torch.tensor([e0.elapsed_time(e1) / 3], dtype=torch.float64, device=eng.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sent = share * (world - 1)
        if rank == 0:
            print(json.dumps({"probe": name, "n_gpus": world, "sent_gb_per_gpu": sent / 1e9, "ms": float(t.item()),
                              "gbs_per_gpu": sent / 1e9 / (float(t.item()) * 1e-3), "clocks": clocks}), flush=True)
        dist.barrier()


def main():
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    eng = pkg.Engine(local)
    nbytes = 2 << 30
    src = torch.arange(nbytes // 8, dtype=torch.int64, device=eng.device)
    if world > 1:
        import torch.distributed._symmetric_memory as symm

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        buf = symm.empty(nbytes // 8, dtype=torch.int64, device=eng.device)
        handle = symm.rendezvous(buf, dist.group.WORLD.group_name)
        dst = int(handle.buffer_ptrs[(rank + 1) % world])
    else:
        buf = torch.empty(nbytes // 8, dtype=torch.int64, device=eng.device)
        dst = buf.data_ptr()
    if world > 2:
        all_to_all(eng, world, rank, local, handle, buf, src, nbytes)
        eng.close()
        dist.destroy_process_group()
        return
    stream = torch.cuda.current_stream().cuda_stream
    variants = [("st.v2.s64 from registers", 0, 0, 0), ("same, 1 CTA/SM", 0, 0, 148)]
    variants += [(f"bulk store {c} B", 1, c, ctas) for c in (256, 1024, 4096, 32768) for ctas in (148, 592)]
    for name, mode, chunk, ctas in variants:
        def run():
            _lib.check(eng.lib.etl_peer_write_probe(eng._h, src.data_ptr(), dst, nbytes, mode, chunk, ctas, stream))
        run()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            run()
        e1.record()
        torch.cuda.synchronize()
        clocks = sampler.stop()
        t = torch.tensor([e0.elapsed_time(e1) / 3], dtype=torch.float64, device=eng.device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            handle.barrier()
        ok = bool(torch.equal(buf, src)) if world == 1 else None
        if world > 1:       # what landed here came from the left neighbour: the same arange
            ok = bool(torch.equal(buf, src))
        if rank == 0:
            print(json.dumps({"probe": name, "ctas": ctas or 592, "n_gpus": world, "gb": nbytes / 1e9, "ms": float(t.item()),
                              "gbs_per_gpu": nbytes / 1e9 / (float(t.item()) * 1e-3), "data_ok": ok, "clocks": clocks}), flush=True)
        buf.zero_()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
