#!/usr/bin/env python3
"""Multi-GPU probe (run under torchrun): P2P reducer vs NCCL all-reduce on an arena-shaped buffer — correctness
for several bucket sizes and steps, then bandwidth.  One rank per GPU."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from hetseq_9cme_b200.parallel.arena import ParamArena
    from hetseq_9cme_b200.parallel.p2p import P2PReducer
    torch.manual_seed(0)
    sizes = [3, 64 * 7, 1000, 65536 * 2, 1 << 20, (1 << 22) + 64, 30522 * 768, 768 * 3072, 768 * 3072, 768 * 3072, 2304 * 768]
    params = [torch.nn.Parameter(torch.zeros(s, device="cuda")) for s in sizes]
    from hetseq_9cme_b200.parallel.p2p import SQ_SLOT, symmetric_grad_alloc
    arena = ParamArena(params, device="cuda", bucket_bytes=8 * 1024 * 1024, grad_alloc=symmetric_grad_alloc)
    red = P2PReducer(arena, world)
    if rank == 0:
        print("multicast base: %#x" % red.mc_base, flush=True)
    ok = True
    for step in range(4):
        g = torch.Generator(device="cuda").manual_seed(100 * step + rank)
        arena.grad.copy_(torch.randn(arena.numel, device="cuda", generator=g))
        ref = arena.grad.clone()
        dist.all_reduce(ref)
        arena.tail.zero_()
        red.begin(True)
        if step % 2 == 0:
            red.ready(list(reversed(arena.params)))      # announce in backward order -> per-bucket launches
        red.finish()
        red.reduce_tail()
        torch.cuda.synchronize()
        nsq, nsq_ref = arena.tail[SQ_SLOT].item(), ref.double().pow(2).sum().item()
        if rank == 0:
            print("  fused sum(g^2) %.6e vs %.6e" % (nsq, nsq_ref), flush=True)
        ok = ok and abs(nsq - nsq_ref) < 1e-4 * nsq_ref
        err = (arena.grad - ref).abs().max().item() / ref.abs().max().item()
        same = torch.tensor([arena.grad.double().sum().item()], device="cuda", dtype=torch.float64)
        lo, hi = same.clone(), same.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        bit_identical = lo.item() == hi.item()
        if rank == 0:
            print("step %d buckets=%d rel_err=%.3e replicas_identical=%s" % (step, len(arena.buckets), err, bit_identical), flush=True)
        ok = ok and err < 1e-5 and bit_identical
    # bandwidth: whole arena, device-timed, max over ranks
    for name, fn in (("p2p", lambda: red.all_reduce_flat()), ("nccl", lambda: dist.all_reduce(arena.grad))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        nbytes = arena.numel * 4
        if rank == 0:
            print("BW %s: %.3f ms for %.1f MB  -> algbw %.1f GB/s, busbw %.1f GB/s" % (
                name, t.item(), nbytes / 1e6, nbytes / t.item() / 1e6, nbytes / t.item() / 1e6 * 2 * (world - 1) / world), flush=True)
    # block-count sweep for the after-backward launches (whole arena, largest bucket = 94 MB word embedding)
    for blocks in (32, 64, 128, 148, 222, 296):
        red.tail_blocks = blocks
        for _ in range(2):
            red.all_reduce_flat()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            red.all_reduce_flat()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print("SWEEP tail_blocks=%3d: %.3f ms  algbw %.1f GB/s" % (blocks, t.item(), arena.numel * 4 / t.item() / 1e6), flush=True)
    if rank == 0:
        print("P2P PROBE", "PASS" if ok else "FAIL", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
