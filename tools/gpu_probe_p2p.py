#!/usr/bin/env python3
"""Multi-GPU probe (run under torchrun, one rank per GPU): the communication runtime and the peer-to-peer kernels
against NCCL on an arena-shaped buffer.

1. all-reduce mode: correctness for several bucket sizes and steps (vs ``ncclAllReduce``), bit-identical replicas,
   fused sum of squares;
2. sharded mode: reduce-scatter per bucket + statistics tail + owned-slice Adam with parameter broadcast, against
   NCCL all-reduce + the replicated ``adam_flat`` kernel on copies (fp32 masters and bf16 working copies);
3. device-timed bandwidth (max over ranks) of each piece.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def identical_across_ranks(t):
    s = t.double().sum().reshape(1)
    q = t.double().pow(2).sum().reshape(1)
    v = torch.cat([s, q, -s, -q])
    dist.all_reduce(v, op=dist.ReduceOp.MAX)
    return v[0].item() == -v[2].item() and v[1].item() == -v[3].item()


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from hetseq_9cme_b200 import optim
    from hetseq_9cme_b200.ops import kernels
    from hetseq_9cme_b200.parallel import p2p
    from hetseq_9cme_b200.parallel.arena import ParamArena
    quick = "--quick" in sys.argv
    torch.manual_seed(0)
    sizes = [3, 64 * 7, 1000, 65536 * 2, 1 << 20, (1 << 22) + 64, 30522 * 768, 768 * 3072, 768 * 3072, 768 * 3072, 2304 * 768]
    ok = True
    rt = p2p.CommRuntime.get()
    if rank == 0:
        print("comm runtime: %s, multicast %s" % ("own (VMM + Unix-socket FD passing)" if rt is not None else "UNAVAILABLE (%s)" % p2p.CommRuntime._failed,
                                                  getattr(rt, "multicast", False)), flush=True)

    def alloc_for(sharded, bf16):
        def alloc(kind, numel, dtype):
            if kind == "grad":
                return p2p.symmetric_grad_alloc(numel)
            if sharded and kind == ("work" if bf16 else "master"):
                return p2p.symmetric_alloc(numel, dtype)
            return None
        return alloc

    # ---- 1. all-reduce mode ------------------------------------------------------------------------------------
    params = [torch.nn.Parameter(torch.randn(s, device="cuda") * 0.02) for s in sizes]
    arena = ParamArena(params, device="cuda", bucket_bytes=8 * 1024 * 1024, alloc=alloc_for(False, False))
    red = p2p.P2PReducer(arena, world)
    if rank == 0:
        print("mapping %s, multicast base %#x, %d buckets" % (red.mapping, red.mc_base, len(arena.buckets)), flush=True)
    for step in range(2 if quick else 4):
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
        nsq, nsq_ref = arena.tail[p2p.SQ_SLOT].item(), ref.double().pow(2).sum().item()
        err = (arena.grad - ref).abs().max().item() / ref.abs().max().item()
        same = identical_across_ranks(arena.grad)
        if rank == 0:
            print("allreduce step %d: rel_err %.3e, replicas identical %s, fused sum(g^2) %.6e vs %.6e" % (step, err, same, nsq, nsq_ref), flush=True)
        ok = ok and err < 1e-5 and same and abs(nsq - nsq_ref) < 1e-4 * nsq_ref
    t_p2p = timed(lambda: red.all_reduce_flat())
    t_nccl = timed(lambda: dist.all_reduce(arena.grad))
    nbytes = arena.numel * 4
    if rank == 0:
        for name, t in (("p2p all-reduce", t_p2p), ("nccl all-reduce", t_nccl)):
            print("BW %-18s %.3f ms for %.1f MB -> algbw %.1f GB/s, busbw %.1f GB/s" % (
                name, t, nbytes / 1e6, nbytes / t / 1e6, nbytes / t / 1e6 * 2 * (world - 1) / world), flush=True)

    # ---- 2. sharded mode ---------------------------------------------------------------------------------------
    if rt is None:
        if rank == 0:
            print("P2P PROBE", "PASS (all-reduce mode only: no comm runtime)" if ok else "FAIL", flush=True)
        dist.barrier(); dist.destroy_process_group()
        return
    for bf16 in (False, True):
        torch.manual_seed(1)
        params = [torch.nn.Parameter(torch.randn(s, device="cuda") * 0.02) for s in sizes]
        dt = torch.bfloat16 if bf16 else torch.float32
        ar = ParamArena(params, device="cuda", bucket_bytes=8 * 1024 * 1024, compute_dtype=dt, alloc=alloc_for(True, bf16))
        rs = p2p.P2PReducer(ar, world, sharded=True)
        opt = optim.Adam(ar.params, lr=1e-3, weight_decay=0.01)
        opt.attach_arena(ar)
        ws = kernels.Workspace.get("cuda")
        # replicated reference state
        r_master, r_m, r_v = ar.master.clone(), torch.zeros_like(ar.master), torch.zeros_like(ar.master)
        r_work = ar.work.clone() if bf16 else None
        ss = torch.tensor([float(world * 4)], device="cuda")
        for step in range(1, 3 if quick else 4):
            g = torch.Generator(device="cuda").manual_seed(1000 * step + rank)
            ar.grad.copy_(torch.randn(ar.numel, device="cuda", generator=g) * (3.0 if step == 2 else 0.01))
            ref = ar.grad.clone()
            dist.all_reduce(ref)
            nsq_ref = torch.zeros(1, device="cuda")
            kernels.sumsq(ref, nsq_ref)
            kernels.adam_flat(r_master, ref, r_m, r_v, ar.numel, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01,
                              step=step, max_norm=1.0, norm_sq=nsq_ref, sample_size=ss, zero_grad=True, p_bf16=r_work)
            ar.tail.zero_()
            rs.begin(True)
            rs.ready(list(reversed(ar.params)))
            rs.finish()
            rs.reduce_tail()
            opt.sharded_step(rs, max_norm=1.0, norm_sq=ar.tail[p2p.SQ_SLOT:p2p.SQ_SLOT + 1], sample_size=ss, stats_out=ws.step_stats)
            torch.cuda.synchronize()
            clipped = ws.step_stats[2].item()
            if bf16:
                w_ok = torch.equal(ar.work, r_work) or (ar.work.float() - r_work.float()).abs().max().item() <= 2 ** -7 * r_work.float().abs().max().item()
                same = identical_across_ranks(ar.work.float())
            else:
                w_ok = (ar.master - r_master).abs().max().item() <= 2e-6 * r_master.abs().max().item()
                same = identical_across_ranks(ar.master)
            own_ok = True
            for bi in range(len(ar.buckets)):
                lo, hi = rs.slice_of(bi, rank)
                for mine, theirs in ((opt._flat_state["exp_avg"], r_m), (opt._flat_state["exp_avg_sq"], r_v), (ar.master, r_master)):
                    if hi > lo:
                        d = (mine[lo:hi] - theirs[lo:hi]).abs().max().item()
                        own_ok = own_ok and d <= 2e-6 * max(theirs[lo:hi].abs().max().item(), 1e-20) + 1e-12
            zero_ok = ar.grad.abs().max().item() == 0.0
            if rank == 0:
                print("sharded %s step %d: params match replicated Adam %s, replicas identical %s, owned moments/masters %s, "
                      "grads cleared %s, clipped %.0f" % ("bf16" if bf16 else "fp32", step, w_ok, same, own_ok, zero_ok, clipped), flush=True)
            ok = ok and w_ok and same and own_ok and zero_ok
        # after gather: every rank holds the full optimizer state
        flats = opt.flat_state_tensors() + ([ar.master] if bf16 else [])
        rs.gather_sharded(flats)
        full = all((a_ - b_).abs().max().item() <= 2e-6 * max(b_.abs().max().item(), 1e-20) for a_, b_ in zip(flats, [r_m, r_v] + ([r_master] if bf16 else [])))
        ok = ok and full
        if rank == 0:
            print("sharded %s: gathered state equals the replicated state: %s" % ("bf16" if bf16 else "fp32", full), flush=True)

        def rs_all():
            rs.begin(True)
            rs.finish()

        def sharded_update():
            opt.sharded_step(rs, max_norm=0.0, norm_sq=None, sample_size=ss, stats_out=ws.step_stats)

        def replicated_update():
            kernels.adam_flat(r_master, ar.grad, r_m, r_v, ar.numel, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01,
                              step=5, sample_size=ss, zero_grad=True, p_bf16=r_work)
        t_rs, t_sh, t_rep = timed(rs_all), timed(sharded_update), timed(replicated_update)
        nb = ar.numel * 4
        if rank == 0:
            print("BW %s: reduce-scatter %.3f ms (%.1f MB) | sharded Adam + broadcast %.3f ms | replicated Adam %.3f ms" % (
                "bf16" if bf16 else "fp32", t_rs, nb / 1e6, t_sh, t_rep), flush=True)
        if not quick:
            # sweep of the data-path knobs (all ranks change them in lock-step)
            for peer in (False, True):
                rs.rs_peer = peer
                for tune in ((4, 8, 20, 24) if not peer else (1, 2, 18)):
                    os.environ["B200_P2P_TUNE"] = str(tune)
                    for blocks in (32, 64, 128, 296):
                        rs.tail_blocks = blocks
                        t = timed(rs_all, iters=6, warm=2)
                        if rank == 0:
                            print("SWEEP %s reduce-scatter %-5s U=%d %-11s blocks=%3d: %.3f ms" % (
                                "bf16" if bf16 else "fp32", "peer" if peer else "nvls", tune & 15, "contiguous" if tune & 16 else "interleaved", blocks, t), flush=True)
            os.environ.pop("B200_P2P_TUNE", None)
            rs.rs_peer, rs.tail_blocks = False, 128
            for peer in (False, True):
                rs.bcast_peer = peer
                for blocks in (148 * 2, 148 * 4, 148 * 8, 148 * 16):
                    os.environ["B200_SHARDED_ADAM_BLOCKS"] = str(blocks)
                    t = timed(sharded_update, iters=6, warm=2)
                    if rank == 0:
                        print("SWEEP %s sharded Adam + broadcast %-9s blocks=%4d: %.3f ms" % (
                            "bf16" if bf16 else "fp32", "peer st" if peer else "multicast", blocks, t), flush=True)
            os.environ.pop("B200_SHARDED_ADAM_BLOCKS", None)
            rs.bcast_peer = False
    if rank == 0:
        print("P2P PROBE", "PASS" if ok else "FAIL", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
