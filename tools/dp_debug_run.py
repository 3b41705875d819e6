#!/usr/bin/env python3
"""Run under torchrun: data-parallel training steps of a small BERT through the Controller, with checks.

    torchrun --nproc-per-node N tools/dp_debug_run.py OUT [--mode sharded|replicated|torch] [--graph] [--update-freq U]
             [--steps K] [--uneven] [--precision tf32|bf16] [--compare-single] [--checkpoint]

* B200_P2P_DEBUG=1 in the environment switches the reducer's own verification on (eager steps only): every bucket is
  compared with the NCCL sum of its pre-launch copy and the device-side epochs with the host's count.
* every run checks that all ranks end with bit-identical parameters;
* ``--compare-single``: rank 0 repeats the run as ONE process with ``--update-freq N*U`` on the same batches (dropout
  is 0, so both runs compute the same global-batch gradient) and compares the parameters — a missing cross-rank
  sum of any gradient (e.g. the tied word-embedding / decoder weight) shows up here;
* ``--uneven``: the shards have an odd number of batches, so the last step of the epoch gives one rank the empty
  filler batch — that rank steps eagerly while its peer may replay a CUDA graph;
* ``--checkpoint``: save at the end (sharded optimizer state is gathered), reload into a fresh Controller, compare.

Used by tests/test_gpu_multi.py."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

CFG = dict(vocab_size=1000, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
           max_position_embeddings=128, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)


def build(out, a, world, rank, local, update_freq, graph, comm, extra=()):
    from hetseq_9cme_b200 import options, tasks
    from hetseq_9cme_b200.controller import Controller
    d = os.path.join(out, "data")
    argv = ["--task", "bert", "--data", d, "--dict", os.path.join(out, "v.txt"), "--config_file", os.path.join(out, "c.json"),
            "--max-sentences", "8", "--lr", "1e-3", "--warmup-updates", "2", "--total-num-update", "50", "--clip-norm", "1.0",
            "--weight-decay", "0.01", "--valid-subset", "test", "--save-dir", os.path.join(out, "ck"),
            "--distributed-world-size", str(world), "--num-workers", "0", "--log-format", "none", "--fast-stat-sync",
            "--seed", "5", "--max-predictions-per-seq", "20", "--comm", comm, "--update-freq", str(update_freq),
            "--precision", a.precision] + list(extra)
    if not a.no_device_batch:
        argv.append("--device-batch-assembly")
    if graph:
        argv.append("--cuda-graph")
    else:
        argv.append("--no-cuda-graph")
    args = options.parse_training_args(argv)
    os.makedirs(args.save_dir, exist_ok=True)
    args.distributed_rank, args.device_id = (rank if world > 1 else 0), local
    torch.manual_seed(args.seed)
    task = tasks.LanguageModelingTask.setup_task(args)
    task.load_dataset("test")
    c = Controller(args, task, task.build_model(args))
    return args, c


def run(c, steps, update_freq, fault=None):
    from hetseq_9cme_b200.data import iterators
    done = 0
    epoch = 0
    while done < steps:
        it = c.get_train_iterator(epoch=epoch, load_dataset=(epoch == 0))
        c.lr_step(it.epoch)
        itr = iterators.GroupedIterator(it.next_epoch_itr(shuffle=False), update_freq)
        for samples in itr:
            if fault is not None:
                if fault[0] == int(os.environ["RANK"]) and done == fault[1]:
                    print("fault injection: rank %d exits before step %d" % fault, flush=True)
                    os._exit(17)
                try:
                    c.train_step(samples)
                    torch.cuda.synchronize()
                except RuntimeError as e:
                    print("WATCHDOG REPORT: %s  (%s)" % (c.watchdog_report(), str(e)[:80].replace("\n", " ")), flush=True)
                    os._exit(86)
            else:
                c.train_step(samples)
            done += 1
            if done >= steps:
                break
        epoch += 1
    c.flush_stats()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--mode", default="sharded", choices=["sharded", "replicated", "torch"])
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--update-freq", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--uneven", action="store_true")
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--compare-single", action="store_true")
    ap.add_argument("--checkpoint", action="store_true")
    ap.add_argument("--no-device-batch", action="store_true")
    ap.add_argument("--fault-inject", default=None, metavar="RANK:STEP", help="RANK exits before its STEP-th step; the others "
                    "must leave through the watchdog (launch the ranks as independent processes, not under torchrun)")
    a = ap.parse_args()
    out = a.out
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if a.mode == "replicated":
        os.environ["B200_NO_SHARDED"] = "1"
    from hetseq_9cme_b200.data import synthetic
    d = os.path.join(out, "data")
    if rank == 0:
        # batches of 8: --uneven gives world*U*k + 1 batches per epoch, so the last step leaves ranks without data
        per_file = 128 if not a.uneven else 4 * (world * a.update_freq * 3 + 1)
        synthetic.make_bert_shards(d, n_train_files=2, samples_per_file=per_file, seq_len=128, max_pred=20, vocab_size=1000, test_samples=8)
        synthetic.write_vocab(os.path.join(out, "v.txt"), 1000)
        synthetic.write_bert_config(os.path.join(out, "c.json"), **CFG)
    dist.barrier()
    comm = "torch" if a.mode == "torch" else "p2p"
    args, c = build(out, a, world, rank, local, a.update_freq, a.graph, comm)
    fault = tuple(int(x) for x in a.fault_inject.split(":")) if a.fault_inject else None
    run(c, a.steps, a.update_freq, fault)
    red = c.reducer
    want = "TorchReducer" if a.mode == "torch" else "P2PReducer"
    assert type(red).__name__ == want, type(red).__name__
    if a.mode != "torch":
        assert red.sharded == (a.mode == "sharded"), (red.sharded, a.mode)
        if red.debug and not a.graph:
            assert red.debug_checked >= a.steps * len(c.arena.buckets), red.debug_checked
    if a.graph:
        assert c._graph is not None, "no CUDA graph was captured"
    assert c.get_num_updates() == a.steps
    c.prepare_checkpoint()                      # sharded: gathers moments (+ masters in bf16 mode)
    opt = c.optimizer.optimizer
    for name, t in [("master", c.arena.master)] + [(n, opt._flat_state[n]) for n in opt._state_names] + \
                   ([("work", c.arena.work.float())] if c.arena.work is not None else []):
        s = t.double().sum().reshape(1)
        q = t.double().pow(2).sum().reshape(1)
        v = torch.cat([s, q, -s, -q])
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        assert v[0].item() == -v[2].item() and v[1].item() == -v[3].item(), "ranks diverged in " + name
    msg = "buckets %d, loss %.4f, gnorm %.4f" % (len(c.arena.buckets), c.get_meter("train_loss").avg, c.get_meter("gnorm").avg)

    if a.checkpoint:
        from hetseq_9cme_b200 import checkpoint_utils
        it = c.get_train_iterator(epoch=0, load_dataset=False)
        args.no_save = False
        checkpoint_utils.save_checkpoint(args, c, it, None)
        args2, c2 = build(out, a, world, rank, local, a.update_freq, a.graph, comm)
        extra, _ = checkpoint_utils.load_checkpoint(args2, c2)
        assert c2.get_num_updates() == a.steps, c2.get_num_updates()
        assert torch.equal(c2.arena.master, c.arena.master), "master weights changed through the checkpoint"
        o2 = c2.optimizer.optimizer
        for n in opt._state_names:
            assert torch.equal(o2._flat_state[n], opt._flat_state[n]), "optimizer state %s changed through the checkpoint" % n
        run(c2, 2, a.update_freq)
        msg += ", checkpoint round trip ok"

    if a.compare_single:
        ref_master = None
        if rank == 0:
            os.environ["B200_NO_SHARDED"] = "1"
            _, single = build(out, a, 1, 0, local, a.update_freq * world, False, "torch")
            run(single, a.steps, a.update_freq * world)
            x, y = c.arena.master, single.arena.master
            # Adam normalises every update to ~lr whatever the gradient's size, so compare the mean displacement per
            # tensor against the distance travelled rather than the worst element
            travelled = a.steps * 1e-3
            tol = 0.05 if a.precision == "tf32" else 0.25
            bad = []
            named = dict(single.get_model().named_parameters())
            pa = dict(c.get_model().named_parameters())
            for k, p in named.items():
                if k.endswith("key.bias"):
                    continue
                diff = (pa[k].master - p.master).abs().mean().item()
                if diff > tol * travelled:
                    bad.append((k, diff / travelled))
            assert not bad, "differs from the single-process run: %s" % bad[:5]
            gl = abs(c.get_meter("train_loss").avg - single.get_meter("train_loss").avg)
            assert gl < 2e-2 * abs(single.get_meter("train_loss").avg) + 1e-3, ("loss", gl)
            msg += ", matches the single-process run"
        dist.barrier()
    if rank == 0:
        print("DP DEBUG RUN PASS  (%s %s uf=%d %s: %s)" % (a.mode, "graph" if a.graph else "eager", a.update_freq, a.precision, msg), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
