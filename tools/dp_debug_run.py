#!/usr/bin/env python3
"""Run under torchrun: a few data-parallel training steps of a small BERT with the P2P reducer's debug verification on
(B200_P2P_DEBUG=1), then check that all ranks hold identical parameters.  Used by tests/test_gpu_multi.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    out = sys.argv[1]
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from hetseq_9cme_b200 import options, tasks
    from hetseq_9cme_b200.controller import Controller
    from hetseq_9cme_b200.data import synthetic
    d = os.path.join(out, "data")
    if rank == 0:
        synthetic.make_bert_shards(d, n_train_files=2, samples_per_file=128, seq_len=128, max_pred=20, vocab_size=1000, test_samples=8)
        synthetic.write_vocab(os.path.join(out, "v.txt"), 1000)
        synthetic.write_bert_config(os.path.join(out, "c.json"), vocab_size=1000, hidden_size=128, num_hidden_layers=2,
                                    num_attention_heads=2, intermediate_size=512, max_position_embeddings=128)
    dist.barrier()
    argv = ["--task", "bert", "--data", d, "--dict", os.path.join(out, "v.txt"), "--config_file", os.path.join(out, "c.json"),
            "--max-sentences", "8", "--lr", "1e-3", "--warmup-updates", "2", "--total-num-update", "50", "--clip-norm", "1.0",
            "--valid-subset", "test", "--save-dir", os.path.join(out, "ck"), "--distributed-world-size", str(world),
            "--num-workers", "0", "--log-format", "none", "--fast-stat-sync", "--seed", "5", "--max-predictions-per-seq", "20",
            "--comm", "p2p", "--device-batch-assembly"]
    args = options.parse_training_args(argv)
    args.distributed_rank, args.device_id = rank, local
    torch.manual_seed(args.seed)
    task = tasks.LanguageModelingTask.setup_task(args)
    task.load_dataset("test")
    c = Controller(args, task, task.build_model(args))
    it = c.get_train_iterator(epoch=0, load_dataset=True)
    c.lr_step(it.epoch)
    itr = it.next_epoch_itr(shuffle=True)
    for _ in range(8):
        c.train_step([next(itr)])
    c.flush_stats()
    red = c.reducer
    assert type(red).__name__ == "P2PReducer" and red.debug and red.debug_checked >= 8 * len(c.arena.buckets), (
        type(red).__name__, getattr(red, "debug_checked", None))
    s = c.arena.master.double().sum().reshape(1)
    lo, hi = s.clone(), s.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert lo.item() == hi.item(), "ranks diverged"
    if rank == 0:
        print("DP DEBUG RUN PASS  (buckets verified: %d, loss %.4f)" % (red.debug_checked, c.get_meter("train_loss").avg), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
