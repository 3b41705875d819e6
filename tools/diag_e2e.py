"""Where does the end-to-end step spend host time?  Replays bench.py's sequence (warm-up, device-resident loop,
pinned-host loop) with wall clocks around the phases of Controller._graph_train_step and reports the slowest step.

    python tools/diag_e2e.py [--steps 30] [--precision bf16]
"""
import os
import sys
import tempfile
import time

import torch

sys.path.insert(0, ".")
import bench                                            # noqa: E402


def main():
    a = bench.parse(sys.argv[1:])
    from hetseq_9cme_b200 import tasks
    from hetseq_9cme_b200.controller import Controller
    from hetseq_9cme_b200.data import synthetic
    from hetseq_9cme_b200.data.packed import PackedBertBatch
    torch.cuda.set_device(0)
    tmp = tempfile.mkdtemp(prefix="b200diag_")
    per_rank = (a.steps + a.warmup + 4) * a.batch
    synthetic.make_bert_shards(tmp, n_train_files=2, samples_per_file=(per_rank + 1) // 2 + a.batch, seq_len=a.seq_len,
                               max_pred=20, fmt="b2s", seed=1234, test_samples=a.batch)
    cfg = synthetic.write_bert_config(os.path.join(tmp, "c.json"), **bench.MODELS[a.model])
    vocab = synthetic.write_vocab(os.path.join(tmp, "v.txt"))
    args = bench.make_args(a, tmp, cfg, vocab, 0, 0, 1)
    torch.manual_seed(args.seed)
    task = tasks.LanguageModelingTask.setup_task(args)
    task.load_dataset("test")
    c = Controller(args, task, task.build_model(args))
    from hetseq_9cme_b200.data import iterators as _it
    make_log = []
    _orig_make = _it.BatchLoader._make

    def _timed_make(self, idxs):
        t = time.perf_counter()
        out = _orig_make(self, idxs)
        make_log.append((t, time.perf_counter(), min(idxs) if len(idxs) else -1, max(idxs) if len(idxs) else -1))
        return out
    _it.BatchLoader._make = _timed_make
    ep = c.get_train_iterator(epoch=0, load_dataset=True)
    c.lr_step(ep.epoch)
    itr = ep.next_epoch_itr(shuffle=True)
    staged = []
    for _ in range(a.warmup):
        s = next(itr)
        c.train_step([s])
        staged.append(PackedBertBatch([s[0].cuda(), s[1]]) if isinstance(s, PackedBertBatch) else [t.cuda() for t in s])
    c.flush_stats()
    torch.cuda.synchronize()
    for i in range(a.steps):
        c.train_step([staged[i % len(staged)]])
    c.flush_stats()
    torch.cuda.synchronize()

    # instrument the phases
    log = []
    cur = {}

    def wrap(obj, name, label):
        fn = getattr(obj, name)

        def timed(*args, **kw):
            t = time.perf_counter()
            try:
                return fn(*args, **kw)
            finally:
                cur[label] = cur.get(label, 0.0) + (time.perf_counter() - t) * 1e3
        setattr(obj, name, timed)

    wrap(c, "_set_seed", "set_seed")
    wrap(c, "_stage_hyper", "stage_hyper")
    wrap(c._graph, "replay", "replay")
    wrap(c._mailbox, "post", "post")
    wrap(c, "_consume_stats", "consume")
    wrap(c, "_settle_gc", "gc")
    wrap(c, "_graphable", "graphable")
    kept = []
    t0 = time.perf_counter()
    for i in range(a.steps):
        ta = time.perf_counter()
        s = next(itr)
        kept.append(PackedBertBatch([s[0].clone(), s[1]]) if isinstance(s, PackedBertBatch) else [t.clone() for t in s])
        tb = time.perf_counter()
        c.train_step([s])
        tc = time.perf_counter()
        cur["fetch"] = (tb - ta) * 1e3
        cur["total"] = (tc - ta) * 1e3
        cur["t_start"], cur["t_end"] = ta, tc
        log.append(dict(cur))
        cur.clear()
    c.flush_stats()
    torch.cuda.synchronize()
    tot = (time.perf_counter() - t0) * 1e3
    worst = max(range(len(log)), key=lambda i: log[i]["total"])
    med = sorted(x["total"] for x in log)[len(log) // 2]
    print("e2e %.3f ms/step, median step %.3f ms, graph=%s" % (tot / a.steps, med, c._graph is not None))
    for i in sorted({0, 1, max(worst - 1, 0), worst, min(worst + 1, len(log) - 1)}):
        print("step %2d: " % i + "  ".join("%s %.3f" % (k, v) for k, v in sorted(log[i].items()) if not k.startswith("t_")))
    w = log[worst]
    print("loader activity overlapping the slowest step (ms relative to its start):")
    for (m0, m1, lo, hi) in make_log:
        if m1 >= w["t_start"] - 0.005 and m0 <= w["t_end"] + 0.005:
            print("   _make rows %d..%d  start %+.3f  dur %.3f ms" % (lo, hi, (m0 - w["t_start"]) * 1e3, (m1 - m0) * 1e3))
    # is it the DATA of a batch?  replay single batches with a sync on both sides
    for i in sorted({0, max(worst - 2, 0), max(worst - 1, 0), worst}):
        ts = []
        for _ in range(3):
            torch.cuda.synchronize()
            t = time.perf_counter()
            c.train_step([kept[i]])
            c.flush_stats()
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t) * 1e3)
        b = kept[i]
        raw = b[0] if isinstance(b, PackedBertBatch) else b[0]
        print("batch %2d alone: %s ms   (distinct token ids %d, zeros %d)" % (i, " ".join("%.2f" % x for x in ts),
              int(raw[:, :a.seq_len].unique().numel()), int((raw[:, :a.seq_len] == 0).sum())))


if __name__ == "__main__":
    main()
