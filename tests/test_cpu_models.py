"""Model-level parity on CPU: state-dict keys and losses against the reference BERT file, datasets, collators,
NER pipeline, and a tiny BERT pre-training run through the CLI."""
import math
import os
import sys
import types

import numpy as np
import pytest
import torch

REF = "/root/reference"
CFG = dict(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, intermediate_size=128, max_position_embeddings=64,
           hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)


def _ref_bert():
    if not os.path.isdir(os.path.join(REF, "hetseq")):
        pytest.skip("reference mount not present")
    for stub in ("h5py", "boto3", "botocore", "botocore.exceptions"):
        if stub not in sys.modules:
            m = types.ModuleType(stub)
            if stub == "botocore.exceptions":
                m.ClientError = type("ClientError", (Exception,), {})
            sys.modules[stub] = m
    if REF not in sys.path:
        sys.path.append(REF)
        sys.path.append(os.path.join(REF, "hetseq"))
    try:
        import hetseq.bert_modeling as rb
        return rb
    except Exception as e:
        pytest.skip("reference bert_modeling not importable: %r" % (e,))


def test_state_dict_keys_and_losses_match_reference():
    from hetseq_9cme_b200 import bert_modeling as bm
    rb = _ref_bert()
    torch.manual_seed(0)
    ours = bm.BertForPreTraining(bm.BertConfig(500, **CFG))
    theirs = rb.BertForPreTraining(rb.BertConfig(500, **CFG))
    assert list(ours.state_dict().keys()) == list(theirs.state_dict().keys())
    for (k, a), (_, b) in zip(ours.state_dict().items(), theirs.state_dict().items()):
        assert a.shape == b.shape, k
    theirs.load_state_dict(ours.state_dict())
    ours.eval(); theirs.eval()
    B, L = 3, 32
    ids = torch.randint(1, 500, (B, L)); tts = torch.randint(0, 2, (B, L))
    mask = torch.ones(B, L, dtype=torch.long); mask[1, 20:] = 0
    mlm = torch.full((B, L), -1); mlm[:, 3:9] = ids[:, 3:9]
    nsp = torch.tensor([0, 1, 0])
    assert ours(ids, tts, mask, mlm, nsp).item() == pytest.approx(theirs(ids, tts, mask, mlm, nsp).item(), rel=2e-5)
    ours_tc = bm.BertForTokenClassification(bm.BertConfig(500, **CFG), 5)
    theirs_tc = rb.BertForTokenClassification(rb.BertConfig(500, **CFG), 5)
    assert list(ours_tc.state_dict().keys()) == list(theirs_tc.state_dict().keys())
    theirs_tc.load_state_dict(ours_tc.state_dict()); ours_tc.eval(); theirs_tc.eval()
    lab = torch.randint(0, 5, (B, L)); lab[:, 0] = -100
    assert ours_tc(ids, tts, mask, lab).item() == pytest.approx(theirs_tc(ids, tts, mask, lab).item(), rel=2e-5)
    for name in ("BertForMaskedLM", "BertForNextSentencePrediction", "BertForQuestionAnswering"):
        a = getattr(bm, name)(bm.BertConfig(500, **CFG)); b = getattr(rb, name)(rb.BertConfig(500, **CFG))
        assert list(a.state_dict().keys()) == list(b.state_dict().keys()), name
    a = bm.BertForSequenceClassification(bm.BertConfig(500, **CFG), 3); b = rb.BertForSequenceClassification(rb.BertConfig(500, **CFG), 3)
    assert list(a.state_dict().keys()) == list(b.state_dict().keys())
    a = bm.BertForMultipleChoice(bm.BertConfig(500, **CFG), 4); b = rb.BertForMultipleChoice(rb.BertConfig(500, **CFG), 4)
    assert list(a.state_dict().keys()) == list(b.state_dict().keys())


def test_bert_config_json_roundtrip(tmp_path):
    from hetseq_9cme_b200 import bert_modeling as bm
    from hetseq_9cme_b200.data import synthetic
    p = synthetic.write_bert_config(str(tmp_path / "c.json"), hidden_size=128, num_hidden_layers=3)
    c = bm.BertConfig.from_json_file(p)
    assert c.hidden_size == 128 and c.num_hidden_layers == 3 and c.vocab_size == 30522
    assert bm.BertConfig.from_dict(c.to_dict()).to_json_string() == c.to_json_string()


def test_shard_formats_and_dataset_contract(tmp_path):
    from hetseq_9cme_b200.data import BertH5pyData, ConBertH5pyData, shard_io, synthetic
    d = str(tmp_path)
    synthetic.make_bert_shards(d, n_train_files=2, samples_per_file=10, seq_len=32, max_pred=5, vocab_size=2000, fmt="b2s", test_samples=4)
    shards = [BertH5pyData(os.path.join(d, "train_shard_%d.b2s" % i)) for i in range(2)]
    ds = ConBertH5pyData(shards)
    assert len(ds) == 20 and ds.num_tokens(3) == 512 and list(ds.ordered_indices()) == list(range(20))
    raw = shard_io.open_shard(os.path.join(d, "train_shard_1.b2s"))
    ids, seg, mask, lab, nsl = ds[13]                                  # global 13 -> shard 1, row 3
    assert torch.equal(ids, torch.from_numpy(np.asarray(raw["input_ids"][3]).astype(np.int64)))
    assert ids.dtype == seg.dtype == mask.dtype == lab.dtype == nsl.dtype == torch.int64
    pos = np.asarray(raw["masked_lm_positions"][3]); mids = np.asarray(raw["masked_lm_ids"][3])
    k = int(np.argmax(pos == 0)) if (pos == 0).any() else len(pos)
    want = torch.full((32,), -1, dtype=torch.int64); want[torch.from_numpy(pos[:k].astype(np.int64))] = torch.from_numpy(mids[:k].astype(np.int64))
    assert torch.equal(lab, want)
    batch = ds.fetch_batch([8, 9, 10, 11])                             # straddles the shard boundary
    coll = ds.collater([ds[i] for i in (8, 9, 10, 11)])
    for a, b in zip(batch, coll):
        assert torch.equal(a, b)
    assert ds.collater([]) is None
    # hdf5 round trip through the built-in minimal reader/writer
    synthetic.make_bert_shards(d + "/h5", n_train_files=1, samples_per_file=6, seq_len=32, max_pred=5, vocab_size=2000, fmt="hdf5", test_samples=2)
    h = BertH5pyData(os.path.join(d, "h5", "train_shard_0.hdf5"))
    b2 = synthetic.bert_shard_arrays(6, 32, 5, 2000, 0)
    assert torch.equal(h[2][0], torch.from_numpy(b2["input_ids"][2].astype(np.int64)))


def test_collators_pad_to_longest():
    from hetseq_9cme_b200.data_collator import YD_DataCollatorForELClassification, YD_DataCollatorForTokenClassification
    feats = [{"input_ids": [2, 5, 6, 3], "labels": [-100, 1, 0, -100], "token_type_ids": [0] * 4, "attention_mask": [1] * 4},
             {"input_ids": torch.tensor([2, 7, 3]), "labels": torch.tensor([-100, 2, -100]), "token_type_ids": [0] * 3, "attention_mask": [1] * 3}]
    out = YD_DataCollatorForTokenClassification(tokenizer=None)(feats)
    assert out["input_ids"].tolist() == [[2, 5, 6, 3], [2, 7, 3, 0]] and out["labels"].tolist()[1] == [-100, 2, -100, -100]
    assert out["attention_mask"].tolist()[1] == [1, 1, 1, 0] and out["input_ids"].dtype == torch.int64
    for f in feats:
        f["entity_labels"] = [0] * len(f["labels"])
    out = YD_DataCollatorForELClassification(tokenizer=None, pad_to_multiple_of=8)(feats)
    assert out["entity_labels"].shape == (2, 8) and out["entity_labels"][1, 3:].tolist() == [-100] * 5


def test_ner_and_bert_cli_runs_on_cpu(tmp_path):
    from hetseq_9cme_b200 import options, train
    from hetseq_9cme_b200.data import synthetic
    d = str(tmp_path)
    paths = synthetic.make_conll(d, n_train=48, n_valid=8, n_test=8, vocab_words=60)
    cfg = synthetic.write_bert_config(os.path.join(d, "c.json"), vocab_size=65, **CFG)
    argv = ["--task", "BertForTokenClassification", "--optimizer", "adam", "--data", d, "--dict", paths["vocab"], "--config_file", cfg,
            "--train_file", paths["train"], "--validation_file", paths["validation"], "--test_file", paths["test"],
            "--cpu", "--max-sentences", "8", "--lr", "1e-3", "--valid-subset", "test", "--save-dir", d + "/ck", "--max-epoch", "2",
            "--distributed-world-size", "1", "--num-workers", "0", "--log-format", "none", "--warmup-updates", "2", "--total-num-update", "100"]
    c = train.main(options.parse_training_args(argv))
    assert c.get_num_updates() == 12 and c.get_model().num_labels == 9
    first = torch.load(d + "/ck/checkpoint1.pt", weights_only=False)
    assert "tokenizer" not in vars(first["args"]) and "bert.encoder.layer.0.attention.self.query.weight" in first["model"]
    # warm-start a fresh NER model from that checkpoint (the --hetseq_state_dict path)
    c2 = train.main(options.parse_training_args(argv + ["--hetseq_state_dict", d + "/ck/checkpoint1.pt", "--save-dir", d + "/ck2", "--max-epoch", "1"]))
    assert c2.get_num_updates() == 6
    # BERT pre-training, seq 32
    synthetic.make_bert_shards(d + "/pt", n_train_files=2, samples_per_file=16, seq_len=32, max_pred=5, vocab_size=65, test_samples=8)
    synthetic.write_vocab(d + "/v.txt", 65)
    argv = ["--task", "bert", "--data", d + "/pt", "--dict", d + "/v.txt", "--config_file", cfg, "--cpu", "--max-sentences", "8",
            "--lr", "1e-3", "--valid-subset", "test", "--save-dir", d + "/ck3", "--max-epoch", "2", "--distributed-world-size", "1",
            "--num-workers", "0", "--log-format", "none", "--update-freq", "2", "--fast-stat-sync"]
    c3 = train.main(options.parse_training_args(argv))
    assert c3.get_num_updates() == 4 and os.path.exists(d + "/ck3/checkpoint_best.pt")


def test_seqeval_lite_and_ner_eval(tmp_path):
    from hetseq_9cme_b200 import seqeval_lite as se
    yt = [["O", "B-PER", "I-PER", "O", "B-LOC"], ["B-ORG", "O"]]
    yp = [["O", "B-PER", "I-PER", "O", "O"], ["B-ORG", "B-LOC"]]
    assert se.get_entities(yt[0]) == {("PER", 1, 2), ("LOC", 4, 4)}
    assert se.precision_score(yt, yp) == pytest.approx(2 / 3) and se.recall_score(yt, yp) == pytest.approx(2 / 3)
    assert se.f1_score(yt, yp) == pytest.approx(2 / 3) and se.accuracy_score(yt, yp) == pytest.approx(5 / 7)
    # end-to-end: train a tiny tagger for a few steps, evaluate the checkpoint
    from hetseq_9cme_b200 import eval_bert_fine_tuning_ner, options, train
    from hetseq_9cme_b200.data import synthetic
    d = str(tmp_path)
    paths = synthetic.make_conll(d, n_train=32, n_valid=8, n_test=8, vocab_words=40)
    cfg = synthetic.write_bert_config(os.path.join(d, "c.json"), vocab_size=45, **CFG)
    argv = ["--task", "BertForTokenClassification", "--data", d, "--dict", paths["vocab"], "--config_file", cfg,
            "--train_file", paths["train"], "--test_file", paths["test"], "--cpu", "--max-sentences", "8", "--lr", "1e-3",
            "--valid-subset", "test", "--save-dir", d + "/ck", "--max-epoch", "1", "--distributed-world-size", "1",
            "--num-workers", "0", "--log-format", "none"]
    train.main(options.parse_training_args(argv))
    out = eval_bert_fine_tuning_ner.evaluate(d + "/ck/checkpoint_last.pt", cfg, paths["vocab"], paths["test"], device="cpu",
                                             return_logits=True)
    logits = out.pop("logits")
    assert set(out) == {"accuracy_score", "precision", "recall", "f1"} and 0.0 <= out["accuracy_score"] <= 1.0
    # cross-implementation check (reference: test/TRANFORMERS_test_eval_bert_fine_tuning.py): the same checkpoint under
    # Hugging Face's BertForTokenClassification gives the same logits, hence the same scores
    pytest.importorskip("transformers")
    hf = eval_bert_fine_tuning_ner.evaluate(d + "/ck/checkpoint_last.pt", cfg, paths["vocab"], paths["test"], device="cpu",
                                            impl="transformers", return_logits=True)
    for a_, b_ in zip(logits, hf.pop("logits")):
        assert torch.allclose(a_, b_, atol=2e-4, rtol=1e-3), (a_ - b_).abs().max()
    assert hf == out


def test_transformers_task_variant_steps_through_the_controller(tmp_path):
    """hetseq/transformers_tasks.py:395-477 counterpart: a Hugging Face BertForTokenClassification trained by this
    engine (arena, fused step): the loss goes down and a checkpoint is written."""
    pytest.importorskip("transformers")
    from hetseq_9cme_b200 import options
    from hetseq_9cme_b200.controller import Controller
    from hetseq_9cme_b200.data import synthetic
    from hetseq_9cme_b200.transformers_tasks import TransformersBertForTokenClassificationTask
    d = str(tmp_path)
    paths = synthetic.make_conll(d, n_train=64, n_valid=8, n_test=8, vocab_words=40)
    cfg = synthetic.write_bert_config(os.path.join(d, "c.json"), vocab_size=45, **CFG)
    argv = ["--task", "BertForTokenClassification", "--data", d, "--dict", paths["vocab"], "--config_file", cfg,
            "--train_file", paths["train"], "--test_file", paths["test"], "--cpu", "--max-sentences", "8", "--lr", "3e-3",
            "--valid-subset", "test", "--save-dir", d + "/ck", "--distributed-world-size", "1", "--num-workers", "0",
            "--log-format", "none"]
    args = options.parse_training_args(argv)
    torch.manual_seed(0)
    task = TransformersBertForTokenClassificationTask.setup_task(args)
    model = task.build_model(args)
    assert type(model).__module__.startswith("transformers")
    c = Controller(args, task, model)
    it = c.get_train_iterator(epoch=0, load_dataset=True)
    c.lr_step(it.epoch)
    losses = []
    for epoch in range(3):
        for s in it.next_epoch_itr(shuffle=False):
            out = c.train_step([s])
            losses.append(out["loss"])
    c.flush_stats()
    assert c.get_num_updates() == 24 and losses[-1] < 0.7 * losses[1]
    c.save_checkpoint(d + "/hf.pt", {"train_iterator": it.state_dict()})
    assert any(k.startswith("bert.encoder.layer.0.") for k in torch.load(d + "/hf.pt", weights_only=False)["model"])


def test_file_utils_and_legacy_loader(tmp_path):
    from hetseq_9cme_b200 import file_utils
    from hetseq_9cme_b200.data import synthetic
    from hetseq_9cme_b200.data.BERT_DATA import CombineBertData
    p = tmp_path / "a.bin"; p.write_bytes(b"x")
    assert file_utils.cached_path(str(p)) == str(p) and file_utils.cached_path("file://" + str(p)) == str(p)
    with pytest.raises(EnvironmentError):
        file_utils.cached_path("https://s3.amazonaws.com/models/bert.tar.gz", cache_dir=str(tmp_path))
    name = file_utils.url_to_filename("https://x/y", "etag")
    (tmp_path / name).write_bytes(b"z")
    assert file_utils.cached_path("https://x/y", cache_dir=str(tmp_path)).endswith(name)
    d = synthetic.make_bert_shards(str(tmp_path / "s"), n_train_files=2, samples_per_file=5, seq_len=16, max_pred=3, vocab_size=100, test_samples=2)
    files = sorted(os.path.join(d, f) for f in os.listdir(d) if "train" in f)
    ds = CombineBertData(files)
    assert len(ds) == 10 and len(ds[3]) == 6 and ds[3][0].shape == (16,)


def test_packed_batches_assemble_to_the_same_tensors(tmp_path):
    """--device-batch-assembly ships one packed int32 matrix; assembling it (host restatement of the kernel) must give
    exactly the five tensors the per-sample path builds (ref: hetseq/data/h5pyDataset.py:33-48)."""
    from hetseq_9cme_b200.data import synthetic
    from hetseq_9cme_b200.data.h5pyDataset import BertH5pyData, ConBertH5pyData
    from hetseq_9cme_b200.data.packed import PackedBertBatch, assemble_host
    from hetseq_9cme_b200.utils import pin_sample, sample_is_empty
    d = str(tmp_path / "pk")
    synthetic.make_bert_shards(d, n_train_files=2, samples_per_file=24, seq_len=32, max_pred=6, vocab_size=500, test_samples=4)
    files = sorted(os.path.join(d, f) for f in os.listdir(d) if "train" in f)
    ds = ConBertH5pyData([BertH5pyData(f) for f in files])
    idx = [3, 30, 7, 41, 0, 25]                                  # spans both shards, unsorted
    plain = ds.fetch_batch(idx)
    ds.packed = True
    packed = ds.fetch_batch(idx)
    assert isinstance(packed, PackedBertBatch) and packed.raw.dtype == torch.int32
    assert packed.raw.shape == (len(idx), 3 * 32 + 2 * 6 + 1) and packed.dims == (32, 6)
    assert not sample_is_empty(packed) and isinstance(pin_sample(packed) if torch.cuda.is_available() else packed, PackedBertBatch)
    got = assemble_host(packed.raw, *packed.dims)
    for a, b in zip(got, plain):
        assert a.dtype == torch.int64 and torch.equal(a, b)
    # semantics of the scatter: stop at the first zero position, last duplicate wins
    raw = torch.zeros(1, 3 * 8 + 2 * 4 + 1, dtype=torch.int32)
    raw[0, 24:28] = torch.tensor([5, 5, 0, 3])
    raw[0, 28:32] = torch.tensor([11, 12, 13, 14])
    labels = assemble_host(raw, 8, 4)[3][0].tolist()
    assert labels == [-1, -1, -1, -1, -1, 12, -1, -1]


def test_entity_linking_task_trains_end_to_end(tmp_path):
    """`--task BertForELClassification` (hetseq/tasks/bert_for_el_classification_task.py, hetseq/model/
    bert_for_EL_classification.py; the reference's smoke script is test/test_entity_vocab/test_CONLL_2003_EL.py): CoNLL
    files with an entity column, entity vectors from --ent_vecs_filename, the joint token-classification + cosine
    entity loss through train.main; both losses are finite and fall, the frozen entity table stays untouched."""
    from hetseq_9cme_b200 import options, train
    from hetseq_9cme_b200.data import synthetic
    d = str(tmp_path)
    paths = synthetic.make_conll(d, n_train=64, n_valid=8, n_test=8, vocab_words=60, entities=5, entity_dim=12)
    cfg = synthetic.write_bert_config(os.path.join(d, "c.json"), vocab_size=65, **CFG)
    vecs = torch.load(paths["ent_vecs"])
    argv = ["--task", "BertForELClassification", "--optimizer", "adam", "--data", d, "--dict", paths["vocab"], "--config_file", cfg,
            "--train_file", paths["train"], "--validation_file", paths["validation"], "--test_file", paths["test"],
            "--ent_vecs_filename", paths["ent_vecs"], "--cpu", "--max-sentences", "8", "--lr", "3e-3", "--valid-subset", "test",
            "--save-dir", d + "/ck", "--max-epoch", "1", "--distributed-world-size", "1", "--num-workers", "0", "--log-format", "none"]
    torch.manual_seed(0)
    c = train.main(options.parse_training_args(argv))
    model = c.get_model()
    assert type(model).__name__ == "BertForELClassification" and c.get_num_updates() == 8
    ner0, ent0 = (float(x) for x in model.last_losses)
    c = train.main(options.parse_training_args(argv[:argv.index("--max-epoch") + 1] + ["4"] + argv[argv.index("--max-epoch") + 2:]))
    ner1, ent1 = (float(x) for x in c.get_model().last_losses)
    assert all(math.isfinite(v) for v in (ner0, ent0, ner1, ent1)) and ner1 < ner0 and ent1 < ent0
    sd = torch.load(d + "/ck/checkpoint_last.pt", weights_only=False)["model"]
    assert torch.equal(sd["entity_emb.weight"], vecs) and sd["entity_classifier.weight"].shape == (12, CFG["hidden_size"])
    feats = c.task.dataset("train")[0] if hasattr(c.task.dataset("train"), "__getitem__") else None
    assert feats is not None and "entity_labels" in feats and max(feats["entity_labels"]) <= 5


@pytest.mark.parametrize("compress,chunk_rows", [(False, 1024), (True, 4), (True, 1024)])
def test_hdf5_lite_round_trip(tmp_path, compress, chunk_rows):
    """The built-in HDF5 reader / writer (data/hdf5_lite.py — the image has no h5py / libhdf5): contiguous and chunked +
    gzip layouts, the integer widths NVIDIA's BERT shards use, a chunk size that does not divide the row count."""
    from hetseq_9cme_b200.data import hdf5_lite
    rng = np.random.RandomState(0)
    arrs = {"input_ids": rng.randint(0, 30000, (37, 128)).astype(np.int32),
            "input_mask": rng.randint(0, 2, (37, 128)).astype(np.int8),
            "next_sentence_labels": rng.randint(0, 2, (37,)).astype(np.int8),
            "masked_lm_ids": rng.randint(0, 30000, (37, 20)).astype(np.int64)}
    path = str(tmp_path / "shard.hdf5")
    hdf5_lite.write_hdf5(path, arrs, compress=compress, chunk_rows=chunk_rows)
    back = hdf5_lite.read_hdf5(path)
    assert set(back) == set(arrs)
    for k, v in arrs.items():
        assert back[k].dtype == v.dtype and np.array_equal(back[k], v), k
    assert open(path, "rb").read(8) == b"\x89HDF\r\n\x1a\n"           # a real HDF5 signature: h5py / h5dump can open it


def test_file_cache_and_shard_conversion(tmp_path):
    """`file_utils.cached_path` (hetseq/file_utils.py:78-105 without the download half: local paths, file:// URLs, a
    pre-populated cache; remote URLs that are not cached fail with a message instead of reaching for the network) and
    `shard_io.convert_to_b2shard` (HDF5 shard -> memory-mappable shard with identical content)."""
    from hetseq_9cme_b200 import file_utils
    from hetseq_9cme_b200.data import shard_io, synthetic
    d = str(tmp_path)
    local = os.path.join(d, "archive.bin")
    open(local, "wb").write(b"x")
    assert file_utils.cached_path(local) == local and file_utils.cached_path("file://" + local) == local
    with pytest.raises(EnvironmentError):
        file_utils.cached_path(os.path.join(d, "missing.bin"))
    url = "https://example.org/bert-base-uncased.tar.gz"
    cache = os.path.join(d, "cache")
    os.makedirs(cache)
    with pytest.raises(EnvironmentError, match="never downloads"):
        file_utils.cached_path(url, cache_dir=cache)
    name = file_utils.url_to_filename(url, etag="abc")
    open(os.path.join(cache, name), "wb").write(b"y")
    import json
    json.dump({"url": url, "etag": "abc"}, open(os.path.join(cache, name + ".json"), "w"))
    assert file_utils.cached_path(url, cache_dir=cache) == os.path.join(cache, name)
    assert file_utils.filename_to_url(name, cache_dir=cache) == (url, "abc")
    with pytest.raises(ValueError):
        file_utils.cached_path("ftp://example.org/x")

    synthetic.make_bert_shards(d + "/h5", n_train_files=1, samples_per_file=6, seq_len=32, max_pred=5, vocab_size=2000, fmt="hdf5", test_samples=2)
    src = os.path.join(d, "h5", "train_shard_0.hdf5")
    dst = os.path.join(d, "train_shard_0.b2s")
    shard_io.convert_to_b2shard(src, dst)
    a, b = shard_io.open_shard(src), shard_io.open_shard(dst)
    assert set(a.keys()) == set(b.keys())
    for k in a.keys():
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), k
