"""Native (sm_100a kernels) BERT forward/backward against the plain-PyTorch module tree with identical weights."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _models(cls, cfg_kwargs, extra=(), dtype=torch.float32):
    from hetseq_9cme_b200 import bert_modeling as bm
    from hetseq_9cme_b200.parallel.arena import ParamArena
    torch.manual_seed(0)
    cfg = bm.BertConfig(1000, **cfg_kwargs)
    ref = getattr(bm, cls)(cfg, *extra).cuda()
    # give LayerNorm / biases non-trivial values so their gradients are exercised
    with torch.no_grad():
        for n, p in ref.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.05)
    nat = copy.deepcopy(ref)
    if hasattr(nat, "cls"):
        nat.cls.predictions.decoder.weight = nat.bert.embeddings.word_embeddings.weight
    arena = ParamArena(list(nat.parameters()), groups=nat.arena_groups(), device="cuda", compute_dtype=dtype)
    nat.bind_arena(arena)
    ref.force_torch_path = True
    return cfg, ref, nat, arena


def _rel(a, b):
    return (a.double() - b.double()).abs().max().item() / (b.double().abs().max().item() + 1e-12)


def _check_grads(ref, nat, tol=3e-2):
    """Every parameter gradient within ``tol`` of the torch path, relative to that tensor's largest entry —
    with a floor tied to the largest gradient in the model, because some gradients are analytically zero
    (a key bias shifts all scores of a row equally, softmax ignores it) and only carry rounding noise."""
    rp = dict(ref.named_parameters())
    glob = max(p.grad.abs().max().item() for p in rp.values() if p.grad is not None)
    bad = []
    for n, p in nat.named_parameters():
        g_ref = rp[n].grad
        if g_ref is None:           # e.g. the pooler under a token-classification head
            continue
        err = (p.main_grad.double() - g_ref.double()).abs().max().item()
        denom = max(g_ref.abs().max().item(), 1e-3 * glob)
        # a handful of scalars (the 2-element NSP bias) is a sum of +/- terms that nearly cancel: its relative error
        # is several times that of the logits it comes from
        if err / denom > (3 * tol if p.numel() <= 8 else tol):
            bad.append((n, err / denom))
    assert not bad, bad


def _batch(B, L, V, pad_from=None):
    torch.manual_seed(1)
    ids = torch.randint(5, V, (B, L), device="cuda")
    tts = torch.randint(0, 2, (B, L), device="cuda")
    mask = torch.ones(B, L, dtype=torch.int64, device="cuda")
    if pad_from is not None:
        mask[0, pad_from:] = 0
    mlm = torch.full((B, L), -1, dtype=torch.int64, device="cuda")
    for b in range(B):
        pos = torch.randperm(L - 1, device="cuda")[: max(2, L // 8)] + 1
        mlm[b, pos] = torch.randint(0, V, (pos.numel(),), device="cuda")
    nsp = torch.randint(0, 2, (B,), device="cuda")
    return ids, tts, mask, mlm, nsp


@pytest.mark.parametrize("L", [128, 72])
def test_pretraining_loss_and_grads_match_torch(L):
    cfgk = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=128)
    cfg, ref, nat, arena = _models("BertForPreTraining", cfgk)
    ids, tts, mask, mlm, nsp = _batch(4, L, 1000, pad_from=L - 9)
    torch.backends.cuda.matmul.allow_tf32 = False
    ref.train(); nat.train()
    loss_ref = ref(ids, tts, mask, mlm, nsp)
    loss_ref.backward()
    loss_nat = nat(ids, tts, mask, mlm, nsp)
    loss_nat.backward()
    torch.cuda.synchronize()
    assert abs(loss_nat.item() - loss_ref.item()) / abs(loss_ref.item()) < 2e-3
    _check_grads(ref, nat)
    assert nat.last_mlm_count.tolist()[0] == nat.last_mlm_count.tolist()[1]


def test_token_classification_matches_torch():
    cfgk = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=4, intermediate_size=256,
                hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=128)
    cfg, ref, nat, arena = _models("BertForTokenClassification", cfgk, extra=(5,))
    B, L = 3, 47
    ids, tts, mask, _, _ = _batch(B, L, 1000, pad_from=30)
    labels = torch.randint(0, 5, (B, L), device="cuda")
    labels[:, 0] = -100
    torch.backends.cuda.matmul.allow_tf32 = False
    ref.train(); nat.train()
    lr = ref(ids, tts, mask, labels); lr.backward()
    ln = nat(ids, tts, mask, labels); ln.backward()
    assert abs(ln.item() - lr.item()) / abs(lr.item()) < 2e-3
    _check_grads(ref, nat)


def test_bf16_mode_runs_and_tracks_fp32():
    cfgk = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=128)
    cfg, ref, nat, arena = _models("BertForPreTraining", cfgk, dtype=torch.bfloat16)
    ids, tts, mask, mlm, nsp = _batch(4, 128, 1000)
    ref.train(); nat.train()
    loss_ref = ref(ids, tts, mask, mlm, nsp)
    loss_nat = nat(ids, tts, mask, mlm, nsp)
    loss_nat.backward()
    assert abs(loss_nat.item() - loss_ref.item()) / abs(loss_ref.item()) < 3e-2
    assert torch.isfinite(arena.grad).all() and arena.grad.abs().sum().item() > 0


def test_dropout_training_step_is_finite_and_seed_dependent():
    from hetseq_9cme_b200.ops import kernels as K
    cfgk = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1, max_position_embeddings=128)
    cfg, ref, nat, arena = _models("BertForPreTraining", cfgk)
    ids, tts, mask, mlm, nsp = _batch(4, 128, 1000)
    nat.train()
    ws = K.Workspace.get("cuda")
    ws.set_seed(1)
    l1 = nat(ids, tts, mask, mlm, nsp); l1.backward()
    g1 = arena.grad.clone(); arena.zero_grad()
    ws.set_seed(1)
    l1b = nat(ids, tts, mask, mlm, nsp); l1b.backward()
    g1b = arena.grad.clone(); arena.zero_grad()
    ws.set_seed(2)
    l2 = nat(ids, tts, mask, mlm, nsp)
    assert torch.isfinite(g1).all()
    assert l1.item() == l1b.item()                 # same seed -> same masks
    assert (g1 - g1b).abs().max().item() < 1e-3 * g1.abs().max().item()   # (atomics reorder fp32 adds)
    assert l1.item() != l2.item()


def test_activation_recompute_gives_identical_gradients():
    cfgk = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1, max_position_embeddings=128)
    from hetseq_9cme_b200.ops import kernels as K
    cfg, ref, nat, arena = _models("BertForPreTraining", cfgk)
    ids, tts, mask, mlm, nsp = _batch(4, 128, 1000)
    nat.train()
    ws = K.Workspace.get("cuda")
    ws.set_seed(9)
    l1 = nat(ids, tts, mask, mlm, nsp); l1.backward()
    g1 = arena.grad.clone(); arena.zero_grad()
    nat.checkpoint_activations = True
    ws.set_seed(9)
    l2 = nat(ids, tts, mask, mlm, nsp); l2.backward()
    assert l1.item() == l2.item()
    assert (arena.grad - g1).abs().max().item() <= 1e-3 * g1.abs().max().item()


# ---------------------------------------------------------------------------------------------------
# the shapes bench.py runs: H = 768, 12 heads, I = 3072, V = 30522, B * L = 4096 (one encoder layer + all heads)
# ---------------------------------------------------------------------------------------------------
BASE = dict(hidden_size=768, num_hidden_layers=1, num_attention_heads=12, intermediate_size=3072,
            hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=512)


def _base_models(dtype=torch.float32, vocab=30522):
    from hetseq_9cme_b200 import bert_modeling as bm
    from hetseq_9cme_b200.parallel.arena import ParamArena
    torch.manual_seed(0)
    cfg = bm.BertConfig(vocab, **BASE)
    ref = bm.BertForPreTraining(cfg).cuda()
    with torch.no_grad():
        for n, p in ref.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.05)
    nat = copy.deepcopy(ref)
    nat.cls.predictions.decoder.weight = nat.bert.embeddings.word_embeddings.weight
    arena = ParamArena(list(nat.parameters()), groups=nat.arena_groups(), device="cuda", compute_dtype=dtype)
    nat.bind_arena(arena)
    nat.max_predictions_per_seq = 20
    ref.force_torch_path = True
    return ref, nat, arena


def _base_batch(B, L, V, n_pred=20):
    g = torch.Generator(device="cuda").manual_seed(7)
    ids = torch.randint(5, V, (B, L), device="cuda", generator=g)
    tts = torch.randint(0, 2, (B, L), device="cuda", generator=g)
    mask = torch.ones(B, L, dtype=torch.int64, device="cuda")
    mask[: B // 4, L - L // 5:] = 0
    mlm = torch.full((B, L), -1, dtype=torch.int64, device="cuda")
    for b in range(B):
        pos = torch.randperm(L - 1, device="cuda", generator=g)[:n_pred] + 1
        mlm[b, pos] = torch.randint(0, V, (n_pred,), device="cuda", generator=g)
    nsp = torch.randint(0, 2, (B,), device="cuda", generator=g)
    return ids, tts, mask, mlm, nsp


@pytest.mark.parametrize("variant", ["default", "cta_pairs", "no_cta_pairs"])
def test_bert_base_shapes_tf32_match_fp32_torch(variant, monkeypatch):
    """One BERT-base layer + embeddings + MLM / NSP heads at the benchmark's sizes (4096 tokens, vocabulary 30522:
    decoder data gradient with K = 30522 and split-K, tied-embedding weight gradient with M = 30522 and a ragged last
    tile), TF32 kernels against the fp32 PyTorch path.  The variants force the GEMM's CTA-pair kernel on and off so
    that both sides of the dispatch heuristic are covered at these shapes."""
    if variant != "default":
        monkeypatch.setenv("B200_GEMM_CTA2", "1" if variant == "cta_pairs" else "0")
    ref, nat, arena = _base_models()
    ids, tts, mask, mlm, nsp = _base_batch(32, 128, 30522)
    torch.backends.cuda.matmul.allow_tf32 = False
    ref.train(); nat.train()
    loss_ref = ref(ids, tts, mask, mlm, nsp); loss_ref.backward()
    loss_nat = nat(ids, tts, mask, mlm, nsp); loss_nat.backward()
    torch.cuda.synchronize()
    assert abs(loss_nat.item() - loss_ref.item()) / abs(loss_ref.item()) < 5e-3
    _check_grads(ref, nat, tol=1e-2)
    used, found = nat.last_mlm_count.tolist()
    assert used == found == 32 * 20


def test_bert_base_shapes_bf16_gradients_track_fp32():
    """Same sizes in bf16 (fp32 master weights and gradient accumulators): every parameter gradient within bf16
    tolerance of the fp32 PyTorch gradient, measured against that tensor's largest entry."""
    ref, nat, arena = _base_models(dtype=torch.bfloat16)
    ids, tts, mask, mlm, nsp = _base_batch(32, 128, 30522)
    torch.backends.cuda.matmul.allow_tf32 = False
    ref.train(); nat.train()
    loss_ref = ref(ids, tts, mask, mlm, nsp); loss_ref.backward()
    loss_nat = nat(ids, tts, mask, mlm, nsp); loss_nat.backward()
    torch.cuda.synchronize()
    assert abs(loss_nat.item() - loss_ref.item()) / abs(loss_ref.item()) < 2e-2
    _check_grads(ref, nat, tol=6e-2)


def test_phase2_shapes_long_sequence_attention():
    """Sequence 512 with 80 labelled positions per sequence (phase 2): the long-sequence attention path, with ragged
    padding, against the fp32 PyTorch path."""
    ref, nat, arena = _base_models(vocab=4096)
    nat.max_predictions_per_seq = 80
    ids, tts, mask, mlm, nsp = _base_batch(4, 512, 4096, n_pred=80)
    torch.backends.cuda.matmul.allow_tf32 = False
    ref.train(); nat.train()
    loss_ref = ref(ids, tts, mask, mlm, nsp); loss_ref.backward()
    loss_nat = nat(ids, tts, mask, mlm, nsp); loss_nat.backward()
    torch.cuda.synchronize()
    assert abs(loss_nat.item() - loss_ref.item()) / abs(loss_ref.item()) < 5e-3
    _check_grads(ref, nat, tol=1e-2)


def test_tied_embedding_gradient_is_announced_last():
    """The decoder weight is the word-embedding table: the reducer may only be told that its gradient is final after the
    embedding backward has added the token-lookup part (otherwise that part is never summed across ranks)."""
    from hetseq_9cme_b200.models.bert_native import GradSink
    cfgk = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=128)
    cfg, ref, nat, arena = _models("BertForPreTraining", cfgk)

    class Recorder(GradSink):
        def __init__(self):
            self.calls = []

        def ready(self, params):
            self.calls.append([id(p) for p in params])

    rec = Recorder()
    nat.bind_arena(arena, rec)
    ids, tts, mask, mlm, nsp = _batch(4, 128, 1000)
    nat.train()
    nat(ids, tts, mask, mlm, nsp).backward()
    emb_id = id(nat.bert.embeddings.word_embeddings.weight)
    hits = [i for i, c in enumerate(rec.calls) if emb_id in c]
    assert hits == [len(rec.calls) - 1], (hits, len(rec.calls))
    every = {pid for c in rec.calls for pid in c}
    assert every == {id(p) for p in arena.params}


def test_strict_fp32_mode_matches_fp32_torch_tightly():
    """--precision fp32 on the native path: loss and gradients against the fp32 PyTorch path at fp32 tolerances (the
    TF32 policy needs 2e-3 / 3e-2 on the same model)."""
    from hetseq_9cme_b200.ops import gemm as G
    cfgk = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=128)
    cfg, ref, nat, arena = _models("BertForPreTraining", cfgk)
    ids, tts, mask, mlm, nsp = _batch(4, 128, 1000, pad_from=119)
    torch.backends.cuda.matmul.allow_tf32 = False
    ref.train(); nat.train()
    loss_ref = ref(ids, tts, mask, mlm, nsp); loss_ref.backward()
    G.set_strict_fp32(True)
    try:
        loss_nat = nat(ids, tts, mask, mlm, nsp); loss_nat.backward()
        torch.cuda.synchronize()
    finally:
        G.set_strict_fp32(False)
    assert abs(loss_nat.item() - loss_ref.item()) / abs(loss_ref.item()) < 2e-5
    _check_grads(ref, nat, tol=5e-4)
