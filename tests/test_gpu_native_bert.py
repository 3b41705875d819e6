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
        if err / denom > tol:
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
