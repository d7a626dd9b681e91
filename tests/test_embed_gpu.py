"""The fused BERTweet / RoBERTa embedding block (csrc/embed.cu through ops.embed) against the Hugging Face module it
replaces -- `bertweet.embeddings` is a transformers RobertaEmbeddings (src/in_loop_train.py:440,486-496) -- and the
patchify kernel against the einops pattern of meant/meant.py:193-195."""
import pytest
import torch

from meant_b200 import ops

pytestmark = pytest.mark.gpu


def _block(vocab=1000, hidden=768, max_pos=130):
    from transformers import RobertaConfig
    from transformers.models.roberta.modeling_roberta import RobertaEmbeddings
    cfg = RobertaConfig(vocab_size=vocab, hidden_size=hidden, max_position_embeddings=max_pos, type_vocab_size=1,
                        pad_token_id=1, layer_norm_eps=1e-5, hidden_dropout_prob=0.1)
    torch.manual_seed(0)
    m = RobertaEmbeddings(cfg).cuda()
    with torch.no_grad():
        m.LayerNorm.weight.add_(torch.randn_like(m.LayerNorm.weight) * 0.1)
        m.LayerNorm.bias.add_(torch.randn_like(m.LayerNorm.bias) * 0.1)
        for e in (m.word_embeddings, m.position_embeddings, m.token_type_embeddings):
            e.weight.normal_(0, 1.0)           # larger than the 0.02 init so that LayerNorm statistics matter
    return m


@pytest.mark.parametrize("n,s,hidden", [(6, 128, 768), (3, 77, 768), (2, 300, 256), (1, 5, 1024)])
@pytest.mark.parametrize("bf16", [False, True])
def test_embedding_block_matches_hf_module(n, s, hidden, bf16):
    m = _block(hidden=hidden, max_pos=max(130, s + 4)).eval()
    torch.manual_seed(n * 100 + s)
    ids = torch.randint(2, 1000, (n, s), device="cuda")
    for r in range(n):                            # padded tails (and one interior pad) exercise the position scan
        cut = int(torch.randint(0, s // 4 + 1, (1,)))
        if cut:
            ids[r, s - cut:] = 1
    ids[0, min(3, s - 1)] = 1
    probe = torch.randn(n, s, hidden, device="cuda")
    ref = m(ids)
    (ref * probe).sum().backward()
    want = {k: p.grad.clone() for k, p in m.named_parameters()}
    m.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
        out = ops.embed(m, ids)
    assert out.dtype == (torch.bfloat16 if bf16 else torch.float32)
    (out.float() * probe).sum().backward()
    tol = 2e-2 if bf16 else 2e-5
    assert float((out.detach().float() - ref.detach()).abs().max()) <= tol * max(1.0, float(ref.detach().abs().max()))
    for k, p in m.named_parameters():
        a, b = p.grad.double().flatten(), want[k].double().flatten()
        cos = float(torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-300))
        assert cos >= (0.999 if bf16 else 0.99999), (k, cos)
        assert abs(float(a.norm() / b.norm()) - 1.0) <= (3e-2 if bf16 else 1e-4), k
    assert float(m.word_embeddings.weight.grad[1].abs().max()) == 0.0      # padding_idx row stays frozen


def test_embedding_block_dropout_is_inverted_dropout():
    """Dropout probabilities are realised in 1/256 steps (csrc/common.cuh drop_mask4): p = 0.1 -> 26/256, survivors scaled
    by 256/230 so that the expectation is unchanged."""
    m = _block().train()
    ids = torch.randint(2, 1000, (4, 64), device="cuda")
    torch.manual_seed(1)
    out = ops.embed(m, ids)
    ref = m.eval()(ids)
    kept = out != 0
    keep_p = 230.0 / 256.0
    assert abs(float(kept.float().mean()) - keep_p) < 0.01
    assert torch.allclose(out[kept], ref[kept] / keep_p, rtol=1e-4, atol=1e-5)


def test_embed_rejects_ids_on_another_device():
    m = _block()
    with pytest.raises(RuntimeError):
        ops.embed(m, torch.randint(2, 1000, (2, 8)))


@pytest.mark.parametrize("b,c,hw,p,dtype", [(3, 4, 224, 16, torch.float32), (2, 3, 64, 16, torch.float32), (5, 4, 32, 8, torch.bfloat16)])
def test_patchify_matches_the_einops_pattern(b, c, hw, p, dtype):
    torch.manual_seed(b)
    img = torch.randn(b, c, hw, hw, device="cuda").to(dtype).requires_grad_(True)
    ref_in = img.detach().clone().requires_grad_(True)
    ref = ref_in.reshape(b, c, hw // p, p, hw // p, p).permute(0, 2, 4, 3, 5, 1).reshape(b, (hw // p) ** 2, p * p * c)
    out = ops.patchify(img, p)
    assert torch.equal(out, ref)
    probe = torch.randn_like(ref)
    (ref * probe).sum().backward()
    (out * probe).sum().backward()
    assert torch.equal(img.grad, ref_in.grad)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        assert torch.equal(ops.patchify(img.detach().float(), p), ref.detach().float().bfloat16())
