"""DeepSpeedTransformerLayer vs a plain-PyTorch BERT layer with identical weights (dropout 0): forward and
per-parameter gradients (oracle strategy of reference tests/unit/test_cuda_forward.py / test_cuda_backward.py)."""
import pytest
import torch

from deeperspeed_b200.models.bert import BertConfig, BertLayer, BertLayerPreLN
from deeperspeed_b200.module_inject import replace_transformer_layer, revert_transformer_layer


def _case(batch, hidden, seq, heads, preln, device, dtype, atol, invertible=False, attn_dropout=0.0):
    torch.manual_seed(0)
    cfg = BertConfig(hidden_size=hidden, num_hidden_layers=2, num_attention_heads=heads, hidden_dropout_prob=0.0,
                     attention_probs_dropout_prob=0.0)
    cls = BertLayerPreLN if preln else BertLayer
    ref = torch.nn.Sequential(cls(cfg)).to(device)
    import copy
    model = copy.deepcopy(ref).to(device, dtype)
    model = replace_transformer_layer(cls, model, micro_batch_size=batch, bert_config=cfg, seed=1, max_seq_length=seq,
                                      preln=preln, fp16=dtype == torch.float16)
    ds_layer = model[0]
    assert type(ds_layer).__name__ == "DeepSpeedTransformerLayer"
    ds_layer.config.normalize_invertible = invertible
    # `use_flash = False` routes attention through the GEMM + softmax + dropout chain; the Philox kernel is then told to
    # keep everything so that the result stays comparable with the dropout-free reference
    ds_layer.config.attn_dropout_ratio = attn_dropout
    if attn_dropout > 0.0:
        ds_layer.use_flash = False
    x = (torch.randn(batch, seq, hidden, device=device) * 0.5)
    mask = torch.zeros(batch, 1, 1, seq, device=device)
    mask[:, :, :, -3:] = -10000.0
    xr = x.clone().requires_grad_()
    xd = x.clone().to(dtype).requires_grad_()
    ref.train(); ds_layer.train()
    yr = ref[0](xr, mask)
    yd = ds_layer(xd, mask.to(dtype))
    assert (yd.float() - yr).abs().max().item() < atol, (yd.float() - yr).abs().max().item()
    g = torch.randn_like(yr)
    yr.backward(g)
    yd.backward(g.to(dtype))

    def close(a, b, name):
        tol = atol * max(b.abs().mean().item(), 1e-3) * 8 + atol * 0.1
        assert (a.float() - b).abs().max().item() < tol, (name, (a.float() - b).abs().max().item(), tol)

    close(xd.grad, xr.grad, "dx")
    r = ref[0]
    qkv_w = torch.cat((r.attention.self.query.weight.grad, r.attention.self.key.weight.grad, r.attention.self.value.weight.grad))
    close(ds_layer.attn_qkvw.grad, qkv_w, "qkv_w")
    close(ds_layer.attn_ow.grad, r.attention.output.dense.weight.grad, "attn_ow")
    inter = r.intermediate.dense_act if preln else r.intermediate.dense
    close(ds_layer.inter_w.grad, inter.weight.grad, "inter_w")
    close(ds_layer.inter_b.grad, inter.bias.grad, "inter_b")
    close(ds_layer.output_w.grad, r.output.dense.weight.grad, "output_w")
    ln = r.PreAttentionLayerNorm if preln else r.output.LayerNorm
    close(ds_layer.norm_w.grad, ln.weight.grad, "norm_w")
    # revert gives back an equivalent torch layer
    back = revert_transformer_layer(cls, model, cfg, preln=preln)
    assert type(back[0]).__name__ == cls.__name__


@pytest.mark.parametrize("preln", [True, False])
def test_layer_cpu(preln):
    _case(2, 64, 24, 4, preln, "cpu", torch.float32, 2e-4)


# shapes of reference tests/unit/test_cuda_forward.py:200-235 / test_cuda_backward.py:259-270 (odd sequence lengths,
# small batches, head dims 64 / 80 / 128, hidden up to 8192)
_SHAPES = [(8, 1024, 128, 16, True), (8, 1024, 128, 16, False), (3, 1024, 119, 16, True), (8, 1600, 128, 25, True),
           (8, 160, 128, 2, False), (2, 2048, 512, 16, True), (8, 256, 52, 4, True), (3, 1024, 51, 16, True),
           (3, 1024, 54, 16, False), (8, 1024, 381, 16, True), (8, 1024, 384, 16, False), (8, 1024, 119, 16, True),
           (8, 1024, 509, 16, True), (3, 1024, 511, 16, False), (1, 1536, 128, 24, True), (1, 2560, 128, 40, False),
           (1, 128, 2048, 2, True), (1, 8192, 128, 64, True), (64, 1024, 56, 16, False), (64, 1024, 21, 16, True)]


@pytest.mark.gpu
@pytest.mark.parametrize("batch,hidden,seq,heads,preln", _SHAPES)
def test_layer_gpu(batch, hidden, seq, heads, preln):
    """Attention on the tcgen05 flash kernel (mask as key bias, head dims padded to 128, sequences padded to 128)."""
    _case(batch, hidden, seq, heads, preln, "cuda", torch.float16, 2e-2)


@pytest.mark.gpu
@pytest.mark.parametrize("batch,hidden,seq,heads,preln", [(8, 1024, 128, 16, True), (3, 1024, 119, 16, False),
                                                         (2, 2048, 512, 16, True)])
def test_layer_gpu_unfused_attention_chain(batch, hidden, seq, heads, preln):
    """The GEMM + masked softmax + dropout chain that serves attention dropout (keep-probability 1 - 1e-7 ~ identity)."""
    _case(batch, hidden, seq, heads, preln, "cuda", torch.float16, 2e-2, attn_dropout=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("batch,hidden,seq,heads", [(8, 1024, 128, 16), (3, 1024, 119, 16), (2, 2048, 512, 16)])
def test_layer_gpu_invertible_layernorm(batch, hidden, seq, heads):
    """normalize_invertible: LayerNorm backward from the layer OUTPUT; same gradients as the regular layer."""
    _case(batch, hidden, seq, heads, True, "cuda", torch.float16, 2e-2, invertible=True)


@pytest.mark.gpu
def test_layernorm_invertible_kernel_vs_fp32():
    from deeperspeed_b200.ops import kernels as K
    torch.manual_seed(3)
    x = torch.randn(300, 1024, device="cuda").half().requires_grad_()
    g = (torch.rand(1024, device="cuda") + 0.5).half().requires_grad_()
    b = (torch.randn(1024, device="cuda") * 0.1).half().requires_grad_()
    y = K.layer_norm_invertible(x, g, b, 1e-5)
    dy = torch.randn_like(y)
    y.backward(dy)
    xf, gf, bf = (t.detach().float().requires_grad_() for t in (x, g, b))
    yf = torch.nn.functional.layer_norm(xf, (1024,), gf, bf, 1e-5)
    yf.backward(dy.float())
    rel = lambda a, r: ((a.float() - r).abs().max() / r.abs().max()).item()
    assert rel(y, yf) < 2e-3 and rel(x.grad, xf.grad) < 1e-2 and rel(g.grad, gf.grad) < 1e-2 and rel(b.grad, bf.grad) < 1e-2


@pytest.mark.gpu
def test_cuda_graph_capture_of_neox_layer():
    """Forward + backward of a GPT-NeoX block replayed from CUDA graphs equals eager execution."""
    import copy
    from deeperspeed_b200.models.gpt_neox import NeoXConfig, ParallelTransformerLayerPipe
    from deeperspeed_b200.runtime.cuda_graphs import graph_module
    torch.manual_seed(0)
    cfg = NeoXConfig.tiny()
    layer = ParallelTransformerLayerPipe(cfg).cuda().half()
    eager = copy.deepcopy(layer)
    B, S, H = 2, 128, cfg.hidden_size
    sample = torch.randn(B, S, H, device="cuda", dtype=torch.float16, requires_grad=True)
    graphed = graph_module(layer, (sample,))
    for it in range(3):   # replays with fresh data
        x = torch.randn(B, S, H, device="cuda", dtype=torch.float16)
        g = torch.randn(B, S, H, device="cuda", dtype=torch.float16)
        xe, xg = x.clone().requires_grad_(), x.clone().requires_grad_()
        for p in list(layer.parameters()) + list(eager.parameters()):
            p.grad = None
        ye = eager(xe)
        ye.backward(g)
        yg = graphed(xg)
        yg.backward(g)
        assert torch.allclose(yg.float(), ye.float(), atol=2e-3, rtol=2e-3)
        assert torch.allclose(xg.grad.float(), xe.grad.float(), atol=2e-3, rtol=2e-3)
        for (n, a), b in zip(layer.named_parameters(), eager.parameters()):
            assert a.grad is not None and torch.allclose(a.grad.float(), b.grad.float(), atol=5e-3, rtol=5e-3), n


@pytest.mark.parametrize("preln", [False, True])
def test_module_injection_replace_and_revert_cpu(preln):
    """module_inject: every BERT layer of a model is swapped for a DeepSpeedTransformerLayer carrying the same weights
    (same outputs), and reverted to layers with the original weights (reference module_inject/replace_module.py,
    ops/module_inject.py:8-216)."""
    from deeperspeed_b200.models.bert import BertConfig, BertLayer, BertLayerPreLN
    from deeperspeed_b200.module_inject import replace_transformer_layer, revert_transformer_layer
    from deeperspeed_b200.ops.transformer import DeepSpeedTransformerLayer
    torch.manual_seed(0)
    bc = BertConfig(hidden_size=32, num_hidden_layers=2, num_attention_heads=4, hidden_dropout_prob=0.0,
                    attention_probs_dropout_prob=0.0, layer_norm_eps=1e-5)
    layer_cls = BertLayerPreLN if preln else BertLayer

    class Encoder(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.layers = torch.nn.ModuleList([layer_cls(bc) for _ in range(2)])

        def forward(self, x, mask):
            for l in self.layers:
                x = l(x, mask)
                x = x[0] if isinstance(x, tuple) else x
            return x

    model = Encoder().eval()
    original = {k: v.detach().clone() for k, v in model.state_dict().items()}
    x = torch.randn(2, 16, 32)
    mask = torch.zeros(2, 1, 1, 16)
    want = model(x, mask)
    replace_transformer_layer(layer_cls, model, micro_batch_size=2, bert_config=bc, seed=1, max_seq_length=16,
                              preln=preln, fp16=False)
    assert all(isinstance(l, DeepSpeedTransformerLayer) for l in model.layers)
    got = model.eval()(x, mask)
    assert (got - want).abs().max().item() < 2e-4, (got - want).abs().max().item()
    revert_transformer_layer(layer_cls, model, bc, preln=preln)
    assert all(isinstance(l, layer_cls) for l in model.layers)
    for k, v in model.state_dict().items():
        assert torch.equal(v, original[k]), k
    assert (model.eval()(x, mask) - want).abs().max().item() < 1e-6


def test_module_injection_into_a_hugging_face_bert_cpu():
    """A BERT built by the `transformers` package in this image: every BertLayer becomes a DeepSpeedTransformerLayer
    with the same weights; hidden states agree on the attended positions although that encoder hands its layers a
    BOOLEAN [B, 1, S, S] mask and extra keyword arguments (`past_key_values`, ...); reverting restores the original."""
    transformers = pytest.importorskip("transformers")
    import copy
    from transformers.models.bert.modeling_bert import BertLayer
    from deeperspeed_b200.module_inject import replace_transformer_layer, revert_transformer_layer
    from deeperspeed_b200.ops.transformer import DeepSpeedTransformerLayer
    torch.manual_seed(0)
    cfg = transformers.BertConfig(num_hidden_layers=2, num_attention_heads=2, hidden_size=32, intermediate_size=64,
                                  vocab_size=128, max_position_embeddings=32, hidden_dropout_prob=0.0,
                                  attention_probs_dropout_prob=0.0)
    model = transformers.BertModel(cfg).eval()
    ids = torch.randint(0, 128, (2, 16))
    keep = torch.ones(2, 16, dtype=torch.long)
    keep[1, 10:] = 0                                    # second sequence is padded
    want = model(input_ids=ids, attention_mask=keep).last_hidden_state
    swapped = replace_transformer_layer(BertLayer, copy.deepcopy(model), micro_batch_size=2, bert_config=cfg, seed=1,
                                        max_seq_length=16, preln=False, fp16=False, huggingface=False)
    assert all(isinstance(layer, DeepSpeedTransformerLayer) for layer in swapped.encoder.layer)
    got = swapped(input_ids=ids, attention_mask=keep).last_hidden_state
    assert torch.allclose(got[0], want[0], atol=1e-5) and torch.allclose(got[1, :10], want[1, :10], atol=1e-5)
    back = revert_transformer_layer(BertLayer, swapped, cfg, preln=False)
    assert all(isinstance(layer, BertLayer) for layer in back.encoder.layer)
    assert torch.allclose(back(input_ids=ids, attention_mask=keep).last_hidden_state, want, atol=1e-6)
