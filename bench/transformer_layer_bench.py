"""Secondary: DeepSpeedTransformerLayer (BERT-large: hidden 1024, 16 heads, FF 4096, fp16, pre-LN) forward + backward
against the plain-PyTorch BERT layer with the same weights (the oracle of the numerics tests; torch eager, fp16).
The reference's own transformer_op cannot be built for compute capability 10.0, so torch eager is the stand-in
baseline on this box; the reference's published figures (V100) are in BASELINE.md.
Shapes from the reference's BERT-large recipe: batch 64 x seq 128 and batch 16 x seq 512; dropout 0.1."""
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeperspeed_b200.models.bert import BertConfig, BertLayerPreLN
from deeperspeed_b200.module_inject import replace_transformer_layer

dev = "cuda"
dt = torch.float16
LAYERS = 4


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


res = []
for batch, seq in ((64, 128), (16, 512)):
    torch.manual_seed(0)
    cfg = BertConfig(hidden_size=1024, num_hidden_layers=LAYERS, num_attention_heads=16, intermediate_size=4096,
                     hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1)
    ref = torch.nn.Sequential(*[BertLayerPreLN(cfg) for _ in range(LAYERS)]).to(dev, dt)
    ours = replace_transformer_layer(BertLayerPreLN, copy.deepcopy(ref), micro_batch_size=batch, bert_config=cfg, seed=1,
                                     max_seq_length=seq, preln=True, fp16=True)
    x = torch.randn(batch, seq, 1024, device=dev, dtype=dt)
    mask = torch.zeros(batch, 1, 1, seq, device=dev, dtype=dt)
    g = torch.randn(batch, seq, 1024, device=dev, dtype=dt)

    def step(model):
        xin = x.clone().requires_grad_()
        h = xin
        for layer in model:
            h = layer(h, mask)
        h.backward(g)

    ref.train()
    ours.train()
    t_ref = timed(lambda: step(ref))
    t_ours = timed(lambda: step(ours))
    # launch-bound: replay forward and backward of the whole stack from CUDA graphs (dropout off, see cuda_graphs.py)
    from deeperspeed_b200.runtime.cuda_graphs import graph_module
    cfg0 = BertConfig(hidden_size=1024, num_hidden_layers=LAYERS, num_attention_heads=16, intermediate_size=4096,
                      hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    ref0 = torch.nn.Sequential(*[BertLayerPreLN(cfg0) for _ in range(LAYERS)]).to(dev, dt)
    ours0 = replace_transformer_layer(BertLayerPreLN, ref0, micro_batch_size=batch, bert_config=cfg0, seed=1,
                                      max_seq_length=seq, preln=True, fp16=True)
    ours0.train()
    t_ours0 = timed(lambda: step(ours0))

    class Stack(torch.nn.Module):
        def __init__(self, layers):
            super().__init__()
            self.layers = layers

        def forward(self, h, m):
            for layer in self.layers:
                h = layer(h, m)
            return h
    graphed = graph_module(Stack(ours0), (x.clone().requires_grad_(), mask))

    def step_g():
        xin = x.clone().requires_grad_()
        graphed(xin, mask).backward(g)
    t_graph = timed(step_g)
    tokens = batch * seq
    # 12 h^2 weights per layer (QKV 3h^2 + out h^2 + FF 8h^2), 6 FLOPs per weight per token, + attention 12 S h
    flops = LAYERS * tokens * (6 * 12 * 1024 * 1024 + 12 * seq * 1024)
    res.append({"batch": batch, "seq": seq, "layers": LAYERS, "torch_eager_ms": round(t_ref, 3), "ours_ms": round(t_ours, 3),
                "ours_nodropout_ms": round(t_ours0, 3), "ours_nodropout_cuda_graph_ms": round(t_graph, 3),
                "speedup": round(t_ref / t_ours, 2), "ours_tflops": round(flops / t_ours / 1e9, 1),
                "ours_samples_per_s_per_24_layers": round(batch / (t_ours * 24 / LAYERS) * 1e3, 1)})
    print(res[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/transformer_layer_bench.json", "w"), indent=1)
