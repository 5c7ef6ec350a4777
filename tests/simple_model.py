"""Tiny fixtures (role of reference tests/unit/simple_model.py)."""
import argparse
import json
import os

import torch

import deeperspeed_b200 as deepspeed
from deeperspeed_b200.pipe import PipelineModule, LayerSpec


class SimpleModel(torch.nn.Module):
    def __init__(self, hidden_dim, empty_grad=False, rank=0):
        super().__init__()
        self.linear = torch.nn.Linear(hidden_dim, hidden_dim)
        if empty_grad:
            self.linear2 = torch.nn.Linear(hidden_dim, hidden_dim)
        self.cross_entropy_loss = torch.nn.CrossEntropyLoss()
        self.rank = rank
        self.empty_grad = empty_grad

    def forward(self, x, y):
        hidden = self.linear(x)
        if self.rank == 0 and self.empty_grad:
            hidden = self.linear2(hidden) + hidden
        return self.cross_entropy_loss(hidden.float(), y)


class LinearStack(torch.nn.Module):
    def __init__(self, input_dim=128, hidden_dim=128, output_dim=128, num_layers=4):
        super().__init__()
        self.input_layer = torch.nn.Linear(input_dim, hidden_dim)
        self.layers = torch.nn.ModuleList([torch.nn.Linear(hidden_dim, hidden_dim, bias=False) for _ in range(num_layers)])
        self.output_layer = torch.nn.Linear(hidden_dim, output_dim)
        self.cross_entropy_loss = torch.nn.CrossEntropyLoss()

    def forward(self, x, y):
        x = self.input_layer(x)
        for layer in self.layers:
            x = layer(x)
        x = self.output_layer(x)
        return self.cross_entropy_loss(x.float(), y)


class LinearStackPipe(PipelineModule):
    def __init__(self, input_dim=128, hidden_dim=128, output_dim=128, num_layers=4, **kwargs):
        layers = [LayerSpec(torch.nn.Linear, input_dim, hidden_dim)]
        for _ in range(num_layers):
            layers.append(LayerSpec(torch.nn.Linear, hidden_dim, hidden_dim, bias=False))
            layers.append(lambda x: x)
        layers.append(LayerSpec(torch.nn.Linear, hidden_dim, output_dim))
        super().__init__(layers=layers, loss_fn=lambda out, y: torch.nn.functional.cross_entropy(out.float(), y), **kwargs)


def random_dataset(total_samples, hidden_dim, device="cpu", dtype=torch.float32, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(total_samples, hidden_dim, generator=g).to(dtype)
    y = torch.randint(0, hidden_dim, (total_samples,), generator=g)
    return torch.utils.data.TensorDataset(x.to(device), y.to(device))


def random_dataloader(model, total_samples, hidden_dim, device, dtype=torch.float32, seed=0):
    batch_size = model.train_micro_batch_size_per_gpu()
    ds = random_dataset(total_samples, hidden_dim, device, dtype, seed)
    return torch.utils.data.DataLoader(ds, batch_size=batch_size)


def args_from_dict(tmpdir, config_dict):
    config_path = os.path.join(str(tmpdir), "temp_config_{}_{}.json".format(os.environ.get("RANK", 0), os.getpid()))
    with open(config_path, "w") as fd:
        json.dump(config_dict, fd)
    parser = argparse.ArgumentParser()
    args = parser.parse_args(args="")
    args.deepspeed = True
    args.deepspeed_config = config_path
    args.local_rank = int(os.environ.get("LOCAL_RANK", 0))
    return args
