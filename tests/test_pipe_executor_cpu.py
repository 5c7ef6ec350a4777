"""The pipeline executor's own pieces: plan compilation, edge-split slot layout and its numerical equivalence with the
plain schedule (gloo, 2 and 3 stages)."""
import pytest
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from deeperspeed_b200.runtime.pipe import schedule
from deeperspeed_b200.runtime.pipe.executor import compile_plan, slot_rows, SlotFeed
from deeperspeed_b200.runtime.pipe.topology import PipeDataParallelTopology, ProcessTopology

from common import distributed_test
from simple_model import LinearStackPipe, args_from_dict, random_dataset

HID = 16


def test_compile_plan_folds_each_tick():
    plan = compile_plan(schedule.TrainSchedule(micro_batches=4, stages=2, stage_id=0))
    assert len(plan) == 2 * (4 + 2 - 1)
    assert plan[0].load == 0 and plan[0].compute == ("fwd", 0) and not plan[0].recvs and not plan[0].sends
    # tick 3 of stage 0: activation of micro-batch 1 goes down while the gradient of micro-batch 0 comes up
    assert plan[3].sends == (("act", 1),) and plan[3].recvs == (("grad", 0),) and plan[3].compute == ("bwd", 0)
    assert plan[-1].tail == ("reduce_tied", "reduce", "step")
    last = compile_plan(schedule.TrainSchedule(micro_batches=4, stages=2, stage_id=1))
    assert last[1].recvs == (("act", 0),) and last[1].load == 0 and last[1].compute == ("fwd", 0)
    # every send of a tick has its receive in the same tick of the neighbour
    for a, b in zip(plan, last):
        assert [k for k, _ in a.sends if k == "act"] == [k for k, _ in b.recvs if k == "act"]
        assert [k for k, _ in b.sends if k == "grad"] == [k for k, _ in a.recvs if k == "grad"]


def test_compile_plan_rejects_unknown_instruction():
    class Bogus(schedule.PipeInstruction):
        pass

    with pytest.raises(RuntimeError):
        compile_plan([[Bogus()]])


def test_slot_rows_and_feed():
    assert slot_rows(4, 4, 1) == [4, 4, 4, 4]
    assert slot_rows(4, 4, 2) == [2, 2, 4, 4, 2, 2]
    assert slot_rows(1, 4, 2) == [4]                 # a single micro-batch is not split
    assert slot_rows(4, 3, 2) == [3, 3, 3, 3]        # rows not divisible
    rows = slot_rows(3, 4, 2)
    feed = SlotFeed(rows, 4)
    batches = iter([(torch.arange(4) + 10 * i, torch.arange(4) - 10 * i) for i in range(3)])
    got = [feed.next(lambda: next(batches)) for _ in rows]
    assert [g[0].tolist() for g in got] == [[0, 1], [2, 3], [10, 11, 12, 13], [20, 21], [22, 23]]
    assert got[3][1].tolist() == [-20, -19]


def test_topology_is_stride_arithmetic():
    topo = ProcessTopology(axes=["pipe", "data", "model"], dims=[2, 3, 2])
    assert topo.world_size() == 12 and topo.get_rank(pipe=1, data=2, model=1) == 11
    assert topo.get_coord(7) == topo.ProcessCoord(pipe=1, data=0, model=1)
    assert topo.get_axis_comm_lists("data") == [[0, 2, 4], [1, 3, 5], [6, 8, 10], [7, 9, 11]]
    assert topo.get_axis_comm_lists("model")[0] == [0, 1] and topo.get_axis_comm_lists("nope") == []
    assert topo.filter_match(pipe=0, model=1) == [1, 3, 5] and topo.get_axis_list("pipe", 1) == list(range(6, 12))
    assert topo.get_rank_repr(rank=3) == "model_01" and topo.mapping[topo.ProcessCoord(0, 1, 1)] == 3
    with pytest.raises(ValueError):
        topo.get_rank(pipe=0)


def _train_losses(tmp_path, stages, edge_split, steps=3):
    cfg = {"train_batch_size": 16, "train_micro_batch_size_per_gpu": 4, "steps_per_print": 100,
           "optimizer": {"type": "Adam", "params": {"lr": 0.01}}, "fp16": {"enabled": True, "type": "fp32"},
           "zero_optimization": {"stage": 1}, "pipeline": {"edge_split": edge_split}}
    torch.manual_seed(3)
    pipe = LinearStackPipe(HID, HID, HID, 4, topology=PipeDataParallelTopology(num_pp=stages, num_dp=1),
                           partition_method="uniform")
    engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=pipe,
                                           model_parameters=[p for p in pipe.parameters()])
    assert engine.edge_split == edge_split
    it = iter(deepspeed.utils.RepeatingLoader(torch.utils.data.DataLoader(random_dataset(64, HID, seed=11), batch_size=4)))
    losses = [float(engine.train_batch(it)) for _ in range(steps)]
    return losses, [p.detach().clone() for p in pipe.parameters()]


@pytest.mark.parametrize("stages", [2, 3])
def test_edge_split_matches_plain_schedule(stages, tmp_path):
    """Feeding the first and last micro-batch as two half-size slices changes the bubble, not the mathematics."""
    @distributed_test(world_size=stages)
    def run():
        plain, w0 = _train_losses(tmp_path, stages, 1)
        split, w1 = _train_losses(tmp_path, stages, 2)
        for a, b in zip(plain, split):
            assert abs(a - b) < 1e-5, (plain, split)
        for a, b in zip(w0, w1):
            assert torch.allclose(a, b, atol=1e-5)
    run()


def test_compute_partitioning_ignores_lookup_tables(tmp_path):
    """`partition_method="compute"`: an embedding costs nothing, the output projection costs what its weight says, so a
    decoder with a large vocabulary gets more blocks on its first stage than `parameters` would give it."""
    @distributed_test(world_size=2)
    def run():
        from deeperspeed_b200.models.gpt_neox import GPTNeoXPipe, NeoXConfig
        cfg = NeoXConfig.tiny()
        cfg.num_layers, cfg.vocab_size, cfg.hidden_size = 6, 4096, 64
        topo = PipeDataParallelTopology(num_pp=2, num_dp=1)
        by_compute = GPTNeoXPipe(cfg, topology=topo, partition_method="compute")
        by_params = GPTNeoXPipe(cfg, topology=topo, partition_method="parameters")
        # specs: embedding, 6 blocks, final norm, head. head = 4096 * 64 = 262144 ~ 5.3 blocks of 49k parameters
        assert by_params.parts[1] < by_compute.parts[1]
        first, last = by_compute.parts[1] - 1, 6 - (by_compute.parts[1] - 1)
        assert first > last
    run()


def test_comm_stub_skips_transfers_without_desynchronising(tmp_path):
    """utils/comm_stub (the bench's exposed-communication measurement): with the stub on, a pipeline step runs to the
    end without any activation / gradient transfer and without the ZeRO collectives; switching it off again leaves the
    ranks in step (same number of steps everywhere), and training goes on."""
    @distributed_test(world_size=2)
    def run():
        from deeperspeed_b200.utils import comm_stub
        from deeperspeed_b200.runtime.pipe import p2p
        cfg = {"train_batch_size": 16, "train_micro_batch_size_per_gpu": 4, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 0.01}}, "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 1}}
        torch.manual_seed(3)
        pipe = LinearStackPipe(HID, HID, HID, 4, topology=PipeDataParallelTopology(num_pp=2, num_dp=1),
                               partition_method="uniform")
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=pipe,
                                               model_parameters=[p for p in pipe.parameters()])
        it = iter(deepspeed.utils.RepeatingLoader(torch.utils.data.DataLoader(random_dataset(64, HID, seed=11), batch_size=4)))
        first = float(engine.train_batch(it))                    # real step: meta handshake and buffers are set up
        posted = []
        real_post = p2p.Link.post
        p2p.Link.post = lambda self, tensor: posted.append(comm_stub.enabled()) or real_post(self, tensor)
        try:
            assert comm_stub.set_enabled(True) is False
            for _ in range(2):
                engine.train_batch(it)                           # must not hang although nothing is exchanged
            assert posted and all(posted)                        # every post happened under the stub ...
            links = p2p.links()
            link = links.act_out or links.act_in
            assert link.post(torch.zeros(1)) is comm_stub.DONE   # ... and a post under the stub transfers nothing
            assert comm_stub.set_enabled(False) is True
        finally:
            p2p.Link.post = real_post
            comm_stub.set_enabled(False)
        dist.barrier()
        after = float(engine.train_batch(it))                    # transfers are back and both stages still agree
        assert after == after and first == first
        t = torch.tensor([engine.global_steps], dtype=torch.float32)
        dist.all_reduce(t)
        assert int(t.item()) == 2 * engine.global_steps == 8
    run()


def test_pipeline_engine_data_plumbing_api(tmp_path):
    """set_dataloader / set_dataiterator / set_batch_fn / inference_batch / set_has_attention_mask / the log helpers
    (reference pipe/engine.py:380-470,560-650): train_batch() without an iterator uses the one installed, the batch
    function sees every micro-batch of the stages that load data, inference_batch returns the last stage's output on
    every stage and equals the unpipelined model, and the engine can go on training afterwards."""
    import deeperspeed_b200 as deepspeed
    from simple_model import LinearStackPipe, args_from_dict, random_dataset

    @distributed_test(world_size=2)
    def run():
        hid = 16
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 1,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}}
        torch.manual_seed(3)
        model = LinearStackPipe(input_dim=hid, hidden_dim=hid, output_dim=hid, num_layers=2, num_stages=2)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=[p for p in model.parameters() if p.requires_grad])
        data = random_dataset(32, hid)
        calls = []

        def batch_fn(batch):
            calls.append(batch[0].shape[0])
            return batch[0] * 1.0, batch[1]

        engine.set_batch_fn(batch_fn)
        engine.set_dataloader(deepspeed.utils.RepeatingLoader(torch.utils.data.DataLoader(data, batch_size=2)))
        first = engine.train_batch()                           # no iterator argument: the installed loader is used
        assert calls == [2, 2]                                 # both stages load data (inputs / labels), 2 micro-batches
        engine.set_dataiterator(iter(deepspeed.utils.RepeatingLoader(torch.utils.data.DataLoader(data, batch_size=2))))
        assert engine.training_dataloader is None
        second = engine.train_batch()
        assert torch.isfinite(first) and torch.isfinite(second) and len(calls) == 4 and engine.global_steps == 2

        engine.set_has_attention_mask(True)
        assert engine.has_attention_mask is True
        with pytest.raises(AssertionError):
            engine.set_has_attention_mask("yes")
        engine.set_has_attention_mask(False)

        # inference: one micro-batch through both stages; every stage gets the logits
        engine.set_batch_fn(None)
        x, y = data[:2]
        logits, presents = engine.inference_batch(iter([(x, y)]))
        assert presents is None and logits.shape == (2, hid)
        pieces = [None, None]
        dist.all_gather_object(pieces, {k: v.detach().clone() for k, v in model.state_dict().items()})
        whole = {}
        for piece in pieces:
            whole.update(piece)
        h = x
        for idx in sorted({int(k.split(".")[0]) for k in whole}):
            h = torch.nn.functional.linear(h, whole[f"{idx}.weight"], whole.get(f"{idx}.bias"))
        assert torch.allclose(logits, h, atol=1e-5)
        other = logits.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(other, logits)                      # same tensor on both stages

        it = iter(deepspeed.utils.RepeatingLoader(torch.utils.data.DataLoader(data, batch_size=2)))
        assert torch.isfinite(engine.train_batch(data_iter=it)) and engine.global_steps == 3
        engine.tput_log("throughput line")
        engine.log_for_device("device line")
        engine.mem_status("memory line")

    run()
