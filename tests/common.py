"""Fork-N-processes harness for distributed tests (role of reference tests/unit/common.py:16-104): gloo on CPU,
NCCL when CUDA devices are present; hung or failed workers become pytest failures."""
import os
import socket
import time
import traceback

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

DEEPSPEED_UNIT_WORKER_TIMEOUT = 180


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, args, kwargs, errq, use_cuda):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        os.environ["LOCAL_RANK"] = str(rank)
        os.environ["RANK"] = str(rank)
        os.environ["WORLD_SIZE"] = str(world)
        if not use_cuda:
            os.environ["CUDA_VISIBLE_DEVICES"] = ""
        # forked children must not touch the parent's OpenMP pool (it does not survive fork once the parent has run a
        # threaded op such as a conv): one thread => parallel regions run inline
        torch.set_num_threads(2 if use_cuda else 1)
        import deeperspeed_b200 as deepspeed
        # a test that initialised torch.distributed in the parent process leaves a group behind that fork copies;
        # its sockets and threads do not exist here, so it can neither be used nor torn down
        assert not dist.is_initialized(), "the parent process holds an initialised process group: run tests that " \
            "call init_distributed() in-process after the fork-based ones (or in their own pytest process)"
        deepspeed.init_distributed(dist_backend="nccl" if use_cuda else "gloo", verbose=False)
        if use_cuda:
            torch.cuda.set_device(rank)
        fn(*args, **kwargs)
        dist.barrier()
        dist.destroy_process_group()
    except Exception:
        errq.put((rank, traceback.format_exc()))
        raise


def distributed_test(world_size=2, use_cuda=False):
    def deco(fn):
        def run(*args, **kwargs):
            sizes = world_size if isinstance(world_size, (list, tuple)) else [world_size]
            for ws in sizes:
                if use_cuda and torch.cuda.device_count() < ws:
                    pytest.skip(f"needs {ws} GPUs")
                ctx = mp.get_context("spawn" if use_cuda else "fork")
                errq = ctx.Queue()
                port = _free_port()
                procs = [ctx.Process(target=_worker, args=(r, ws, port, fn, args, kwargs, errq, use_cuda)) for r in range(ws)]
                for p in procs:
                    p.start()
                deadline = time.time() + DEEPSPEED_UNIT_WORKER_TIMEOUT
                for p in procs:
                    p.join(max(0.0, deadline - time.time()))
                hung = [p for p in procs if p.is_alive()]
                for p in hung:
                    p.terminate()
                if not errq.empty():
                    rank, tb = errq.get()
                    pytest.fail(f"worker {rank} failed:\n{tb}", pytrace=False)
                if hung:
                    pytest.fail(f"{len(hung)} worker(s) hung", pytrace=False)
                for p in procs:
                    if p.exitcode != 0:
                        pytest.fail(f"worker exited with code {p.exitcode}", pytrace=False)
        run.__name__ = fn.__name__
        return run
    return deco


def run_distributed(fn, world_size=2, use_cuda=False, args=()):
    """Run a MODULE-LEVEL function on `world_size` ranks (spawn-safe: needed once CUDA is initialised in the parent)."""
    if use_cuda and torch.cuda.device_count() < world_size:
        pytest.skip(f"needs {world_size} GPUs")
    ctx = mp.get_context("spawn" if use_cuda else "fork")
    errq = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, fn, args, {}, errq, use_cuda)) for r in range(world_size)]
    for p in procs:
        p.start()
    deadline = time.time() + DEEPSPEED_UNIT_WORKER_TIMEOUT
    for p in procs:
        p.join(max(0.0, deadline - time.time()))
    hung = [p for p in procs if p.is_alive()]
    for p in hung:
        p.terminate()
    if not errq.empty():
        rank, tb = errq.get()
        pytest.fail(f"worker {rank} failed:\n{tb}", pytrace=False)
    if hung:
        pytest.fail(f"{len(hung)} worker(s) hung", pytrace=False)
    for p in procs:
        if p.exitcode != 0:
            pytest.fail(f"worker exited with code {p.exitcode}", pytrace=False)
