"""init_distributed with env / MPI discovery (reference deepspeed/utils/distributed.py:12-142).

B200-native difference: the backend follows the device — `nccl` when CUDA is present, `gloo` on CPU-only hosts — so
the whole engine is runnable on CPU for CI (the reference hard-codes nccl)."""
import os
from datetime import timedelta

import torch
import torch.distributed as dist

from ..runtime.constants import TORCH_DISTRIBUTED_DEFAULT_PORT, DEFAULT_PROCESS_GROUP_TIMEOUT_MINUTES
from .logging import logger

default_pg_timeout = timedelta(minutes=DEFAULT_PROCESS_GROUP_TIMEOUT_MINUTES)


def default_backend():
    return "nccl" if torch.cuda.is_available() else "gloo"


def init_distributed(dist_backend=None, auto_mpi_discovery=True, distributed_port=TORCH_DISTRIBUTED_DEFAULT_PORT,
                     verbose=True, timeout=default_pg_timeout, init_method=None):
    if dist_backend is None or (dist_backend == "nccl" and not torch.cuda.is_available()):
        dist_backend = default_backend()
    required_env = ["RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "LOCAL_RANK"]
    if auto_mpi_discovery and not all(e in os.environ for e in required_env):
        if in_aml() or in_mpi():
            if verbose:
                logger.info("Not using the DeepSpeed or torch.distributed launchers, attempting to detect MPI environment...")
            if in_aml() and not in_dlts():
                patch_aml_env_for_torch_nccl_backend(verbose=verbose)
            else:
                mpi_discovery(distributed_port=distributed_port, verbose=verbose)
        elif not dist.is_initialized():
            # single-process convenience (tests, notebooks)
            os.environ.setdefault("RANK", "0")
            os.environ.setdefault("LOCAL_RANK", "0")
            os.environ.setdefault("WORLD_SIZE", "1")
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", str(distributed_port))
    if not dist.is_initialized():
        if verbose:
            logger.info("Initializing torch distributed with backend: {}".format(dist_backend))
        kwargs = {}
        if dist_backend == "nccl" and torch.cuda.is_available():
            local_rank = int(os.environ.get("LOCAL_RANK", 0))
            torch.cuda.set_device(local_rank)
            try:
                kwargs["device_id"] = torch.device("cuda", local_rank)
            except Exception:
                pass
        dist.init_process_group(backend=dist_backend, timeout=timeout, init_method=init_method, **kwargs)


def in_mpi():
    return "OMPI_COMM_WORLD_RANK" in os.environ or "PMI_RANK" in os.environ or "MV2_COMM_WORLD_RANK" in os.environ


def mpi_discovery(distributed_port=TORCH_DISTRIBUTED_DEFAULT_PORT, verbose=True):
    """Derive RANK/WORLD_SIZE/LOCAL_RANK/MASTER_* from mpi4py, or from OpenMPI/PMI/MVAPICH env when mpi4py is absent."""
    try:
        from mpi4py import MPI
        import subprocess
        comm = MPI.COMM_WORLD
        rank, world_size = comm.Get_rank(), comm.Get_size()
        master_addr = None
        if rank == 0:
            result = subprocess.check_output("hostname -I", shell=True)
            master_addr = result.decode("utf-8").split()[0]
        master_addr = comm.bcast(master_addr, root=0)
        proc_name = MPI.Get_processor_name()
        all_procs = comm.allgather(proc_name)
        local_rank = sum(i == proc_name for i in all_procs[:rank])
    except ImportError:
        env = os.environ
        rank = int(env.get("OMPI_COMM_WORLD_RANK", env.get("PMI_RANK", env.get("MV2_COMM_WORLD_RANK", 0))))
        world_size = int(env.get("OMPI_COMM_WORLD_SIZE", env.get("PMI_SIZE", env.get("MV2_COMM_WORLD_SIZE", 1))))
        local_rank = int(env.get("OMPI_COMM_WORLD_LOCAL_RANK", env.get("MV2_COMM_WORLD_LOCAL_RANK", 0)))
        master_addr = env.get("MASTER_ADDR", "127.0.0.1")
    os.environ["RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world_size)
    os.environ["LOCAL_RANK"] = str(local_rank)
    os.environ["MASTER_ADDR"] = master_addr
    os.environ["MASTER_PORT"] = str(distributed_port)
    if verbose:
        logger.info("Discovered MPI settings of world_rank={}, local_rank={}, world_size={}, master_addr={}, "
                    "master_port={}".format(rank, local_rank, world_size, master_addr, distributed_port))
    if dist.is_initialized():
        assert dist.get_rank() == rank and dist.get_world_size() == world_size


def in_aml():
    return "AZUREML_EXPERIMENT_ID" in os.environ


def in_dlts():
    return "DLTS_JOB_ID" in os.environ


def patch_aml_env_for_torch_nccl_backend(master_port=6105, verbose=True):
    """AzureML OpenMPI env -> torch.distributed env (reference utils/distributed.py:108-142)."""
    os.environ["RANK"] = os.environ["OMPI_COMM_WORLD_RANK"]
    os.environ["WORLD_SIZE"] = os.environ["OMPI_COMM_WORLD_SIZE"]
    single_node = int(os.environ["OMPI_COMM_WORLD_LOCAL_SIZE"]) == int(os.environ["WORLD_SIZE"])
    if not single_node:
        master_node_params = os.environ["AZ_BATCH_MASTER_NODE"].split(":")
        os.environ["MASTER_ADDR"] = master_node_params[0]
        os.environ.setdefault("MASTER_PORT", str(master_port))
    else:
        os.environ["MASTER_ADDR"] = os.environ["AZ_BATCHAI_MPI_MASTER_NODE"]
        os.environ["MASTER_PORT"] = "54965"
    os.environ["NCCL_SOCKET_IFNAME"] = "^docker0,lo"
    os.environ["LOCAL_RANK"] = os.environ["OMPI_COMM_WORLD_LOCAL_RANK"]
    if verbose:
        logger.info("NCCL_SOCKET_IFNAME original value = {}".format(os.environ["NCCL_SOCKET_IFNAME"]))
