"""Multi-node transports: how the per-node launcher (or the user script itself) is started on every host.

Command shapes per reference deepspeed/launcher/multinode_runner.py:14-292: pdsh / ssh fan-out,
OpenMPI `mpirun`, MVAPICH2-GDR `mpirun`, Slurm `srun`, and the MosaicML platform. A transport is described by a few
class attributes (the binary that proves it is installed, how exported variables are spelled, which runner options it
cannot honour) plus one `get_cmd`; the argv pieces shared by several transports are built once in the base class.
"""
import os
import shutil
import subprocess
import sys
import warnings
from abc import ABC, abstractmethod

from ..utils.logging import logger
from .constants import PDSH_MAX_FAN_OUT, MVAPICH_TMP_HOSTFILE

LAUNCH_MODULE = "deeperspeed_b200.launcher.launch"


class MultiNodeRunner(ABC):
    probe = None               # executable whose presence means the transport is installed
    label = None               # value of `name` (launcher option / log messages)
    unsupported = ()           # (runner attribute, neutral value): options this transport cannot honour

    def __init__(self, args, world_info_base64, resource_pool=None):
        self.args = args
        self.user_arguments = self.parse_user_args()
        self.user_script = args.user_script
        self.world_info_base64 = world_info_base64
        self.resource_pool = resource_pool
        self.exports = {}

    def backend_exists(self):
        """Return whether the corresponding backend exists"""
        return shutil.which(self.probe)

    @abstractmethod
    def get_cmd(self, environment, active_resources):
        """Return the command to execute on node"""

    def add_export(self, key, var):
        self.exports[key.strip()] = var.strip()

    def parse_user_args(self):
        return self.args.user_args

    @property
    def name(self):
        return self.label or self.__class__.__name__

    # ---- shared argv pieces
    def _reject_unsupported(self):
        for attr, neutral in self.unsupported:
            what = {"include": "worker include/exclusion", "exclude": "worker include/exclusion",
                    "num_nodes": "limiting num nodes", "num_gpus": "limiting num gpus"}[attr]
            assert getattr(self.args, attr) == neutral, f"{self.name} backend does not support {what}"

    def _total_processes(self):
        return sum(self.resource_pool.values())

    def _script_argv(self):
        """python -u <user script> <user args>: transports that start every rank themselves."""
        return [sys.executable, "-u", self.user_script] + self.user_arguments

    def _launcher_argv(self, node_rank, master_addr=None, master_port=None):
        """python -u -m <per-node launcher> ...: transports that start one process per node."""
        return [sys.executable, "-u", "-m", LAUNCH_MODULE, f"--world_info={self.world_info_base64}",
                f"--node_rank={node_rank}", f"--master_addr={master_addr or self.args.master_addr}",
                f"--master_port={master_port or self.args.master_port}"]


class PDSHRunner(MultiNodeRunner):
    probe, label = "pdsh", "pdsh"

    def parse_user_args(self):
        # the command line passes through a remote shell: quote everything that is not an option
        return [a if a.startswith("-") else f"'{a}'" for a in self.args.user_args]

    def get_cmd(self, environment, active_resources):
        environment["PDSH_RCMD_TYPE"] = "ssh"
        workers = ",".join(active_resources.keys())
        logger.info("Running on the following workers: %s" % workers)
        exports = "".join(f"export {k}={v}; " for k, v in self.exports.items())
        remote = [exports, f"cd {os.path.abspath('.')};"] + self._launcher_argv("%n")     # pdsh expands %n per host
        if getattr(self.args, "detect_nvlink_pairs", False):
            remote.append("--detect_nvlink_pairs")
        return ["pdsh", "-f", str(PDSH_MAX_FAN_OUT), "-w", workers] + remote + [self.user_script] + self.user_arguments


class _MpirunRunner(MultiNodeRunner):
    """Transports where an MPI launcher starts every rank and the user script discovers its rank from MPI."""
    unsupported = (("include", ""), ("exclude", ""), ("num_nodes", -1), ("num_gpus", -1))
    export_flag = "-x"

    def _export_argv(self):
        return [tok for k, v in self.exports.items() for tok in (self.export_flag, f"{k}={v}")]


class OpenMPIRunner(_MpirunRunner):
    probe, label = "ompi_info", "openmpi"

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64, resource_pool)
        self.add_export("UCX_TLS", "tcp")

    def get_cmd(self, environment, active_resources):
        self._reject_unsupported()
        head = ["mpirun", "-n", f"{self._total_processes()}", "-hostfile", f"{self.args.hostfile}",
                "--mca", "btl", "^openib", "--mca", "btl_tcp_if_include", "eth0"]
        return head + self._export_argv() + self._script_argv()


class MVAPICHRunner(_MpirunRunner):
    label = "mvapich"
    export_flag = "-env"
    _ENV = {"MV2_SMP_USE_CMA": "0",             # CMA is unavailable in containers
            "MV2_DEBUG_SHOW_BACKTRACE": "1", "MV2_USE_CUDA": "1", "MV2_SUPPORT_DL": "1",
            "MV2_ENABLE_AFFINITY": "0",        # leave thread placement to the framework (OpenMP CPU-Adam)
            "MV2_INTER_ALLGATHER_TUNING": "5", "MV2_CUDA_USE_NAIVE": "0"}

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64, resource_pool)
        for k, v in self._ENV.items():
            self.add_export(k, v)

    def backend_exists(self):
        if not shutil.which("mpiname"):
            warnings.warn("mpiname does not exist, mvapich is not installed properly")
            return False
        flavour = subprocess.check_output("mpiname", shell=True).decode("utf-8").strip()
        if "MVAPICH2-GDR" not in flavour:
            warnings.warn(f"Expected MVAPICH2-GDR as return for mpiname but received {flavour}")
            return False
        return True

    def get_cmd(self, environment, active_resources):
        self._reject_unsupported()
        per_node = list(self.resource_pool.values())
        assert all(n == per_node[0] for n in per_node), "mvapich requires same number of devices per node"
        with open(MVAPICH_TMP_HOSTFILE, "w") as fd:
            fd.writelines(f"{host}\n" for host in self.resource_pool)
        head = ["mpirun", "-np", f"{sum(per_node)}", "-ppn", f"{per_node[0]}", "--hostfile", f"{MVAPICH_TMP_HOSTFILE}"]
        return head + self._export_argv() + self._script_argv()


class SlurmRunner(MultiNodeRunner):
    probe, label = "sinfo", "slurm"

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64, resource_pool)

    def get_cmd(self, environment, active_resources):
        assert not getattr(self.args, "detect_nvlink_pairs", False), \
            "slurm backend does not support remapping visible devices"
        cmd = ["srun", "-n", f"{self._total_processes()}"]
        optional = (("--comment", getattr(self.args, "comment", "")), ("--include", self.args.include),
                    ("--exclude", self.args.exclude),
                    ("--nodes", self.args.num_nodes if self.args.num_nodes > 0 else ""),
                    ("--gpus", self.args.num_gpus if self.args.num_gpus > 0 else ""))
        for flag, value in optional:
            if value != "":
                cmd += [flag, f"{value}"]
        exports = ",".join(["--export=ALL"] + [f"{k}={v}" for k, v in self.exports.items()])
        return cmd + [exports] + self._script_argv()


class MosaicMLRunner(MultiNodeRunner):
    """MosaicML platform: the platform starts one container per node and tells it NODE_RANK / MASTER_ADDR / MASTER_PORT
    through the environment; each node then runs the per-node launcher."""
    label = "mosaicml"

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64, resource_pool)

    def backend_exists(self):
        return "NODE_RANK" in os.environ or shutil.which("mcli")

    def get_cmd(self, environment, active_resources):
        env = os.environ
        return self._launcher_argv(env.get("NODE_RANK", "0"), env.get("MASTER_ADDR"), env.get("MASTER_PORT")) + \
            [self.user_script] + self.user_arguments
