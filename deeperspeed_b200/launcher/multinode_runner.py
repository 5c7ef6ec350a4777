"""Multi-node command builders: pdsh/ssh, OpenMPI, MVAPICH, Slurm, MosaicML
(reference deepspeed/launcher/multinode_runner.py:14-292, command shapes in SURVEY C.12)."""
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
    def __init__(self, args, world_info_base64):
        self.args = args
        self.user_arguments = self.parse_user_args()
        self.user_script = args.user_script
        self.world_info_base64 = world_info_base64
        self.exports = {}

    @abstractmethod
    def backend_exists(self):
        """Return whether the corresponding backend exists"""

    @abstractmethod
    def get_cmd(self, environment, active_resources):
        """Return the command to execute on node"""

    def add_export(self, key, var):
        self.exports[key.strip()] = var.strip()

    def parse_user_args(self):
        return self.args.user_args

    @property
    def name(self):
        return self.__class__.__name__


class PDSHRunner(MultiNodeRunner):
    def backend_exists(self):
        return shutil.which("pdsh")

    @property
    def name(self):
        return "pdsh"

    def parse_user_args(self):
        return list(map(lambda x: x if x.startswith("-") else "'{}'".format(x), self.args.user_args))

    def get_cmd(self, environment, active_resources):
        environment["PDSH_RCMD_TYPE"] = "ssh"
        active_workers = ",".join(active_resources.keys())
        logger.info("Running on the following workers: %s" % active_workers)
        pdsh_cmd_args = ["pdsh", "-f", str(PDSH_MAX_FAN_OUT), "-w", active_workers]
        exports = ""
        for key, val in self.exports.items():
            exports += "export {}={}; ".format(key, val)
        deepspeed_launch = [exports, "cd {};".format(os.path.abspath(".")), sys.executable, "-u", "-m", LAUNCH_MODULE,
                            "--world_info={}".format(self.world_info_base64), "--node_rank=%n",
                            "--master_addr={}".format(self.args.master_addr),
                            "--master_port={}".format(self.args.master_port)]
        if getattr(self.args, "detect_nvlink_pairs", False):
            deepspeed_launch.append("--detect_nvlink_pairs")
        return pdsh_cmd_args + deepspeed_launch + [self.user_script] + self.user_arguments


class OpenMPIRunner(MultiNodeRunner):
    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool
        self.add_export("UCX_TLS", "tcp")

    def backend_exists(self):
        return shutil.which("ompi_info")

    @property
    def name(self):
        return "openmpi"

    def get_cmd(self, environment, active_resources):
        assert self.args.include == "", "openmpi backend does not support worker include/exclusion"
        assert self.args.exclude == "", "openmpi backend does not support worker include/exclusion"
        assert self.args.num_nodes == -1, "openmpi backend does not support limiting num nodes"
        assert self.args.num_gpus == -1, "openmpi backend does not support limiting num gpus"
        total_process_count = sum(self.resource_pool.values())
        mpirun_cmd = ["mpirun", "-n", f"{total_process_count}", "-hostfile", f"{self.args.hostfile}", "--mca", "btl",
                      "^openib", "--mca", "btl_tcp_if_include", "eth0"]
        export_cmd = []
        for k, v in self.exports.items():
            export_cmd += ["-x", f"{k}={v}"]
        python_exec = [sys.executable, "-u"]
        return mpirun_cmd + export_cmd + python_exec + [self.user_script] + self.user_arguments


class MVAPICHRunner(MultiNodeRunner):
    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool
        self.add_export("MV2_SMP_USE_CMA", "0")
        self.add_export("MV2_DEBUG_SHOW_BACKTRACE", "1")
        self.add_export("MV2_USE_CUDA", "1")
        self.add_export("MV2_SUPPORT_DL", "1")
        self.add_export("MV2_ENABLE_AFFINITY", "0")
        self.add_export("MV2_INTER_ALLGATHER_TUNING", "5")
        self.add_export("MV2_CUDA_USE_NAIVE", "0")

    def backend_exists(self):
        mpiname_exists = shutil.which("mpiname")
        exists = False
        if not mpiname_exists:
            warnings.warn("mpiname does not exist, mvapich is not installed properly")
        else:
            results = subprocess.check_output("mpiname", shell=True)
            mpiname_results = results.decode("utf-8").strip()
            if "MVAPICH2-GDR" in mpiname_results:
                exists = True
            else:
                warnings.warn(f"Expected MVAPICH2-GDR as return for mpiname but received {mpiname_results}")
        return exists

    @property
    def name(self):
        return "mvapich"

    def get_cmd(self, environment, active_resources):
        assert self.args.include == "", "mvapich backend does not support worker include/exclusion"
        assert self.args.exclude == "", "mvapich backend does not support worker include/exclusion"
        assert self.args.num_nodes == -1, "mvapich backend does not support limiting num nodes"
        assert self.args.num_gpus == -1, "mvapich backend does not support limiting num gpus"
        devices_per_node = self.resource_pool.values()
        total_process_count = sum(devices_per_node)
        process_per_node = list(devices_per_node)[0]
        assert all([n == process_per_node for n in devices_per_node]), "mvapich requires same number of devices per node"
        with open(MVAPICH_TMP_HOSTFILE, "w") as fd:
            for host in self.resource_pool.keys():
                fd.write(f"{host}\n")
        mpirun_cmd = ["mpirun", "-np", f"{total_process_count}", "-ppn", f"{process_per_node}", "--hostfile",
                      f"{MVAPICH_TMP_HOSTFILE}"]
        export_cmd = []
        for k, v in self.exports.items():
            export_cmd += ["-env", f"{k}={v}"]
        python_exec = [sys.executable, "-u"]
        return mpirun_cmd + export_cmd + python_exec + [self.user_script] + self.user_arguments


class SlurmRunner(MultiNodeRunner):
    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool

    def backend_exists(self):
        return shutil.which("sinfo")

    @property
    def name(self):
        return "slurm"

    def get_cmd(self, environment, active_resources):
        assert not getattr(self.args, "detect_nvlink_pairs", False), "slurm backend does not support remapping visible devices"
        total_process_count = sum(self.resource_pool.values())
        srun_cmd = ["srun", "-n", f"{total_process_count}"]
        if getattr(self.args, "comment", ""):
            srun_cmd += ["--comment", self.args.comment]
        if self.args.include != "":
            srun_cmd += ["--include", f"{self.args.include}"]
        if self.args.exclude != "":
            srun_cmd += ["--exclude", f"{self.args.exclude}"]
        if self.args.num_nodes > 0:
            srun_cmd += ["--nodes", f"{self.args.num_nodes}"]
        if self.args.num_gpus > 0:
            srun_cmd += ["--gpus", f"{self.args.num_gpus}"]
        exports = "--export=ALL"
        for key, val in self.exports.items():
            exports += f",{key}={val}"
        python_exec = [sys.executable, "-u"]
        return srun_cmd + [exports] + python_exec + [self.user_script] + self.user_arguments


class MosaicMLRunner(MultiNodeRunner):
    """MosaicML platform: ranks are started by the platform; NODE_RANK / MASTER_ADDR come from its environment."""

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool

    def backend_exists(self):
        return "NODE_RANK" in os.environ or shutil.which("mcli")

    @property
    def name(self):
        return "mosaicml"

    def get_cmd(self, environment, active_resources):
        node_rank = os.environ.get("NODE_RANK", "0")
        master_addr = os.environ.get("MASTER_ADDR", self.args.master_addr)
        master_port = os.environ.get("MASTER_PORT", str(self.args.master_port))
        launch = [sys.executable, "-u", "-m", LAUNCH_MODULE, "--world_info={}".format(self.world_info_base64),
                  "--node_rank={}".format(node_rank), "--master_addr={}".format(master_addr),
                  "--master_port={}".format(master_port)]
        return launch + [self.user_script] + self.user_arguments
