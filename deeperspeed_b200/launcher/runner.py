"""`deepspeed` CLI: hostfile parsing, include/exclude filters, world-info encoding and dispatch to the single-node
spawner or a multi-node runner (reference deepspeed/launcher/runner.py:33-383)."""
import argparse
import base64
import collections
import copy
import json
import os
import subprocess
import sys
from copy import deepcopy

import torch

from ..runtime.constants import TORCH_DISTRIBUTED_DEFAULT_PORT
from ..utils.logging import logger
from .constants import PDSH_LAUNCHER, OPENMPI_LAUNCHER, MVAPICH_LAUNCHER, SLURM_LAUNCHER, MOSAICML_LAUNCHER
from .multinode_runner import PDSHRunner, OpenMPIRunner, MVAPICHRunner, SlurmRunner, MosaicMLRunner, LAUNCH_MODULE

DLTS_HOSTFILE = "/job/hostfile"
EXPORT_ENVS = ["NCCL", "PYTHON", "MV2", "UCX"]
DEEPSPEED_ENVIRONMENT_NAME = ".deepspeed_env"
DEEPSPEED_ENVIRONMENT_PATHS = [os.path.expanduser("~"), "."]
PDSH_MAX_FAN_OUT = 1024


def parse_args(args=None):
    parser = argparse.ArgumentParser(description="DeepSpeed runner to help launch distributed multi-node/multi-gpu training jobs.")
    parser.add_argument("-H", "--hostfile", type=str, default=DLTS_HOSTFILE,
                        help="Hostfile path (in MPI style) that defines the resource pool available to the job (e.g., worker-0 slots=4)")
    parser.add_argument("-i", "--include", type=str, default="",
                        help="Specify hardware resources to use during execution: NODE_SPEC[@NODE_SPEC ...] where NODE_SPEC=NAME[:SLOT[,SLOT ...]]")
    parser.add_argument("-e", "--exclude", type=str, default="",
                        help="Specify hardware resources to NOT use during execution. Mutually exclusive with --include.")
    parser.add_argument("--num_nodes", type=int, default=-1, help="Total number of worker nodes to run on (top N hosts of the hostfile)")
    parser.add_argument("--num_gpus", type=int, default=-1, help="Max number of GPUs to use on each node")
    parser.add_argument("--master_port", default=TORCH_DISTRIBUTED_DEFAULT_PORT, type=int,
                        help="(optional) Port used by PyTorch distributed for communication during training.")
    parser.add_argument("--master_addr", default="", type=str, help="(optional) IP address of node 0")
    parser.add_argument("--launcher", default=PDSH_LAUNCHER, type=str,
                        help="(optional) choose launcher backend for multi-node training: pdsh, openmpi, mvapich, slurm, mosaicml")
    parser.add_argument("--launcher_args", default="", type=str, help="(optional) pass launcher specific arguments as a single quoted argument.")
    parser.add_argument("--force_multi", action="store_true", help="Force multi-node launcher mode (e.g. a single remote node)")
    parser.add_argument("--detect_nvlink_pairs", action="store_true",
                        help="(optional) autodetects nvlink pairs and remaps CUDA_VISIBLE_DEVICES along the fastest connections")
    parser.add_argument("--comment", default="", type=str, help="(optional) Adds a --comment to the srun command (slurm launcher)")
    parser.add_argument("user_script", type=str, help="User script to launch, followed by any required arguments.")
    parser.add_argument("user_args", nargs=argparse.REMAINDER)
    return parser.parse_args(args=args)


def fetch_hostfile(hostfile_path):
    if not os.path.isfile(hostfile_path):
        logger.warning("Unable to find hostfile, will proceed with training with local resources only.")
        return None
    resource_pool = collections.OrderedDict()
    with open(hostfile_path, "r") as fd:
        for line in fd.readlines():
            line = line.strip()
            if line == "" or line.startswith("#"):
                continue
            try:
                hostname, slots = line.split()
                _, slot_count = slots.split("=")
                slot_count = int(slot_count)
            except ValueError as err:
                logger.error("Hostfile is not formatted correctly, unable to proceed with training.")
                raise err
            if hostname in resource_pool:
                logger.error("Hostfile contains duplicate hosts, unable to proceed with training.")
                raise ValueError("host {} is already defined".format(hostname))
            resource_pool[hostname] = slot_count
    return resource_pool


def parse_resource_filter(host_info, include_str="", exclude_str=""):
    """Apply 'worker-0@worker-1:0,2' style include / exclude filters to {host: [slots]}."""
    NODE_SEP, SLOT_LIST_START, SLOT_SEP = "@", ":", ","
    if include_str == "" and exclude_str == "":
        return host_info
    if include_str != "" and exclude_str != "":
        raise ValueError("include_str and exclude_str are mutually exclusive.")
    filtered_hosts = dict()
    if include_str:
        parse_str = include_str
    if exclude_str != "":
        filtered_hosts = deepcopy(host_info)
        parse_str = exclude_str
    for node_config in parse_str.split(NODE_SEP):
        if SLOT_LIST_START in node_config:
            hostname, slots = node_config.split(SLOT_LIST_START)
            slots = [int(x) for x in slots.split(SLOT_SEP)]
            if hostname not in host_info:
                raise ValueError("Hostname '{}' not found in hostfile".format(hostname))
            for s in slots:
                if s not in host_info[hostname]:
                    raise ValueError("No slot '{}' specified on host '{}'".format(s, hostname))
            if include_str:
                filtered_hosts[hostname] = slots
            elif exclude_str:
                for s in slots:
                    logger.info("removing {} from {}".format(s, hostname))
                    filtered_hosts[hostname].remove(s)
        else:
            hostname = node_config
            if hostname not in host_info:
                raise ValueError("Hostname '{}' not found in hostfile".format(hostname))
            if include_str:
                filtered_hosts[hostname] = host_info[hostname]
            elif exclude_str:
                filtered_hosts[hostname] = []
    del_keys = []
    for hostname in filtered_hosts:
        filtered_hosts[hostname] = list(set(filtered_hosts[hostname]))
        if len(filtered_hosts[hostname]) == 0:
            del_keys.append(hostname)
    for name in del_keys:
        del filtered_hosts[name]
    ordered_hosts = collections.OrderedDict()
    for host in host_info:
        if host in filtered_hosts:
            ordered_hosts[host] = sorted(filtered_hosts[host])
    return ordered_hosts


def parse_inclusion_exclusion(resource_pool, inclusion, exclusion):
    active_resources = collections.OrderedDict()
    for hostname, slots in resource_pool.items():
        active_resources[hostname] = list(range(slots))
    return parse_resource_filter(active_resources, include_str=inclusion, exclude_str=exclusion)


def encode_world_info(world_info):
    world_info_json = json.dumps(world_info).encode("utf-8")
    return base64.urlsafe_b64encode(world_info_json).decode("utf-8")


def main(args=None):
    args = parse_args(args)
    if (args.num_nodes >= 0 or args.num_gpus >= 0) and (args.include != "" or args.exclude != ""):
        raise ValueError("Cannot specify num_nodes/gpus with include/exclude")
    multi_node_exec = True
    resource_pool = fetch_hostfile(args.hostfile)
    if not resource_pool:
        resource_pool = {}
        device_count = torch.cuda.device_count()
        if device_count == 0 and args.num_gpus > 0:
            # CPU-only host (gloo back end): --num_gpus then simply says how many local processes to start
            device_count = args.num_gpus
        if device_count == 0:
            raise RuntimeError("Unable to proceed, no GPU resources available")
        resource_pool["localhost"] = device_count
        args.master_addr = "127.0.0.1"
        multi_node_exec = False
    if not multi_node_exec and args.num_nodes > 1:
        raise ValueError("Num nodes is >1 but no extra nodes available via hostfile")
    active_resources = parse_inclusion_exclusion(resource_pool, args.include, args.exclude)
    env = os.environ.copy()
    if not args.master_addr:
        first_host = list(active_resources.keys())[0]
        hostname_cmd = ["ssh {} hostname -I".format(first_host)]
        result = subprocess.check_output(hostname_cmd, shell=True)
        args.master_addr = result.decode("utf-8").split()[0]
        logger.info("Using IP address of {} for node {}".format(args.master_addr, first_host))
    if args.num_nodes > 0:
        updated = collections.OrderedDict()
        for count, hostname in enumerate(active_resources.keys()):
            if args.num_nodes == count:
                break
            updated[hostname] = active_resources[hostname]
        active_resources = updated
    if args.num_gpus > 0:
        updated = collections.OrderedDict()
        for hostname in active_resources.keys():
            updated[hostname] = list(range(args.num_gpus))
        active_resources = updated
    world_info_base64 = encode_world_info(active_resources)
    multi_node_exec = args.force_multi or len(active_resources) > 1

    if not multi_node_exec:
        deepspeed_launch = [sys.executable, "-u", "-m", LAUNCH_MODULE, "--world_info={}".format(world_info_base64),
                            "--master_addr={}".format(args.master_addr), "--master_port={}".format(args.master_port)]
        if args.detect_nvlink_pairs:
            deepspeed_launch += ["--detect_nvlink_pairs"]
        cmd = deepspeed_launch + [args.user_script] + args.user_args
    else:
        args.launcher = args.launcher.lower()
        if args.launcher == PDSH_LAUNCHER:
            runner = PDSHRunner(args, world_info_base64)
        elif args.launcher == OPENMPI_LAUNCHER:
            runner = OpenMPIRunner(args, world_info_base64, resource_pool)
        elif args.launcher == MVAPICH_LAUNCHER:
            runner = MVAPICHRunner(args, world_info_base64, resource_pool)
        elif args.launcher == SLURM_LAUNCHER:
            runner = SlurmRunner(args, world_info_base64, resource_pool)
        elif args.launcher == MOSAICML_LAUNCHER:
            runner = MosaicMLRunner(args, world_info_base64, resource_pool)
        else:
            raise NotImplementedError(f"Unknown launcher {args.launcher}")
        if not runner.backend_exists():
            raise RuntimeError(f"launcher '{args.launcher}' not installed.")
        curr_path = os.path.abspath(".")
        if "PYTHONPATH" in env:
            env["PYTHONPATH"] = curr_path + ":" + env["PYTHONPATH"]
        else:
            env["PYTHONPATH"] = curr_path
        exports = ""
        for var in env.keys():
            if any([var.startswith(name) for name in EXPORT_ENVS]):
                runner.add_export(var, env[var])
        for environ_path in DEEPSPEED_ENVIRONMENT_PATHS:
            environ_file = os.path.join(environ_path, DEEPSPEED_ENVIRONMENT_NAME)
            if os.path.isfile(environ_file):
                with open(environ_file, "r") as fd:
                    for var in fd.readlines():
                        key, val = var.split("=", 1)
                        runner.add_export(key, val)
        cmd = runner.get_cmd(env, active_resources)
    logger.info("cmd = {}".format(" ".join(cmd)))
    result = subprocess.Popen(cmd, env=env)
    result.wait()
    if result.returncode > 0:
        sys.exit(result.returncode)


if __name__ == "__main__":
    main()
