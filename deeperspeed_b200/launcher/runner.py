"""`deepspeed` command line front end.

Turns a hostfile (or the local device count) plus include / exclude filters into the {host: [slots]} map of the job,
encodes it for the per-node spawner and either starts that spawner directly (one node) or hands the job to a
multi-node transport (pdsh / OpenMPI / MVAPICH / Slurm / MosaicML). Behaviour follows the reference CLI
(deepspeed/launcher/runner.py:33-383: option names, hostfile grammar `name slots=N`, filter grammar
`host[:slot,slot][@host...]`, exported environment prefixes, `.deepspeed_env`); its single-node path referenced an
undefined `runner` - here the two paths are separate functions.
"""
import argparse
import base64
import collections
import json
import os
import re
import subprocess
import sys

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

_TRANSPORTS = {PDSH_LAUNCHER: PDSHRunner, OPENMPI_LAUNCHER: OpenMPIRunner, MVAPICH_LAUNCHER: MVAPICHRunner,
               SLURM_LAUNCHER: SlurmRunner, MOSAICML_LAUNCHER: MosaicMLRunner}
_HOST_LINE = re.compile(r"^(\S+)\s+slots=(\d+)$")


def parse_args(args=None):
    p = argparse.ArgumentParser(description="Start a (multi-node) multi-GPU training job.")
    add = p.add_argument
    add("-H", "--hostfile", type=str, default=DLTS_HOSTFILE, help="MPI-style hostfile, one `hostname slots=N` per line")
    add("-i", "--include", type=str, default="", help="use only these resources: host[:slot,slot...][@host...]")
    add("-e", "--exclude", type=str, default="", help="leave these resources out (same grammar; not with --include)")
    add("--num_nodes", type=int, default=-1, help="use the first N hosts of the hostfile")
    add("--num_gpus", type=int, default=-1, help="processes per node (GPUs; on a CPU-only host, gloo ranks)")
    add("--master_port", default=TORCH_DISTRIBUTED_DEFAULT_PORT, type=int, help="rendezvous port of rank 0")
    add("--master_addr", default="", type=str, help="address of the first node (looked up over ssh if empty)")
    add("--launcher", default=PDSH_LAUNCHER, type=str, help="multi-node transport: " + ", ".join(_TRANSPORTS))
    add("--launcher_args", default="", type=str, help="extra arguments for the transport, as one quoted string")
    add("--force_multi", action="store_true", help="use the multi-node transport even for a single (remote) node")
    add("--detect_nvlink_pairs", action="store_true", help="order local ranks so that NVLink partners are adjacent")
    add("--comment", default="", type=str, help="--comment passed on to srun (slurm transport)")
    add("user_script", type=str, help="training script, followed by its own arguments")
    add("user_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def fetch_hostfile(hostfile_path):
    """{hostname: slot count} in file order, or None when there is no hostfile (local resources are used then)."""
    if not os.path.isfile(hostfile_path):
        logger.warning("Unable to find hostfile, will proceed with training with local resources only.")
        return None
    pool = collections.OrderedDict()
    with open(hostfile_path) as f:
        entries = [ln.strip() for ln in f if ln.strip() and not ln.lstrip().startswith("#")]
    for entry in entries:
        m = _HOST_LINE.match(entry)
        if m is None:
            logger.error("Hostfile is not formatted correctly, unable to proceed with training.")
            raise ValueError(f"bad hostfile line: {entry!r} (expected `hostname slots=N`)")
        host, count = m.group(1), int(m.group(2))
        if host in pool:
            logger.error("Hostfile contains duplicate hosts, unable to proceed with training.")
            raise ValueError("host {} is already defined".format(host))
        pool[host] = count
    return pool


def _parse_node_specs(spec, host_info):
    """`a:0,2@b` -> [("a", [0, 2]), ("b", None)], validated against the known hosts and slots."""
    out = []
    for part in spec.split("@"):
        host, _, slot_txt = part.partition(":")
        if host not in host_info:
            raise ValueError("Hostname '{}' not found in hostfile".format(host))
        slots = None
        if slot_txt:
            slots = [int(s) for s in slot_txt.split(",")]
            for s in slots:
                if s not in host_info[host]:
                    raise ValueError("No slot '{}' specified on host '{}'".format(s, host))
        out.append((host, slots))
    return out


def parse_resource_filter(host_info, include_str="", exclude_str=""):
    """Apply an include OR an exclude filter to {host: [slots]}; hosts keep their hostfile order, emptied hosts go."""
    if include_str and exclude_str:
        raise ValueError("include_str and exclude_str are mutually exclusive.")
    if not include_str and not exclude_str:
        return host_info
    if include_str:
        chosen = {host: set(host_info[host] if slots is None else slots)
                  for host, slots in _parse_node_specs(include_str, host_info)}
    else:
        chosen = {host: set(slots) for host, slots in host_info.items()}
        for host, slots in _parse_node_specs(exclude_str, host_info):
            dropped = set(host_info[host] if slots is None else slots)
            logger.info("removing {} from {}".format(sorted(dropped), host))
            chosen[host] -= dropped
    return collections.OrderedDict((host, sorted(chosen[host])) for host in host_info if chosen.get(host))


def parse_inclusion_exclusion(resource_pool, inclusion, exclusion):
    slots_per_host = collections.OrderedDict((host, list(range(n))) for host, n in resource_pool.items())
    return parse_resource_filter(slots_per_host, include_str=inclusion, exclude_str=exclusion)


def encode_world_info(world_info):
    return base64.urlsafe_b64encode(json.dumps(world_info).encode("utf-8")).decode("utf-8")


def _resource_pool(args):
    """(pool, came_from_hostfile). Without a hostfile the pool is this machine: its GPUs, or - on a CPU-only host with
    an explicit --num_gpus - that many gloo ranks."""
    pool = fetch_hostfile(args.hostfile)
    if pool:
        return pool, True
    local = torch.cuda.device_count()
    if local == 0 and args.num_gpus > 0:
        local = args.num_gpus
    if local == 0:
        raise RuntimeError("Unable to proceed, no GPU resources available")
    args.master_addr = "127.0.0.1"
    return {"localhost": local}, False


def _trim(active, args):
    """--num_nodes keeps the first N hosts, --num_gpus the first N slots of every host."""
    if args.num_nodes > 0:
        active = collections.OrderedDict(list(active.items())[:args.num_nodes])
    if args.num_gpus > 0:
        active = collections.OrderedDict((host, list(range(args.num_gpus))) for host in active)
    return active


def _single_node_cmd(args, world_info_base64):
    cmd = [sys.executable, "-u", "-m", LAUNCH_MODULE, "--world_info={}".format(world_info_base64),
           "--master_addr={}".format(args.master_addr), "--master_port={}".format(args.master_port)]
    if args.detect_nvlink_pairs:
        cmd.append("--detect_nvlink_pairs")
    return cmd + [args.user_script] + args.user_args


def _multi_node_cmd(args, world_info_base64, resource_pool, active, env):
    args.launcher = args.launcher.lower()
    if args.launcher not in _TRANSPORTS:
        raise NotImplementedError(f"Unknown launcher {args.launcher}")
    cls = _TRANSPORTS[args.launcher]
    runner = cls(args, world_info_base64) if cls is PDSHRunner else cls(args, world_info_base64, resource_pool)
    if not runner.backend_exists():
        raise RuntimeError(f"launcher '{args.launcher}' not installed.")
    here = os.path.abspath(".")
    env["PYTHONPATH"] = here + ":" + env["PYTHONPATH"] if "PYTHONPATH" in env else here
    for name, value in env.items():          # communication / interpreter settings travel with the job
        if any(name.startswith(prefix) for prefix in EXPORT_ENVS):
            runner.add_export(name, value)
    for folder in DEEPSPEED_ENVIRONMENT_PATHS:   # plus whatever the user lists in ~/.deepspeed_env or ./.deepspeed_env
        path = os.path.join(folder, DEEPSPEED_ENVIRONMENT_NAME)
        if os.path.isfile(path):
            with open(path) as f:
                for line in f:
                    if "=" in line:
                        key, val = line.split("=", 1)
                        runner.add_export(key, val)
    return runner.get_cmd(env, active)


def main(args=None):
    args = parse_args(args)
    if (args.num_nodes >= 0 or args.num_gpus >= 0) and (args.include != "" or args.exclude != ""):
        raise ValueError("Cannot specify num_nodes/gpus with include/exclude")
    resource_pool, from_hostfile = _resource_pool(args)
    if not from_hostfile and args.num_nodes > 1:
        raise ValueError("Num nodes is >1 but no extra nodes available via hostfile")
    active = parse_inclusion_exclusion(resource_pool, args.include, args.exclude)
    env = os.environ.copy()
    if not args.master_addr:
        first_host = next(iter(active))
        out = subprocess.check_output(["ssh {} hostname -I".format(first_host)], shell=True)
        args.master_addr = out.decode("utf-8").split()[0]
        logger.info("Using IP address of {} for node {}".format(args.master_addr, first_host))
    active = _trim(active, args)
    world_info_base64 = encode_world_info(active)
    if args.force_multi or len(active) > 1:
        cmd = _multi_node_cmd(args, world_info_base64, resource_pool, active, env)
    else:
        cmd = _single_node_cmd(args, world_info_base64)
    logger.info("cmd = {}".format(" ".join(cmd)))
    result = subprocess.Popen(cmd, env=env)
    result.wait()
    if result.returncode > 0:
        sys.exit(result.returncode)


if __name__ == "__main__":
    main()
