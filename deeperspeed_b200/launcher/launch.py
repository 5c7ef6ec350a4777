"""Per-node process spawner: one training process per local GPU with RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* set,
SIGINT/SIGTERM fan-out, and kill-all-on-first-failure (reference deepspeed/launcher/launch.py:25-175).

  python -m deeperspeed_b200.launcher.launch --world_info=<b64 json> --node_rank=N --master_addr=A --master_port=P script args..."""
import base64
import json
import os
import signal
import subprocess
import sys
import time
from argparse import ArgumentParser, REMAINDER
from collections import defaultdict

from ..runtime.constants import TORCH_DISTRIBUTED_DEFAULT_PORT
from ..utils.logging import logger


def parse_args(args=None):
    parser = ArgumentParser(description="DeepSpeed distributed training launch utility that creates multiple "
                                        "distributed processes on a single node")
    parser.add_argument("--node_rank", type=int, default=0, help="The rank of the node for multi-node distributed training")
    parser.add_argument("--master_addr", default="127.0.0.1", type=str, help="Master node (rank 0)'s address")
    parser.add_argument("--master_port", default=TORCH_DISTRIBUTED_DEFAULT_PORT, type=int, help="Master node (rank 0)'s free port")
    parser.add_argument("--world_info", default="None", type=str, help="world info base64 encoded dictionary")
    parser.add_argument("--detect_nvlink_pairs", action="store_true", help="order local ranks so NVLink pairs are adjacent")
    parser.add_argument("training_script", type=str, help="The full path to the single GPU training program/script")
    parser.add_argument("training_script_args", nargs=REMAINDER)
    return parser.parse_args(args=args)


def decode_world_info(encoded):
    return json.loads(base64.urlsafe_b64decode(encoded))


def build_rank_env(world_info, node_rank, master_addr, master_port, detect_nvlink_pairs=False):
    """Returns (list of per-process env dicts, world size) for this node. Global ranks are host-major in dict order."""
    node_list = list(world_info.keys())
    local_node = node_list[node_rank]
    local_gpu_ids = list(world_info[local_node])
    if detect_nvlink_pairs:
        from .gpu_topology import detect_nvlink_pairs as _detect
        order = [g for g in _detect() if g in local_gpu_ids]
        if len(order) == len(local_gpu_ids):
            local_gpu_ids = order
    global_rank_mapping = defaultdict(list)
    curr_global_rank = 0
    dist_world_size = 0
    for node_id in node_list:
        gids = world_info[node_id]
        dist_world_size += len(gids)
        for _ in gids:
            global_rank_mapping[node_id].append(curr_global_rank)
            curr_global_rank += 1
    envs = []
    for local_rank in range(len(local_gpu_ids)):
        env = {
            "CUDA_VISIBLE_DEVICES": ",".join(map(str, local_gpu_ids)),
            "MASTER_ADDR": str(master_addr),
            "MASTER_PORT": str(master_port),
            "WORLD_SIZE": str(dist_world_size),
            "RANK": str(global_rank_mapping[local_node][local_rank]),
            "LOCAL_RANK": str(local_rank),
        }
        envs.append(env)
    return envs, dist_world_size


def main(args=None):
    args = parse_args(args)
    current_env = os.environ.copy()
    for k in current_env.keys():
        if "NCCL" in k:
            logger.info("%s %s %s", args.node_rank, k, current_env[k])
    assert args.world_info != "None", "must provide world info dict"
    world_info = decode_world_info(args.world_info)
    logger.info("WORLD INFO DICT: {}".format(world_info))
    envs, world = build_rank_env(world_info, args.node_rank, args.master_addr, args.master_port, args.detect_nvlink_pairs)
    logger.info("nnodes={}, num_local_procs={}, node_rank={}, dist_world_size={}".format(len(world_info), len(envs),
                                                                                        args.node_rank, world))
    processes = []
    for env in envs:
        e = dict(current_env)
        e.update(env)
        cmd = [sys.executable, "-u", args.training_script, "--local_rank={}".format(env["LOCAL_RANK"])] + args.training_script_args
        processes.append(subprocess.Popen(cmd, env=e))

    sig_names = {2: "SIGINT", 15: "SIGTERM"}
    last_return_code = None

    def sigkill_handler(signum, frame):
        for process in processes:
            logger.info(f"Killing subprocess {process.pid}")
            try:
                process.kill()
            except Exception:
                pass
        if last_return_code is not None:
            # the reference raises CalledProcessError here, which makes the launcher itself exit with 1; handing the
            # failed rank's own code to the caller (runner / scheduler) is more useful
            logger.error(f"{cmd} exited with code {last_return_code}; siblings killed")
            sys.exit(last_return_code if 0 < last_return_code < 256 else 1)
        if signum in sig_names:
            logger.info(f"Main process received {sig_names[signum]}, exiting")
        sys.exit(1)

    signal.signal(signal.SIGINT, sigkill_handler)
    signal.signal(signal.SIGTERM, sigkill_handler)

    alive_processes = set(processes)
    while len(alive_processes):
        finished_processes = []
        for process in alive_processes:
            if process.poll() is None:
                continue
            if process.returncode != 0:
                last_return_code = process.returncode   # first failure kills the siblings
                sigkill_handler(signal.SIGTERM, None)
            else:
                finished_processes.append(process)
        alive_processes = set(alive_processes) - set(finished_processes)
        time.sleep(1)


if __name__ == "__main__":
    main()
