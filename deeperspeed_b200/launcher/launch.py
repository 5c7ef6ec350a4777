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
    """Returns (list of per-process env dicts, world size) for this node. Global ranks are host-major in dict order:
    the first rank of a node is the number of GPUs listed on the nodes before it."""
    hosts = list(world_info)
    gpus = list(world_info[hosts[node_rank]])
    if detect_nvlink_pairs:
        from .gpu_topology import detect_nvlink_pairs as _detect
        paired = [g for g in _detect() if g in gpus]
        if len(paired) == len(gpus):
            gpus = paired
    first_rank = sum(len(world_info[h]) for h in hosts[:node_rank])
    world_size = sum(len(world_info[h]) for h in hosts)
    shared = {"CUDA_VISIBLE_DEVICES": ",".join(str(g) for g in gpus), "MASTER_ADDR": str(master_addr),
              "MASTER_PORT": str(master_port), "WORLD_SIZE": str(world_size)}
    return [dict(shared, RANK=str(first_rank + local), LOCAL_RANK=str(local)) for local in range(len(gpus))], world_size


class _Children:
    """The training processes of this node: started together, and brought down together - on a signal to the launcher
    or as soon as one of them fails (its exit code becomes the launcher's; the reference exits 1 whatever failed)."""

    def __init__(self, commands_and_envs):
        self.procs = [(subprocess.Popen(cmd, env=env), cmd) for cmd, env in commands_and_envs]
        for sig in (signal.SIGINT, signal.SIGTERM):
            signal.signal(sig, self._on_signal)

    def kill_all(self):
        for proc, _ in self.procs:
            logger.info(f"Killing subprocess {proc.pid}")
            try:
                proc.kill()
            except Exception:
                pass

    def _on_signal(self, signum, frame):
        self.kill_all()
        logger.info(f"Main process received {signal.Signals(signum).name}, exiting")
        sys.exit(1)

    def wait(self, poll_s=1.0):
        running = list(self.procs)
        while running:
            for proc, cmd in list(running):
                code = proc.poll()
                if code is None:
                    continue
                if code != 0:
                    self.kill_all()
                    logger.error(f"{cmd} exited with code {code}; siblings killed")
                    sys.exit(code if 0 < code < 256 else 1)
                running.remove((proc, cmd))
            if running:
                time.sleep(poll_s)


def main(args=None):
    args = parse_args(args)
    for name, value in os.environ.items():
        if "NCCL" in name:
            logger.info("%s %s %s", args.node_rank, name, value)
    assert args.world_info != "None", "must provide world info dict"
    world_info = decode_world_info(args.world_info)
    logger.info("WORLD INFO DICT: {}".format(world_info))
    envs, world = build_rank_env(world_info, args.node_rank, args.master_addr, args.master_port, args.detect_nvlink_pairs)
    logger.info("nnodes={}, num_local_procs={}, node_rank={}, dist_world_size={}".format(len(world_info), len(envs),
                                                                                        args.node_rank, world))
    launches = []
    for env in envs:
        cmd = [sys.executable, "-u", args.training_script, "--local_rank={}".format(env["LOCAL_RANK"])]
        launches.append((cmd + args.training_script_args, dict(os.environ, **env)))
    _Children(launches).wait()


if __name__ == "__main__":
    main()
