"""Process-wide logger and rank-filtered logging. `logger`, `log_dist(message, ranks, level)` and `LoggerFactory` are the
names user code imports from the reference (deepspeed/utils/logging.py:7-60); the log line format is kept so existing
log scrapers keep working."""
import logging
import sys

import torch.distributed as dist

_LINE_FORMAT = "[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d:%(funcName)s] %(message)s"


class LoggerFactory:
    @staticmethod
    def create_logger(name=None, level=logging.INFO):
        if name is None:
            raise ValueError("name for logger cannot be None")
        log = logging.getLogger(name)
        log.setLevel(level)
        log.propagate = False                     # do not duplicate lines through the root logger
        if not log.handlers:                      # idempotent: importing twice must not double every line
            out = logging.StreamHandler(stream=sys.stdout)
            out.setLevel(level)
            out.setFormatter(logging.Formatter(_LINE_FORMAT))
            log.addHandler(out)
        return log


logger = LoggerFactory.create_logger(name="DeepSpeed", level=logging.INFO)


def _global_rank():
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else -1


def log_dist(message, ranks=None, level=logging.INFO):
    """Log on the listed global ranks only. `ranks=[-1]` means every rank; before the process group exists every
    process logs; an empty `ranks` with an initialised group logs nowhere (the reference's behaviour)."""
    me = _global_rank()
    wanted = list(ranks or [])
    emit = me == -1 or (bool(wanted) and (wanted[0] == -1 or me in wanted))
    if emit:
        logger.log(level, f"[Rank {me}] {message}")
