"""Logger + rank-filtered logging (reference deepspeed/utils/logging.py:7-60)."""
import logging
import sys

import torch.distributed as dist


class LoggerFactory:
    @staticmethod
    def create_logger(name=None, level=logging.INFO):
        if name is None:
            raise ValueError("name for logger cannot be None")
        formatter = logging.Formatter("[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d:%(funcName)s] %(message)s")
        logger_ = logging.getLogger(name)
        logger_.setLevel(level)
        logger_.propagate = False
        if not logger_.handlers:
            ch = logging.StreamHandler(stream=sys.stdout)
            ch.setLevel(level)
            ch.setFormatter(formatter)
            logger_.addHandler(ch)
        return logger_


logger = LoggerFactory.create_logger(name="DeepSpeed", level=logging.INFO)


def log_dist(message, ranks=None, level=logging.INFO):
    """Log `message` only on the listed global ranks (-1 = all ranks)."""
    should_log = not (dist.is_available() and dist.is_initialized())
    ranks = ranks or []
    my_rank = dist.get_rank() if (dist.is_available() and dist.is_initialized()) else -1
    if ranks and not should_log:
        should_log = ranks[0] == -1 or my_rank in set(ranks)
    if should_log:
        logger.log(level, "[Rank {}] {}".format(my_rank, message))
