from .logging import logger, log_dist
from .distributed import init_distributed
from .timer import SynchronizedWallClockTimer, ThroughputTimer
from ..runtime.dataloader import RepeatingLoader
