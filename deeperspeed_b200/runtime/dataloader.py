"""Data loading front end of the engine (`training_data=` of `initialize`, `engine.deepspeed_io`).

`DeepSpeedDataLoader` shards a dataset over the data-parallel ranks (DistributedSampler unless a sampler is given),
starts the engine's throughput timer whenever a batch is drawn, and builds its torch DataLoader lazily on `iter()`.
`RepeatingLoader` turns any iterable into an endless stream (pipeline engines pull a fixed number of micro-batches per
step and must never see StopIteration). Surface of reference deepspeed/runtime/dataloader.py:10-101. Host batches are
pinned on CUDA machines so the engine's H2D copies can be asynchronous."""
import torch
from torch.utils.data import DataLoader, RandomSampler
from torch.utils.data.distributed import DistributedSampler


class RepeatingLoader:
    """Endless iterator over `loader`: when it is exhausted it is simply started again."""

    def __init__(self, loader):
        self.loader = loader
        self.data_iter = iter(loader)

    def __iter__(self):
        return self

    def __next__(self):
        while True:
            try:
                return next(self.data_iter)
            except StopIteration:
                self.data_iter = iter(self.loader)
                try:
                    return next(self.data_iter)
                except StopIteration:
                    raise RuntimeError("RepeatingLoader: the wrapped loader is empty") from None


class DeepSpeedDataLoader:
    def __init__(self, dataset, batch_size, pin_memory, local_rank, tput_timer, collate_fn=None, num_local_io_workers=None,
                 data_sampler=None, data_parallel_world_size=None, data_parallel_rank=None):
        self.dataset = dataset
        self.collate_fn = collate_fn
        self.tput_timer = tput_timer
        distributed = local_rank >= 0
        # without a launcher-provided rank one process drives every visible GPU (DataParallel-style batch)
        self.device_count = 1 if distributed else max(torch.cuda.device_count(), 1)
        self.batch_size = batch_size * self.device_count
        if data_sampler is None:
            data_sampler = (DistributedSampler(dataset, num_replicas=data_parallel_world_size, rank=data_parallel_rank)
                            if distributed else RandomSampler(dataset))
        self.data_sampler = data_sampler
        if num_local_io_workers is None:
            num_local_io_workers = 2 * self.device_count if torch.cuda.is_available() else 0
        self.num_local_io_workers = num_local_io_workers
        self.pin_memory = bool(pin_memory) and torch.cuda.is_available()
        self.len = len(data_sampler)
        self.dataloader = None
        self.data = None

    def __len__(self):
        return self.len

    def __iter__(self):
        self._create_dataloader()
        return self

    def __next__(self):
        if self.tput_timer:
            self.tput_timer.start()
        return next(self.data)

    def _create_dataloader(self):
        extra = {} if self.collate_fn is None else {"collate_fn": self.collate_fn}
        self.dataloader = DataLoader(self.dataset, batch_size=self.batch_size, sampler=self.data_sampler,
                                     num_workers=self.num_local_io_workers, pin_memory=self.pin_memory, **extra)
        self.data = iter(self.dataloader)
        return self.dataloader
