"""Secondary (host only): read / write sweep of the async I/O engine (csrc/host/aio.cpp) over block size x queue depth
x threads x single_submit x overlap_events, reporting GB/s per configuration and the best one. Counterpart of the
reference's csrc/aio/py_test/{ds_aio_handle.py, run_read_sweep.sh, run_write_sweep.sh, parse_aio_stats.py}.

  python bench/aio_bench.py [--dir /path/on/nvme] [--mb 256] [--quick] [--validate]
"""
import argparse
import itertools
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeperspeed_b200.ops.aio import aio_handle


def run(args):
    nbytes = args.mb << 20
    src = torch.empty(nbytes, dtype=torch.uint8).random_(0, 255)
    dst = torch.empty_like(src)
    if torch.cuda.is_available():
        src, dst = src.pin_memory(), dst.pin_memory()
    blocks = [128 << 10, 1 << 20, 8 << 20] if not args.quick else [1 << 20]
    depths = [4, 32, 128] if not args.quick else [32]
    threads = [1, 4, 8] if not args.quick else [1, 4]
    flags = [(False, True), (True, True), (False, False)] if not args.quick else [(False, True)]
    rows = []
    path = os.path.join(args.dir, "dsb_aio_bench.bin")
    try:
        for bs, qd, nt, (single, overlap) in itertools.product(blocks, depths, threads, flags):
            h = aio_handle(bs, qd, single, overlap, nt)
            t0 = time.perf_counter()
            h.async_pwrite(src, path)
            h.wait()
            tw = time.perf_counter() - t0
            os.posix_fadvise(os.open(path, os.O_RDONLY), 0, 0, os.POSIX_FADV_DONTNEED)
            t0 = time.perf_counter()
            h.async_pread(dst, path)
            h.wait()
            tr = time.perf_counter() - t0
            if args.validate:
                assert torch.equal(src, dst), "read back differs from what was written"
            rows.append({"block_size": bs, "queue_depth": qd, "threads": nt, "single_submit": single,
                         "overlap_events": overlap, "write_GBps": round(nbytes / tw / 1e9, 3),
                         "read_GBps": round(nbytes / tr / 1e9, 3)})
            del h
    finally:
        if os.path.exists(path):
            os.remove(path)
    best_r = max(rows, key=lambda r: r["read_GBps"])
    best_w = max(rows, key=lambda r: r["write_GBps"])
    return {"bytes": nbytes, "dir": args.dir, "configs": len(rows), "best_read": best_r, "best_write": best_w,
            "rows": rows}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--dir", default=tempfile.gettempdir())
    ap.add_argument("--mb", type=int, default=256)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--validate", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    res = run(a)
    print(json.dumps({k: v for k, v in res.items() if k != "rows"}))
    if a.out:
        json.dump(res, open(a.out, "w"), indent=1)
