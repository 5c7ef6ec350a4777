"""Arena-based mixed-precision optimizer core shared by FP16_Optimizer and ZeRO stages 1/2.

What the reference does in Python per step — flatten grads, per-tensor overflow checks with host syncs, norm,
unscale/clip pass, inner optimizer step on fp32 copies, fp32->fp16 copy pass, all-gather
(reference fp16/fused_optimizer.py:242-350, zero/stage1.py:629-805, zero/stage2.py:1366-1530) — is here:

  1. (sharded) one in-place reduce-scatter per bucket of the gradient arena, or the fused NVLink kernel;
  2. one fused sum-of-squares + non-finite-flag kernel over the local gradient shard (results stay on the device);
  3. one tiny all-reduce of [sumsq, flag]; the combined unscale/clip factor and the skip flag are computed on the
     device, so nothing waits for the host;
  4. one fused Adam kernel per bucket: reads low-precision grads, updates fp32 master/m/v, writes the
     low-precision parameters back into the parameter arena (no unscale pass, no copy pass);
  5. (sharded) one in-place all-gather per bucket;
  6. a single 2-float device->host read at the end for loss-scale bookkeeping.

Layout: the arena of a param group is `count` buckets of dp * sub elements; rank r owns slice r of every bucket
(the reference's stage-1 "comm interval / sub-partition" layout, reference zero/stage1.py:356-410), which makes
every bucket collective in-place and lets stage 2 launch a bucket's reduction as soon as its gradients are final.
"""
import math

import torch
import torch.distributed as dist

from ...ops import kernels as K
from ...utils import comm_stub
from ...utils.logging import logger
from ..flat import FlatArena


def _dist_ready():
    return dist.is_available() and dist.is_initialized()


class FlatMixedPrecisionOptimizer:
    def __init__(self, init_optimizer, dp_process_group=None, shard=False, mpu=None, clip_grad=0.0,
                 max_elements_per_comm=int(5e8), fp32_allreduce=False, overlap_comm=False, allocator=None,
                 gradient_predivide_factor=1.0, postscale_gradients=True, verbose=False, shard_gradients=False):
        self.optimizer = init_optimizer
        # shard_gradients (ZeRO-2 memory mode): no full-size gradient arena; only this rank's reduced slices persist
        # (`grad_shards`), gradients pass through pooled bucket-sized staging buffers (runtime/zero/stage2.py)
        self.shard_gradients = bool(shard_gradients) and shard
        self.grad_shards = []
        self.mpu = mpu
        self.clip_grad = clip_grad
        self.verbose = verbose
        self.shard = shard
        self.fp32_allreduce = fp32_allreduce
        self.overlap_comm = overlap_comm
        self.gradient_predivide_factor = gradient_predivide_factor
        self.postscale_gradients = postscale_gradients
        self.dp_process_group = dp_process_group
        if _dist_ready():
            self.dp_world = dist.get_world_size(group=dp_process_group)
            self.dp_rank = dist.get_rank(group=dp_process_group)
        else:
            self.dp_world, self.dp_rank = 1, 0
        self.partition_count = self.dp_world if shard else 1
        self.overflow = False
        self._global_grad_norm = 0.0
        self.micro_step_id = 0
        self._reduced = False
        self.fused_inner = bool(getattr(self.optimizer, "supports_flat_fused", False))

        self.arenas = []          # low-precision FlatArena per param group
        self.sub_size = []        # elements per (bucket, rank) slice
        self.num_buckets = []     # buckets ("comm intervals") per group
        self.fp32_masters = []    # flat fp32 master of this rank's slices, per group
        self.group_params = []    # original (low-precision) params per group
        pc = self.partition_count
        for group in self.optimizer.param_groups:
            params = list(group["params"])
            self.group_params.append(params)
            n_est = sum((p.numel() + 7) // 8 * 8 for p in params)
            aligned_param = -(-n_est // pc)
            aligned_param = (aligned_param + 7) // 8 * 8
            aligned_comm = max(8, (max_elements_per_comm // pc) // 8 * 8)
            if aligned_param <= aligned_comm:
                count, sub = 1, aligned_param
            else:
                count = -(-aligned_param // aligned_comm)
                sub = aligned_comm
            arena = FlatArena(params, shard_count=pc * count, allocator=allocator, shard_align=8)
            # FlatArena pads to pc*count*8; recompute sub from the final size
            sub = arena.numel // (pc * count)
            if self.shard_gradients:
                self.grad_shards.append(torch.zeros(sub * count, dtype=arena.dtype, device=arena.device))
            else:
                arena.allocate_grads()
            self.arenas.append(arena)
            self.sub_size.append(sub)
            self.num_buckets.append(count)
            master = torch.cat([self._slice(arena.data, len(self.arenas) - 1, k).float() for k in range(count)])
            master = master.detach().clone()
            master.requires_grad = True
            self.fp32_masters.append(master)
            group["params"] = [master]
        self._dev = self.arenas[0].device
        self._sumsq = torch.zeros(2, dtype=torch.float32, device=self._dev)   # [sumsq, nonfinite flag as float]
        self._flag = torch.zeros(1, dtype=torch.int32, device=self._dev)
        self._inv_scale = torch.ones(1, dtype=torch.float32, device=self._dev)
        self._skip = torch.zeros(1, dtype=torch.int32, device=self._dev)
        self._init_inner_state()

    # ------------------------------------------------------------------ layout helpers
    def _slice(self, flat, gi, k, rank=None):
        """Slice owned by `rank` inside bucket k of group gi."""
        rank = self.dp_rank if (rank is None and self.shard) else (rank or 0)
        sub = self.sub_size[gi]
        return flat.narrow(0, (k * self.partition_count + rank) * sub, sub)

    def _bucket(self, flat, gi, k):
        sub = self.sub_size[gi]
        return flat.narrow(0, k * self.partition_count * sub, self.partition_count * sub)

    def _master_slice(self, t, gi, k):
        return t.narrow(0, k * self.sub_size[gi], self.sub_size[gi])

    def _owned_grad(self, gi, k):
        """The gradient the optimizer consumes for bucket k of group gi: this rank's slice (sharded) or the whole
        bucket (replicated) of the gradient arena, or the persistent owned shard when gradients are not kept whole."""
        if self.shard_gradients:
            return self._master_slice(self.grad_shards[gi], gi, k)
        grad = self.arenas[gi].grad
        return self._slice(grad, gi, k) if self.shard else self._bucket(grad, gi, k)

    def _init_inner_state(self):
        """Allocate optimizer state up front (the reference does a dummy step for the same reason)."""
        if self.fused_inner:
            for master in self.fp32_masters:
                st = self.optimizer.state[master]
                if len(st) == 0:
                    st["exp_avg"] = torch.zeros_like(master.data)
                    st["exp_avg_sq"] = torch.zeros_like(master.data)

    # ------------------------------------------------------------------ properties of the wrapper protocol
    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @param_groups.setter
    def param_groups(self, value):
        self.optimizer.param_groups = value

    @property
    def state(self):
        return self.optimizer.state

    @state.setter
    def state(self, value):
        self.optimizer.state = value

    def flat_grad_buffers(self):
        return list(self.grad_shards) if self.shard_gradients else [a.grad for a in self.arenas]

    # ------------------------------------------------------------------ gradients
    def zero_grad(self, set_grads_to_None=True):
        already_zero, self._grads_already_zero = getattr(self, "_grads_already_zero", False), False
        for a in self.arenas:
            if already_zero and a.grad is not None:
                a.bind_grads()      # a fused step cleared the arena inside its kernel
            else:
                a.zero_grads()
        for g in self.grad_shards:
            g.zero_()
        self._reduced = False

    def _loss_scale(self):
        raise NotImplementedError

    def _on_step_result(self, overflow):
        raise NotImplementedError

    def backward(self, loss, retain_graph=False):
        self.micro_step_id += 1
        scaled = loss.float() * self._loss_scale()
        scaled.backward(retain_graph=retain_graph)

    def _avg_op(self):
        return dist.ReduceOp.AVG if (self._dev.type == "cuda" and dist.get_backend(self.dp_process_group) == "nccl") \
            else dist.ReduceOp.SUM

    def reduce_bucket(self, gi, k):
        """In-place reduce-scatter (sharded) or all-reduce of bucket k of group gi, averaging over dp."""
        if self.dp_world == 1 or comm_stub.enabled():
            return
        arena = self.arenas[gi]
        buck = self._bucket(arena.grad, gi, k)
        op = self._avg_op()
        if self.fp32_allreduce and buck.dtype != torch.float32:
            tmp = buck.float()
            if self.shard:
                out = tmp.narrow(0, self.dp_rank * self.sub_size[gi], self.sub_size[gi])
                dist.reduce_scatter_tensor(out, tmp, op=op, group=self.dp_process_group)
                if op == dist.ReduceOp.SUM:
                    out.div_(self.dp_world)
                self._slice(arena.grad, gi, k).copy_(out)
            else:
                dist.all_reduce(tmp, op=op, group=self.dp_process_group)
                if op == dist.ReduceOp.SUM:
                    tmp.div_(self.dp_world)
                buck.copy_(tmp)
            return
        if self.shard:
            out = self._slice(arena.grad, gi, k)
            dist.reduce_scatter_tensor(out, buck, op=op, group=self.dp_process_group)
            if op == dist.ReduceOp.SUM:
                K.scale_(out, 1.0 / self.dp_world)
        else:
            dist.all_reduce(buck, op=op, group=self.dp_process_group)
            if op == dist.ReduceOp.SUM:
                K.scale_(buck, 1.0 / self.dp_world)

    def reduce_gradients(self):
        """Reduce every bucket (idempotent per step)."""
        if self._reduced:
            return
        for a in self.arenas:
            a.absorb_stray_grads()
        for gi in range(len(self.arenas)):
            for k in range(self.num_buckets[gi]):
                self.reduce_bucket(gi, k)
        self._reduced = True

    # reference-compatible entry points
    def reduce_scatter_gradients(self, postscale_gradients=True, gradient_predivide_factor=1.0,
                                 gradient_average=True):
        self.reduce_gradients()

    def allreduce_gradients(self):
        self.reduce_gradients()

    # ------------------------------------------------------------------ step
    def _local_norm_and_overflow(self):
        self._sumsq.zero_()
        self._flag.zero_()
        acc = self._sumsq.narrow(0, 0, 1)
        for gi, arena in enumerate(self.arenas):
            for k in range(self.num_buckets[gi]):
                K.sumsq(self._owned_grad(gi, k), out=acc, flag=self._flag, accumulate=True)
        self._sumsq[1] = self._flag[0].float()

    def _mp_owned_correction(self, reduced=None):
        """With model parallelism replicated (non-MP) params must be counted once; the arena mixes both kinds, so
        subtract the replicated segments on MP ranks > 0. `reduced` maps (group, bucket) to the fp32 tensor holding
        this rank's reduced slice when the reduction did not happen in place in the gradient arena (fused NVLink
        step): the sum of squares was taken over those values, so the correction must read them too."""
        if self.mpu is None or self.mpu.get_model_parallel_world_size() == 1:
            return
        if self.mpu.get_model_parallel_rank() == 0:
            return
        for gi, arena in enumerate(self.arenas):
            for s in arena.segments:
                if getattr(s.param, "model_parallel", False):
                    continue
                for k in range(self.num_buckets[gi]):
                    base = (k * self.partition_count + (self.dp_rank if self.shard else 0)) * self.sub_size[gi]
                    span = self.sub_size[gi] if self.shard else self.partition_count * self.sub_size[gi]
                    lo, hi = max(base, s.offset), min(base + span, s.offset + s.numel)
                    if lo < hi:
                        if reduced is None and self.shard_gradients:
                            seg = self._owned_grad(gi, k).narrow(0, lo - base, hi - lo).float()
                        elif reduced is None:
                            seg = arena.grad.narrow(0, lo, hi - lo).float()
                        else:
                            seg = reduced[(gi, k)].narrow(0, lo - base, hi - lo).float()
                        self._sumsq[0] -= (seg * seg).sum()

    def _global_scalars(self):
        """All-reduce [sumsq, flag]; compute inv combined scale + skip flag on the device."""
        if _dist_ready():
            if self.shard and self.dp_world > 1:
                dist.all_reduce(self._sumsq, group=self.dp_process_group)
            if self.mpu is not None and self.mpu.get_model_parallel_world_size() > 1:
                dist.all_reduce(self._sumsq, group=self.mpu.get_model_parallel_group())
        scale = float(self._loss_scale())
        norm = self._sumsq[0].clamp(min=0).sqrt()
        bad = (self._sumsq[1] > 0) | ~torch.isfinite(norm)
        combined = torch.full_like(norm, scale)
        if self.clip_grad > 0.0:
            clip = (norm / scale + 1e-6) / self.clip_grad
            combined = combined * torch.clamp(clip, min=1.0)
        self._inv_scale.copy_((1.0 / combined).reshape(1))
        self._skip.copy_(bad.to(torch.int32).reshape(1))
        return norm, bad

    def _update_group(self, gi):
        group = self.optimizer.param_groups[gi]
        arena = self.arenas[gi]
        master = self.fp32_masters[gi]
        if self.fused_inner:
            st = self.optimizer.state[master]
            group["step"] = group.get("step", 0)
            step_before = group["step"]
            for k in range(self.num_buckets[gi]):
                group["step"] = step_before
                self.optimizer.step_flat(group, self._master_slice(master.data, gi, k),
                                         self._master_slice(st["exp_avg"], gi, k),
                                         self._master_slice(st["exp_avg_sq"], gi, k),
                                         self._owned_grad(gi, k),
                                         lp=self._slice(arena.data, gi, k) if self.shard else self._bucket(arena.data, gi, k),
                                         inv_scale_dev=self._inv_scale, skip_dev=self._skip)
            return
        # generic inner optimizer: materialise fp32 grads, step, copy back
        g32 = torch.cat([self._owned_grad(gi, k).float() for k in range(self.num_buckets[gi])])
        g32.mul_(self._inv_scale)
        master.grad = g32

    def _writeback_group(self, gi):
        arena = self.arenas[gi]
        master = self.fp32_masters[gi]
        for k in range(self.num_buckets[gi]):
            dst = self._slice(arena.data, gi, k) if self.shard else self._bucket(arena.data, gi, k)
            K.cast(self._master_slice(master.data, gi, k), dst)

    def all_gather_params(self):
        if not self.shard or self.dp_world == 1 or comm_stub.enabled():
            return
        timer = getattr(self, "_comms_timer", None)
        if timer is not None:
            timer.start()
        try:
            self._all_gather_buckets()
        finally:
            if timer is not None:
                timer.stop()

    def _all_gather_buckets(self):
        for gi, arena in enumerate(self.arenas):
            for k in range(self.num_buckets[gi]):
                buck = self._bucket(arena.data, gi, k)
                mine = self._slice(arena.data, gi, k)
                if self.fp32_allreduce and buck.dtype == torch.bfloat16:
                    tmp = buck.float()
                    dist.all_gather_into_tensor(tmp, mine.float(), group=self.dp_process_group)
                    buck.copy_(tmp)
                else:
                    dist.all_gather_into_tensor(buck, mine, group=self.dp_process_group)

    def step(self, closure=None):
        self.reduce_gradients()
        self._local_norm_and_overflow()
        self._mp_owned_correction()
        norm, bad = self._global_scalars()
        if self.fused_inner:
            for gi in range(len(self.arenas)):
                self._update_group(gi)
            self.all_gather_params()
            host = torch.stack([norm, bad.float()]).tolist()   # the only device->host read of the step
            overflow = bool(host[1])
        else:
            host = torch.stack([norm, bad.float()]).tolist()
            overflow = bool(host[1])
            if not overflow:
                for gi in range(len(self.arenas)):
                    self._update_group(gi)
                self.optimizer.step()
                for gi in range(len(self.arenas)):
                    self.fp32_masters[gi].grad = None
                    self._writeback_group(gi)
                self.all_gather_params()
        self.overflow = overflow
        self._global_grad_norm = -1 if overflow else host[0] / float(self._loss_scale())
        prev_scale = self._loss_scale()
        self._on_step_result(overflow)
        if overflow:
            if self.fused_inner:
                for group in self.optimizer.param_groups:   # the kernels were no-ops: undo the step count
                    group["step"] = max(group.get("step", 1) - 1, 0)
            if self.verbose or True:
                logger.info("[deepspeed] OVERFLOW! Rank {} Skipping step. Attempted loss scale: {}, reducing to {}".format(
                    dist.get_rank() if _dist_ready() else 0, prev_scale, self._loss_scale()))
        self.zero_grad()
        return self.overflow

    def get_global_grad_norm(self):
        return self._global_grad_norm

    # ------------------------------------------------------------------ master <-> model sync helpers
    def refresh_fp32_params(self):
        for gi, arena in enumerate(self.arenas):
            m = self.fp32_masters[gi]
            for k in range(self.num_buckets[gi]):
                src = self._slice(arena.data, gi, k) if self.shard else self._bucket(arena.data, gi, k)
                self._master_slice(m.data, gi, k).copy_(src)

    def _restore_lp_from_master(self):
        for gi in range(len(self.arenas)):
            self._writeback_group(gi)
        self.all_gather_params()

    # ------------------------------------------------------------------ elastic (padding-free) shard <-> full
    def _strip(self, gi, k, t, rank=None):
        """Drop arena tail padding from the slice of (bucket k, rank)."""
        rank = self.dp_rank if rank is None else rank
        start = (k * self.partition_count + (rank if self.shard else 0)) * self.sub_size[gi]
        span = t.numel()
        used = self.arenas[gi].numel_used
        keep = max(0, min(span, used - start))
        return t.narrow(0, 0, keep).clone()

    def _local_stripped(self, gi, flat_master_like):
        span = self.sub_size[gi] if self.shard else self.partition_count * self.sub_size[gi]
        return [self._strip(gi, k, flat_master_like.narrow(0, k * span, span)) for k in range(self.num_buckets[gi])]

    def _merge_and_take_mine(self, gi, per_rank_lists):
        """per_rank_lists[r] = list over saved buckets of stripped tensors (saved with dp=len(per_rank_lists)).
        Rebuild the full used range in saved order (bucket-major, rank-minor) and re-slice for the current layout."""
        saved_dp = len(per_rank_lists)
        saved_count = len(per_rank_lists[0])
        pieces = []
        for k in range(saved_count):
            for r in range(saved_dp):
                pieces.append(per_rank_lists[r][k].to(self._dev).float().view(-1))
        full = torch.cat(pieces) if pieces else torch.zeros(0, device=self._dev)
        arena = self.arenas[gi]
        padded = torch.zeros(arena.numel, dtype=torch.float32, device=self._dev)
        n = min(full.numel(), arena.numel)
        padded.narrow(0, 0, n).copy_(full.narrow(0, 0, n))
        outs = []
        for k in range(self.num_buckets[gi]):
            outs.append((self._slice(padded, gi, k) if self.shard else self._bucket(padded, gi, k)).clone())
        return torch.cat(outs)
