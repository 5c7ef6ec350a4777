"""ZeRO stage 2 (+ ZeRO-Offload): gradient + optimizer-state sharding.

Reference behaviour: deepspeed/runtime/zero/stage2.py:68-1894 — per-parameter hooks fill an IPG bucket which is
reduced to the owning ranks while backward is still running (`overlap_comm`), fp32 master/Adam state optionally on
the CPU with DeepSpeedCPUAdam and a fused low-precision copy-back.

B200-native shape of the same idea: buckets are fixed, contiguous, dp-interleaved slices of the gradient arena
(see flat_optimizer.py), so "bucket ready" is a counter per bucket driven by post-accumulate-grad hooks and the
reduction is ONE in-place reduce-scatter on a side stream — no per-slice `dist.reduce` calls, no IPG copy, no
Python bookkeeping of (rank, offset) lists. Offload moves only this rank's reduced slice D2H (pinned, side stream)."""
from contextlib import nullcontext as _nullcontext

import torch
import torch.distributed as dist

from ...ops import kernels as K
from ...utils.logging import logger
from .stage1 import ZeroShardedOptimizer


class FP16_DeepSpeedZeroOptimizer(ZeroShardedOptimizer):
    zero_stage = 2

    def __init__(self, init_optimizer, timers=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None, verbose=True, contiguous_gradients=True, reduce_bucket_size=int(5e8),
                 allgather_bucket_size=int(5e8), dp_process_group=None, reduce_scatter=True, overlap_comm=False,
                 cpu_offload=False, mpu=None, clip_grad=0.0, allreduce_always_fp32=False, postscale_gradients=True,
                 gradient_predivide_factor=1.0, gradient_accumulation_steps=1, elastic_checkpoint=True,
                 precision=torch.float16, pin_memory=True, fused_nvlink=False, allocator=None,
                 shard_gradients=False):
        self.timers = timers
        self.cpu_offload = cpu_offload
        self.contiguous_gradients = contiguous_gradients
        self.reduce_scatter = reduce_scatter
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.is_gradient_accumulation_boundary = True
        self._pin = pin_memory and torch.cuda.is_available()
        if cpu_offload:
            # the inner optimizer must see CPU masters: build the arenas first with a placeholder, then move
            pass
        super().__init__(init_optimizer, static_loss_scale=static_loss_scale, dynamic_loss_scale=dynamic_loss_scale,
                         dynamic_loss_args=dynamic_loss_args, verbose=verbose, dp_process_group=dp_process_group,
                         mpu=mpu, clip_grad=clip_grad, max_elements_per_comm=reduce_bucket_size,
                         elastic_checkpoint=elastic_checkpoint, fp32_allreduce=allreduce_always_fp32,
                         overlap_comm=overlap_comm, precision=precision, allocator=allocator,
                         gradient_predivide_factor=gradient_predivide_factor,
                         postscale_gradients=postscale_gradients,
                         fused_nvlink=fused_nvlink and not cpu_offload and not shard_gradients,
                         shard_gradients=shard_gradients)
        self.single_partition_of_fp32_groups = self.fp32_masters
        if cpu_offload:
            self._move_masters_to_cpu()
        self._side_stream = torch.cuda.Stream() if (overlap_comm and self._dev.type == "cuda") else None
        self._bucket_pending = {}
        self._bucket_done = set()
        self._param_buckets = {}
        if self.shard_gradients:
            self._install_shard_hooks()
        elif overlap_comm:
            self._install_hooks()

    # ------------------------------------------------------------------ offload
    def _move_masters_to_cpu(self):
        self.fused_inner = False
        self._cpu_grad = []
        for gi, master in enumerate(self.fp32_masters):
            cpu_master = torch.empty(master.numel(), dtype=torch.float32, device="cpu", pin_memory=self._pin)
            cpu_master.copy_(master.data)
            cpu_master.requires_grad = True
            st = self.optimizer.state.pop(master, None)
            self.fp32_masters[gi] = cpu_master
            self.optimizer.param_groups[gi]["params"] = [cpu_master]
            self._cpu_grad.append(torch.zeros(master.numel(), dtype=torch.float32, device="cpu", pin_memory=self._pin))
            del st
        self.single_partition_of_fp32_groups = self.fp32_masters
        self.local_sub_partitions_of_fp32_groups = [
            [self._master_slice(m.data, gi, k) for k in range(self.num_buckets[gi])]
            for gi, m in enumerate(self.fp32_masters)]

    def _offload_step(self, norm, bad):
        host = torch.stack([norm, bad.float()]).tolist()
        overflow = bool(host[1])
        if not overflow:
            # unscale on the device, then pinned device -> host copies on a copy stream; the host waits for exactly
            # those copies (one event), not for the whole device
            on_gpu = self._dev.type == "cuda"
            if on_gpu:
                if getattr(self, "_d2h_stream", None) is None:
                    self._d2h_stream = torch.cuda.Stream()
                self._d2h_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._d2h_stream) if on_gpu else _nullcontext():
                for gi, arena in enumerate(self.arenas):
                    for k in range(self.num_buckets[gi]):
                        wide = self._owned_grad(gi, k).float().mul_(self._inv_scale)
                        self._master_slice(self._cpu_grad[gi], gi, k).copy_(wide, non_blocking=True)
                    self.fp32_masters[gi].grad = self._cpu_grad[gi]
                if on_gpu:
                    landed = torch.cuda.Event()
                    landed.record()
            if on_gpu:
                landed.synchronize()
            from ...ops.adam import DeepSpeedCPUAdam
            lowp = all(a.dtype in (torch.float16, torch.bfloat16) for a in self.arenas)
            if isinstance(self.optimizer, DeepSpeedCPUAdam) and lowp and all(n == 1 for n in self.num_buckets):
                # fused copy-back: the CPU optimizer writes the low-precision slice straight to the GPU arena
                self.optimizer.step(fp16_param_groups=[[self._slice(a.data, gi, 0)] for gi, a in enumerate(self.arenas)])
            else:
                self.optimizer.step()
                for gi, arena in enumerate(self.arenas):
                    m = self.fp32_masters[gi]
                    for k in range(self.num_buckets[gi]):
                        self._slice(arena.data, gi, k).copy_(self._master_slice(m.data, gi, k), non_blocking=True)
            for m in self.fp32_masters:
                m.grad = None
            self.all_gather_params()
        return host, overflow

    def step(self, closure=None):
        if not self.cpu_offload:
            return super().step(closure)
        self.reduce_gradients()
        self._local_norm_and_overflow()
        self._mp_owned_correction()
        norm, bad = self._global_scalars()
        host, overflow = self._offload_step(norm, bad)
        self.overflow = overflow
        self._global_grad_norm = -1 if overflow else host[0] / float(self._loss_scale())
        prev = self._loss_scale()
        self._on_step_result(overflow)
        if overflow:
            logger.info("[deepspeed] OVERFLOW! Rank {} Skipping step. Attempted loss scale: {}, reducing to {}".format(
                dist.get_rank() if dist.is_initialized() else 0, prev, self._loss_scale()))
        self.zero_grad()
        return self.overflow

    def refresh_fp32_params(self):
        for gi, arena in enumerate(self.arenas):
            m = self.fp32_masters[gi]
            for k in range(self.num_buckets[gi]):
                self._master_slice(m.data, gi, k).copy_(self._slice(arena.data, gi, k))

    def _writeback_group(self, gi):
        arena = self.arenas[gi]
        master = self.fp32_masters[gi]
        for k in range(self.num_buckets[gi]):
            src = self._master_slice(master.data, gi, k)
            dst = self._slice(arena.data, gi, k)
            if src.device == dst.device:
                K.cast(src, dst)
            else:
                dst.copy_(src)

    # ------------------------------------------------------------------ overlap of reduction with backward
    def _install_hooks(self):
        # bucket readiness is driven by autograd's post-accumulate hooks, so gradients must flow through autograd
        K.ACCUMULATE_INTO_GRAD = False
        for gi, arena in enumerate(self.arenas):
            span = self.partition_count * self.sub_size[gi]
            for s in arena.segments:
                if not s.param.requires_grad:
                    continue
                first, last = s.offset // span, (s.offset + max(s.numel, 1) - 1) // span
                buckets = [(gi, k) for k in range(first, min(last, self.num_buckets[gi] - 1) + 1)]
                self._param_buckets[s.param] = buckets
                for b in buckets:
                    self._bucket_pending[b] = self._bucket_pending.get(b, 0) + 1
                s.param.register_post_accumulate_grad_hook(self._make_hook(s.param))
        self._bucket_total = dict(self._bucket_pending)

    def _make_hook(self, param):
        def hook(p):
            if not self.is_gradient_accumulation_boundary:
                return
            for b in self._param_buckets[param]:
                self._bucket_pending[b] -= 1
                if self._bucket_pending[b] == 0 and b not in self._bucket_done:
                    self._launch_bucket(b)
        return hook

    def _launch_bucket(self, b):
        gi, k = b
        self._bucket_done.add(b)
        if self._side_stream is not None:
            self._side_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._side_stream):
                self.reduce_bucket(gi, k)
        else:
            self.reduce_bucket(gi, k)

    # ------------------------------------------------------------------ sharded gradient memory
    # Only this rank's reduced slices persist (`grad_shards`, 1 / dp of the gradient bytes). During backward every
    # parameter gradient is copied into a bucket-sized staging buffer as soon as autograd has finished it and is freed
    # at once; a full bucket is reduce-scattered (side stream with overlap_comm) and ADDED to the owned slice, and its
    # staging buffer goes back to a pool. Live gradient memory = owned slices + the buckets in flight (normally two)
    # + the few parameter gradients autograd holds. Reduction happens every micro-batch, as in the reference
    # (deepspeed/runtime/zero/stage2.py:505-558, 679-745, 915-991).
    def _install_shard_hooks(self):
        K.ACCUMULATE_INTO_GRAD = False           # gradients must be autograd's own tensors: there is no arena to add to
        self._pieces = {}                        # param -> [(group, bucket, offset in param, offset in bucket, n)]
        self._bucket_total = {}
        self._active = {}                        # (group, bucket) -> staging buffer being filled
        self._pool = {}                          # (dtype, numel) -> [(buffer, event of its last reduction)]
        self.staging_high_water = 0              # buffers ever allocated (for tests / memory reports)
        for gi, arena in enumerate(self.arenas):
            span = self.partition_count * self.sub_size[gi]
            for seg in arena.segments:
                if not seg.param.requires_grad or seg.numel == 0:
                    continue
                lo, hi = seg.offset, seg.offset + seg.numel
                pieces = []
                for k in range(lo // span, min((hi - 1) // span, self.num_buckets[gi] - 1) + 1):
                    a, b = max(lo, k * span), min(hi, (k + 1) * span)
                    pieces.append((gi, k, a - lo, a - k * span, b - a))
                    self._bucket_total[(gi, k)] = self._bucket_total.get((gi, k), 0) + 1
                self._pieces[seg.param] = pieces
                seg.param.register_post_accumulate_grad_hook(self._gradient_arrived)
        self._bucket_pending = dict(self._bucket_total)

    def _staging(self, gi, k):
        buf = self._active.get((gi, k))
        if buf is None:
            arena = self.arenas[gi]
            key = (arena.dtype, self.partition_count * self.sub_size[gi])
            free = self._pool.setdefault(key, [])
            if free:
                buf, ev = free.pop()
                if ev is not None:
                    torch.cuda.current_stream().wait_event(ev)      # its previous reduction may still be running
            else:
                buf = torch.empty(key[1], dtype=key[0], device=arena.device)
                self.staging_high_water += 1
            buf.zero_()
            self._active[(gi, k)] = buf
        return buf

    def _gradient_arrived(self, param):
        grad = param.grad
        if grad is None:
            return
        flat = grad.contiguous().view(-1)
        for gi, k, src, dst, n in self._pieces[param]:
            self._staging(gi, k).narrow(0, dst, n).copy_(flat.narrow(0, src, n))
            self._bucket_pending[(gi, k)] -= 1
            if self._bucket_pending[(gi, k)] == 0:
                self._flush_bucket(gi, k)
        param.grad = None                         # the only full-size copy of this gradient dies here

    def _reduce_into_shard(self, gi, k, staging):
        own = self._owned_grad(gi, k)
        if self.dp_world == 1:
            own.add_(staging)
            return
        from ...utils import comm_stub
        if comm_stub.enabled():
            own.add_(staging.narrow(0, self.dp_rank * self.sub_size[gi], self.sub_size[gi]))
            return
        op = self._avg_op()
        wide = staging.float() if (self.fp32_allreduce and staging.dtype != torch.float32) else staging
        out = torch.empty(self.sub_size[gi], dtype=wide.dtype, device=wide.device)
        dist.reduce_scatter_tensor(out, wide, op=op, group=self.dp_process_group)
        if op == dist.ReduceOp.SUM:
            out.div_(self.dp_world)
        own.add_(out)

    def _flush_bucket(self, gi, k):
        staging = self._active.pop((gi, k))
        self._bucket_pending[(gi, k)] = self._bucket_total[(gi, k)]
        ev = None
        if self._side_stream is not None:
            self._side_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._side_stream):
                self._reduce_into_shard(gi, k, staging)
                ev = torch.cuda.Event()
                ev.record()
        else:
            self._reduce_into_shard(gi, k, staging)
        self._pool[(staging.dtype, staging.numel())].append((staging, ev))

    def _flush_partial_buckets(self):
        """Buckets some of whose parameters got no gradient this micro-batch (zeros stand in for them)."""
        for gi, k in sorted(self._active):
            self._flush_bucket(gi, k)
        if self._side_stream is not None:
            torch.cuda.current_stream().wait_stream(self._side_stream)

    def backward(self, loss, retain_graph=False):
        super().backward(loss, retain_graph=retain_graph)
        if self.shard_gradients:
            self._flush_partial_buckets()         # the next micro-batch must start with empty staging

    def gradient_memory_bytes(self):
        """(persistent, peak staging) gradient bytes of this rank."""
        if not self.shard_gradients:
            return sum(a.grad.numel() * a.grad.element_size() for a in self.arenas), 0
        own = sum(g.numel() * g.element_size() for g in self.grad_shards)
        pooled = sum(b.numel() * b.element_size() for free in self._pool.values() for b, _ in free)
        pooled += sum(b.numel() * b.element_size() for b in self._active.values())
        return own, pooled

    def reduce_gradients(self):
        if self._reduced:
            return
        if self.shard_gradients:
            self._flush_partial_buckets()
            self._reduced = True
            return
        for a in self.arenas:
            a.absorb_stray_grads()
        if not self._use_fused():      # the fused step reduces what is left itself, with the whole GPU
            for gi in range(len(self.arenas)):
                for k in range(self.num_buckets[gi]):
                    if (gi, k) not in self._bucket_done:
                        self.reduce_bucket(gi, k)
        if self._side_stream is not None:
            torch.cuda.current_stream().wait_stream(self._side_stream)
        self._reduced = True

    def overlapping_partition_gradients_reduce_epilogue(self):
        self.reduce_gradients()

    def zero_grad(self, set_grads_to_None=True):
        super().zero_grad(set_grads_to_None)
        self._bucket_done = set()
        if self.overlap_comm or self.shard_gradients:
            self._bucket_pending = dict(self._bucket_total)

    # ------------------------------------------------------------------ checkpoint format (stage-2 key names)
    def state_dict(self):
        sd = {
            "loss_scaler": self.loss_scaler,
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "overflow": self.overflow,
            "zero_stage": self.zero_stage,
            "partition_count": self.partition_count,
        }
        if self.elastic_checkpoint:
            sd["base_optimizer_state"] = self._inner_state_stripped()
            sd["single_partition_of_fp32_groups"] = [self._local_stripped(gi, m.data)
                                                     for gi, m in enumerate(self.fp32_masters)]
        else:
            sd["base_optimizer_state"] = self.optimizer.state_dict()
            sd["single_partition_of_fp32_groups"] = [m.data.clone() for m in self.fp32_masters]
        sd["num_comm_intervals_per_group"] = self.num_comm_intervals_per_group
        return sd

    def load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=False):
        if isinstance(state_dict_list, dict):
            state_dict_list = [state_dict_list]
        # normalise to the stage-1 key so the shared loader can be used
        norm = []
        for sd in state_dict_list:
            sd = dict(sd)
            parts = sd["single_partition_of_fp32_groups"]
            if len(parts) > 0 and torch.is_tensor(parts[0]):
                parts = [[p] for p in parts]
            sd["local_sub_partitions_of_fp32_groups"] = parts
            norm.append(sd)
        rigid = isinstance(norm[0]["base_optimizer_state"], dict)
        if rigid and load_from_fp32_weights:
            mine = norm[min(self.dp_rank, len(norm) - 1)]
            for gi, master in enumerate(self.fp32_masters):
                master.data.copy_(mine["local_sub_partitions_of_fp32_groups"][gi][0])
        super().load_state_dict(norm, load_optimizer_states=load_optimizer_states,
                                load_from_fp32_weights=load_from_fp32_weights and not rigid)
        if rigid and load_from_fp32_weights:
            self._restore_lp_from_master()
