"""ZeRO stage 2 (+ ZeRO-Offload): gradient + optimizer-state sharding.

Reference behaviour: deepspeed/runtime/zero/stage2.py:68-1894 — per-parameter hooks fill an IPG bucket which is
reduced to the owning ranks while backward is still running (`overlap_comm`), fp32 master/Adam state optionally on
the CPU with DeepSpeedCPUAdam and a fused low-precision copy-back.

B200-native shape of the same idea: buckets are fixed, contiguous, dp-interleaved slices of the gradient arena
(see flat_optimizer.py), so "bucket ready" is a counter per bucket driven by post-accumulate-grad hooks and the
reduction is ONE in-place reduce-scatter on a side stream — no per-slice `dist.reduce` calls, no IPG copy, no
Python bookkeeping of (rank, offset) lists. Offload moves only this rank's reduced slice D2H (pinned, side stream)."""
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
                 precision=torch.float16, pin_memory=True, fused_nvlink=False, allocator=None):
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
                         postscale_gradients=postscale_gradients, fused_nvlink=fused_nvlink and not cpu_offload)
        self.single_partition_of_fp32_groups = self.fp32_masters
        if cpu_offload:
            self._move_masters_to_cpu()
        self._side_stream = torch.cuda.Stream() if (overlap_comm and self._dev.type == "cuda") else None
        self._bucket_pending = {}
        self._bucket_done = set()
        self._param_buckets = {}
        if overlap_comm:
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
            inv = float(self._inv_scale.item())
            lp_groups = []
            for gi, arena in enumerate(self.arenas):
                for k in range(self.num_buckets[gi]):
                    g = self._slice(arena.grad, gi, k)
                    dst = self._master_slice(self._cpu_grad[gi], gi, k)
                    dst.copy_(g.float().mul_(inv), non_blocking=True)
                self.fp32_masters[gi].grad = self._cpu_grad[gi]
                lp_groups.append(None)
            if self._dev.type == "cuda":
                torch.cuda.current_stream().synchronize()
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

    def reduce_gradients(self):
        if self._reduced:
            return
        for a in self.arenas:
            a.absorb_stray_grads()
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
        if self.overlap_comm:
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
