"""ZeRO stage 1: optimizer-state sharding on the flat arena.

Semantics and checkpoint keys per reference deepspeed/runtime/zero/stage1.py:100-1197 (sub-partition / comm
interval layout, elastic padding-free checkpoints, DynamicLossScaler). The step itself is the fused sequence of
`FlatMixedPrecisionOptimizer` (in-place reduce-scatter per bucket -> fused norm/overflow -> fused Adam with
low-precision write-back -> in-place all-gather), or — when the data-parallel group lives on one NVSwitch domain —
the single fused peer-memory kernel from deeperspeed_b200.parallel.fused_zero.
"""
import torch
import torch.distributed as dist

from ...utils import comm_stub
from ...utils.logging import logger
from ..fp16.flat_optimizer import FlatMixedPrecisionOptimizer
from ..fp16.loss_scaler import LossScaler, DynamicLossScaler


def get_alignment_padding(flattened_lean_size, sub_partition_id, sub_partition_size):
    sub_partition_high_limit = (sub_partition_id + 1) * sub_partition_size
    if sub_partition_high_limit <= flattened_lean_size:
        return 0
    return min(sub_partition_size, sub_partition_high_limit - flattened_lean_size)


def get_group_alignment_padding(tensor_list, sub_partition_size, sub_partition_count):
    flattened_size = sum(t.numel() for t in tensor_list)
    return [get_alignment_padding(flattened_size, i, sub_partition_size) for i in range(sub_partition_count)]


class ZeroShardedOptimizer(FlatMixedPrecisionOptimizer):
    """Common part of stages 1 and 2 (loss scaler objects, state dict formats)."""

    zero_stage = 1

    def __init__(self, init_optimizer, static_loss_scale=1.0, dynamic_loss_scale=False, dynamic_loss_args=None,
                 verbose=True, dp_process_group=None, partition_size=None, mpu=None, all_gather_partitions=True,
                 allgather_size=int(5e8), clip_grad=0.0, max_elements_per_comm=int(5e8), elastic_checkpoint=True,
                 fp32_allreduce=False, overlap_comm=False, precision=torch.float16, allocator=None,
                 gradient_predivide_factor=1.0, postscale_gradients=True, fused_nvlink=False, shard_gradients=False):
        if dynamic_loss_scale and precision == torch.float16:
            self.loss_scaler = DynamicLossScaler() if dynamic_loss_args is None else DynamicLossScaler(**dynamic_loss_args)
            self.dynamic_loss_scale = True
        else:
            self.loss_scaler = LossScaler(scale=static_loss_scale if precision == torch.float16 else 1.0)
            self.dynamic_loss_scale = False
        self.elastic_checkpoint = elastic_checkpoint
        self.precision = precision
        self.all_gather_partitions = all_gather_partitions
        self._symm_heap = None
        if fused_nvlink and allocator is None and (getattr(init_optimizer, "supports_flat_fused", False)
                                                   or getattr(init_optimizer, "supports_fused_zero_lamb", False)):
            # arenas go into a symmetric (peer-mapped) heap so the fused NVLink kernels can address every rank
            from ...parallel.fused_zero import FusedZeroStep
            try:
                allocator, self._symm_heap = FusedZeroStep.make_allocator(init_optimizer.param_groups, dp_process_group)
            except Exception as e:  # peer mapping unavailable (no P2P / IPC): stay on the NCCL path, loudly
                logger.warning(f"[fused-zero] symmetric heap unavailable ({e!r}); using NCCL collectives")
                allocator, self._symm_heap = None, None
        super().__init__(init_optimizer, dp_process_group=dp_process_group, shard=True, mpu=mpu, clip_grad=clip_grad,
                         max_elements_per_comm=max_elements_per_comm, fp32_allreduce=fp32_allreduce,
                         overlap_comm=overlap_comm, allocator=allocator,
                         gradient_predivide_factor=gradient_predivide_factor,
                         postscale_gradients=postscale_gradients, verbose=verbose, shard_gradients=shard_gradients)
        # reference attribute names
        self.fp16_groups = self.group_params
        self.fp16_groups_flat = [a.data for a in self.arenas]
        self.local_sub_partitions_of_fp32_groups = [
            [self._master_slice(m.data, gi, k) for k in range(self.num_buckets[gi])]
            for gi, m in enumerate(self.fp32_masters)]
        self.num_comm_intervals_per_group = list(self.num_buckets)
        self.sub_partition_sizes = list(self.sub_size)
        self.fused = None
        if fused_nvlink:
            from ...parallel.fused_zero import FusedZeroStep
            self.fused = FusedZeroStep.maybe_create(self)

    # ---- loss scaling
    def _loss_scale(self):
        return self.loss_scaler.loss_scale

    @property
    def loss_scale(self):
        return self.loss_scaler.loss_scale

    @property
    def cur_scale(self):
        return self.loss_scaler.cur_scale

    @cur_scale.setter
    def cur_scale(self, value):
        self.loss_scaler.cur_scale = value

    def _on_step_result(self, overflow):
        self.loss_scaler.update_scale(overflow)

    def _use_fused(self):
        fused = getattr(self, "fused", None)
        if comm_stub.enabled():       # measurement mode: the NCCL path, whose collectives are skipped
            return False
        return fused is not None and (self.fused_inner or fused.is_lamb)

    def reduce_gradients(self):
        if self._use_fused():
            # the fused step pulls the peers' raw gradients itself; an NCCL reduction here would double count
            self._reduced = True
            return
        super().reduce_gradients()

    def reduce_bucket(self, gi, k):
        if self._use_fused():
            # Called per bucket while backward is still running (ZeRO-2 overlap_comm hooks, on the reduction stream):
            # run the fused kernel's reduce phase for this bucket now, with a small grid; the step skips it later.
            if self.overlap_comm:
                self.fused.reduce_early(gi, k)
            return
        super().reduce_bucket(gi, k)

    def step(self, closure=None, comms_timer=None):
        """`comms_timer` (reference stage1.py:688,759-775): a timer object that is started / stopped around the
        parameter all-gather, so a caller can report the communication share of the step."""
        self._comms_timer = comms_timer
        try:
            if self._use_fused():
                return self.fused.step()      # one kernel: there is no separate all-gather to time
            return super().step(closure)
        finally:
            self._comms_timer = None

    # ---- checkpoints -------------------------------------------------------------------------------------
    def _inner_state_stripped(self):
        """Inner optimizer state with every flat tensor cut into padding-free per-bucket pieces."""
        out = []
        for gi, master in enumerate(self.fp32_masters):
            st = self.optimizer.state.get(master, {})
            ent = {}
            for key, val in st.items():
                if torch.is_tensor(val) and val.numel() == master.numel():
                    ent[key] = self._local_stripped(gi, val)
                else:
                    ent[key] = val
            ent["__group_step__"] = self.optimizer.param_groups[gi].get("step", 0)
            # lr / betas / weight decay ... as they were when saved (an LR scheduler may have moved them)
            ent["__group_hparams__"] = {k: v for k, v in self.optimizer.param_groups[gi].items()
                                        if k != "params" and not torch.is_tensor(v)}
            out.append(ent)
        return out

    def _elastic_state_dict(self):
        sd = {
            "loss_scaler": self.loss_scaler,
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "overflow": self.overflow,
            "base_optimizer_state": self._inner_state_stripped(),
            "zero_stage": self.zero_stage,
            "partition_count": self.partition_count,
            "num_comm_intervals_per_group": self.num_comm_intervals_per_group,
            "local_sub_partitions_of_fp32_groups": [self._local_stripped(gi, m.data)
                                                    for gi, m in enumerate(self.fp32_masters)],
        }
        return sd

    def _rigid_state_dict(self):
        return {
            "loss_scaler": self.loss_scaler,
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "overflow": self.overflow,
            "base_optimizer_state": self.optimizer.state_dict(),
            "zero_stage": self.zero_stage,
            "partition_count": self.partition_count,
            "num_comm_intervals_per_group": self.num_comm_intervals_per_group,
            "local_sub_partitions_of_fp32_groups": [
                [self._master_slice(m.data, gi, k).clone() for k in range(self.num_buckets[gi])]
                for gi, m in enumerate(self.fp32_masters)],
        }

    def state_dict(self):
        return self._elastic_state_dict() if self.elastic_checkpoint else self._rigid_state_dict()

    def _restore_from_fp32_weights(self, all_state_dict):
        for gi, master in enumerate(self.fp32_masters):
            per_rank = [sd["local_sub_partitions_of_fp32_groups"][gi] for sd in all_state_dict]
            master.data.copy_(self._merge_and_take_mine(gi, per_rank))

    def _restore_base_optimizer_state(self, all_state_dict):
        for gi, master in enumerate(self.fp32_masters):
            st = self.optimizer.state[master]
            ents = [sd["base_optimizer_state"][gi] for sd in all_state_dict]
            for key, val in ents[0].items():
                if key == "__group_step__":
                    self.optimizer.param_groups[gi]["step"] = val
                elif key == "__group_hparams__":
                    self.optimizer.param_groups[gi].update(val)
                elif isinstance(val, list) and len(val) > 0 and torch.is_tensor(val[0]):
                    merged = self._merge_and_take_mine(gi, [e[key] for e in ents]).to(master.device)
                    if key in st and torch.is_tensor(st[key]) and st[key].numel() == merged.numel():
                        st[key].copy_(merged)
                    else:
                        st[key] = merged
                else:
                    st[key] = val

    def load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=False):
        """`state_dict_list`: the optimizer state dicts of ALL saved data-parallel ranks (elastic re-partitioning)
        or a single dict (rigid)."""
        if isinstance(state_dict_list, dict):
            state_dict_list = [state_dict_list]
        mine = state_dict_list[min(self.dp_rank, len(state_dict_list) - 1)]
        self.loss_scaler = mine["loss_scaler"]
        self.dynamic_loss_scale = mine["dynamic_loss_scale"]
        self.overflow = mine["overflow"]
        rigid = isinstance(mine["base_optimizer_state"], dict)
        if rigid:
            assert mine["partition_count"] == self.partition_count, "rigid ZeRO checkpoints need the same DP degree"
            if load_optimizer_states:
                self.optimizer.load_state_dict(mine["base_optimizer_state"])
            if load_from_fp32_weights:
                for gi, master in enumerate(self.fp32_masters):
                    for k, t in enumerate(mine["local_sub_partitions_of_fp32_groups"][gi]):
                        self._master_slice(master.data, gi, k).copy_(t)
        else:
            if load_optimizer_states:
                self._restore_base_optimizer_state(state_dict_list)
            if load_from_fp32_weights:
                self._restore_from_fp32_weights(state_dict_list)
        if load_from_fp32_weights:
            self._restore_lp_from_master()
        else:
            self.refresh_fp32_params()

    def __repr__(self):
        return repr(self.optimizer)


class FP16_DeepSpeedZeroOptimizer_Stage1(ZeroShardedOptimizer):
    zero_stage = 1
