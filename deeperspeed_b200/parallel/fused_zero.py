"""Fused ZeRO step over NVLink / NVSwitch peer memory.

Adam: ONE cooperative kernel per bucket (csrc/comm/fused_zero_step.cu) does reduce-scatter (in the switch through the
heap's multicast address when there is one - NVLS - else peer loads) + average + norm / overflow, trades the norm terms
with the peers through mailboxes in the heap, derives the unscale / clip factor on the device, runs Adam on the fp32
master, casts, all-gathers by (multicast) stores into every rank's parameter arena and zeroes the gradient bucket.
With a model-parallel / pipeline group the norm also has to be reduced over that group, so the kernel is launched
twice (reduce, then update) around the host-side scalar all-reduce.
LAMB: reduce kernel + the three-phase LAMB kernel of csrc/comm/fused_zero.cu."""
import ctypes as C
import os

import torch
import torch.distributed as dist

from ..ops import kernels as K
from ..ops import native as N
from ..utils.logging import logger
from .symm import SymmetricHeap

_p, _i, _l, _f, _u = C.c_void_p, C.c_int, C.c_long, C.c_float, C.c_uint
N.register_sigs({
    "dsb_fused_rs_norm": [_p, _p, _i, _i, _l, _l, _i, _f, _p, _p, _p, _p, _u, _p],
    "dsb_fused_adam_ag": [_p, _p, _p, _p, _p, _p, _i, _i, _l, _l, _i, _f, _f, _f, _f, _f, _i, _i, _i, _p, _p, _u, _p],
})
N.register_sigs({
    "dsb_fused_lamb_ag": [_p, _p, _p, _p, _p, _p, _p, _i, _i, _l, _l, _p, _p, _i, _i, _i, _f, _f, _f, _f, _f, _f, _f,
                          _i, _i, _i, _p, _p, _p, _u, _p],
})
N.register_sigs({"dsb_fused_zero_scratch_bytes": []}, restype=_l)
N.register_sigs({
    "dsb_fused_zero_step": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _l, _i, _p, _p, _p, _p, _f, _f, _f, _f, _f,
                            _i, _i, _i, _f, _f, _p, _p, _p, _p, _p, _u, _p],
    "dsb_fused_zero_step_grid": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _l, _i, _p, _p, _p, _p, _f, _f, _f, _f,
                                 _f, _i, _i, _i, _f, _f, _p, _p, _p, _p, _p, _u, _i, _p],
})
EARLY_REDUCE_CTAS = 32     # grid of a reduce phase overlapped with backward (NVLink-bound; leaves the SMs to the GEMMs)
PH_REDUCE, PH_UPDATE, PH_EXCHANGE = 1, 2, 4


def _ptr_array(ptrs):
    return (C.c_void_p * len(ptrs))(*ptrs)


class FusedZeroStep:
    """Owns the symmetric heap that backs the arenas of a ZeroShardedOptimizer and runs its step with the fused
    kernels. Created through `maybe_create`, which returns None when the fused path cannot be used (CPU, dp == 1,
    peers not on one NVLink domain, fp32 arenas)."""

    @staticmethod
    def heap_bytes_for(param_groups, dp):
        total = 0
        for g in param_groups:
            n = sum((p.numel() + 7) // 8 * 8 for p in g["params"])
            n = (n + dp * 8 - 1) // (dp * 8) * (dp * 8) + 4096
            total += 2 * n * 2 + 1024
            total += 8 * (len(g["params"]) + 1) + 512   # per-tensor norm table of the LAMB kernel
        return total

    @staticmethod
    def make_allocator(param_groups, group):
        """Returns (allocator, heap): arenas created with this allocator live in the symmetric heap."""
        if not torch.cuda.is_available() or not dist.is_initialized() or dist.get_world_size(group=group) < 2:
            return None, None
        if any(p.dtype not in (torch.float16, torch.bfloat16) for g in param_groups for p in g["params"]):
            return None, None
        dp = dist.get_world_size(group=group)
        heap = SymmetricHeap(FusedZeroStep.heap_bytes_for(param_groups, dp) + (1 << 20), group=group)
        offsets = {}

        def allocator(numel, dtype, device):
            t, off = heap.alloc(numel, dtype)
            offsets[t.data_ptr()] = off
            return t

        heap.offsets = offsets
        return allocator, heap

    @staticmethod
    def maybe_create(opt):
        heap = getattr(opt, "_symm_heap", None)
        if heap is None:
            return None
        return FusedZeroStep(opt, heap)

    def __init__(self, opt, heap):
        self.opt = opt
        self.heap = heap
        self.lib = N.cuda_lib()
        self._prof = [] if os.environ.get("DSB_FUSED_ZERO_PROFILE") == "1" else None
        self.scratch = torch.zeros(self.lib.dsb_fused_zero_scratch_bytes(), dtype=torch.uint8, device=opt._dev)
        self.flag_ptrs = _ptr_array(heap.flag_ptrs())
        self.grad_ptrs, self.param_ptrs, self.mc_grads, self.mc_params = [], [], [], []
        for a in opt.arenas:
            g_off, p_off = heap.offsets[a.grad.data_ptr()], heap.offsets[a.data.data_ptr()]
            self.grad_ptrs.append(_ptr_array(heap.peer_ptrs(g_off)))
            self.param_ptrs.append(_ptr_array(heap.peer_ptrs(p_off)))
            self.mc_grads.append(heap.mc_ptr(g_off))
            self.mc_params.append(heap.mc_ptr(p_off))
        # NVLS reduce (multimem.ld_reduce) is opt-in: a reduce-SCATTER through the switch sends every byte of the local
        # arena out (the switch fetches this rank's own contribution over NVLink too), W / (W - 1) times the bytes of
        # peer loads; measured 2.07 ms vs 1.12 ms for the reduce phase at W = 2 (profiles/fused_zero_bench_w2.json)
        self.use_multicast = bool(heap.mc_base) and os.environ.get("DSB_FUSED_ZERO_MULTICAST", "0") == "1"
        self.host_out = torch.zeros(2, dtype=torch.float32, device=opt._dev)
        self._reds = {}
        self.launches = {"reduce": 0, "reduce_early": 0, "update": 0, "whole_step": 0}      # launch counts
        self._early = set()          # buckets whose reduce phase already ran under this step's backward
        self.is_lamb = bool(getattr(opt.optimizer, "supports_fused_zero_lamb", False))
        if self.is_lamb:
            self._init_lamb()
        logger.info(f"[fused-zero] rank {opt.dp_rank}/{opt.dp_world}: NVLink peer-memory ZeRO step enabled "
                    f"({sum(a.numel for a in opt.arenas)} elements, {'LAMB' if self.is_lamb else 'Adam'}, heap "
                    f"{heap.backend}, {'NVLS multimem' if self.use_multicast else 'peer loads / stores'})")

    # ---- LAMB: per-tensor segment tables of the owned slices + norm tables in the symmetric heap
    @staticmethod
    def owned_segments(arena, sub, count, R, me):
        """Segments of this rank's owned range in LOCAL order (bucket-major): returns (starts, tids) with
        starts[i] .. starts[i+1] belonging to tensor tids[i]; id len(arena.segments) stands for arena padding."""
        pad_tid = len(arena.segments)
        starts, tids = [], []
        for k in range(count):
            lo = (k * R + me) * sub
            hi = lo + sub
            pos = lo
            for ti, sg in enumerate(arena.segments):
                # a tensor owns its alignment padding up to the next tensor's start
                nxt = arena.segments[ti + 1].offset if ti + 1 < len(arena.segments) else arena.numel_used
                a, b = max(lo, sg.offset), min(hi, nxt)
                if a < b:
                    if a > pos:
                        starts.append(pos - lo + k * sub)
                        tids.append(pad_tid)
                    starts.append(a - lo + k * sub)
                    tids.append(ti)
                    pos = b
            if pos < hi:
                starts.append(pos - lo + k * sub)
                tids.append(pad_tid)
        starts.append(count * sub)
        return starts, tids

    def _init_lamb(self):
        o = self.opt
        self.lamb_red, self.lamb_seg, self.lamb_norm_ptrs, self.lamb_coeff = [], [], [], []
        for gi, arena in enumerate(o.arenas):
            sub, count = o.sub_size[gi], o.num_buckets[gi]
            starts, tids = self.owned_segments(arena, sub, count, o.dp_world, o.dp_rank)
            assert all(s % 8 == 0 for s in starts)
            nt = len(arena.segments) + 1
            _, off = self.heap.alloc(2 * nt, torch.float32)
            self.lamb_norm_ptrs.append(_ptr_array(self.heap.peer_ptrs(off)))
            self.lamb_seg.append((torch.tensor(starts, dtype=torch.int64, device=o._dev),
                                  torch.tensor(tids, dtype=torch.int32, device=o._dev), len(tids), nt))
            self.lamb_red.append(torch.empty(count * sub, dtype=torch.float32, device=o._dev))
            self.lamb_coeff.append(torch.ones(nt, dtype=torch.float32, device=o._dev))
            st = o.optimizer.state[o.fp32_masters[gi]]
            if "exp_avg" not in st:
                st["step"] = 0
                st["exp_avg"] = torch.zeros_like(o.fp32_masters[gi].data)
                st["exp_avg_sq"] = torch.zeros_like(o.fp32_masters[gi].data)
        dist.barrier(group=self.heap.group)

    def _lamb_update(self, s):
        o = self.opt
        R, me = o.dp_world, o.dp_rank
        inner = o.optimizer
        del inner.lamb_coeffs[:]
        for gi, arena in enumerate(o.arenas):
            group = inner.param_groups[gi]
            master = o.fp32_masters[gi]
            st = inner.state[master]
            st["step"] = st.get("step", 0) + 1
            beta1, beta2 = group["betas"]
            starts, tids, nseg, nt = self.lamb_seg[gi]
            sub = o.sub_size[gi]
            N.check(self.lib.dsb_fused_lamb_ag(
                N.ptr(self.lamb_red[gi]), N.ptr(master.data), N.ptr(st["exp_avg"]), N.ptr(st["exp_avg_sq"]),
                self.param_ptrs[gi], self.lamb_norm_ptrs[gi], self.flag_ptrs, R, me, sub, sub * o.num_buckets[gi],
                N.ptr(starts), N.ptr(tids), nseg, nt, N.dtype_code(arena.dtype), float(group["lr"]), float(beta1),
                float(beta2), float(group["eps"]), float(group["weight_decay"]), float(group["max_coeff"]),
                float(group["min_coeff"]), int(inner.eps_mode == 0), int(bool(group["bias_correction"])),
                int(st["step"]), N.ptr(o._inv_scale), N.ptr(o._skip), N.ptr(self.lamb_coeff[gi]),
                self.heap.next_epoch(), s), "dsb_fused_lamb_ag")
            K._count()
            inner.lamb_coeffs.append(self.lamb_coeff[gi][:nt - 1])

    def _mark(self, name):
        """DSB_FUSED_ZERO_PROFILE=1: CUDA-event timestamps between the phases of the step (printed by rank 0)."""
        if self._prof is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._prof.append((name, ev))

    def _report(self):
        if self._prof is None or len(self._prof) < 2:
            return
        torch.cuda.synchronize()
        parts = [f"{b[0]}={a[1].elapsed_time(b[1]):.3f}ms" for a, b in zip(self._prof, self._prof[1:])]
        if self.opt.dp_rank == 0:
            logger.info("[fused-zero] " + " ".join(parts))
        self._prof = []

    # ---- Adam: the one-kernel step ----------------------------------------------------------------------------
    def reduce_early(self, gi, k):
        """Reduce phase of one bucket on the CURRENT stream with a small grid, as soon as its gradients are final
        (called from ZeRO-2's per-bucket hooks on the reduction stream while backward is still running). Every rank
        calls it for the same buckets in the same order; the step then skips the reduce of these buckets."""
        if self.is_lamb or (gi, k) in self._early:
            return False
        self._launch(gi, k, PH_REDUCE, N.stream_ptr(), max_ctas=EARLY_REDUCE_CTAS)
        self._early.add((gi, k))
        self.launches["reduce_early"] += 1
        self.launches["reduce"] -= 1
        return True

    def _launch(self, gi, k, phases, s, epoch=None, max_ctas=0):
        o = self.opt
        arena = o.arenas[gi]
        group = o.optimizer.param_groups[gi]
        master = o.fp32_masters[gi]
        st = o.optimizer.state[master]
        R, me, sub = o.dp_world, o.dp_rank, o.sub_size[gi]
        beta1, beta2 = group["betas"]
        mc = self.use_multicast
        N.check(self.lib.dsb_fused_zero_step_grid(
            self.grad_ptrs[gi], self.param_ptrs[gi], self.flag_ptrs, self.mc_grads[gi] if mc else None,
            self.mc_params[gi] if mc else None, R, me, phases, 1, (k * R + me) * sub, k * R * sub, R * sub, sub,
            N.dtype_code(arena.dtype), N.ptr(self._red(gi, k)), N.ptr(o._master_slice(master.data, gi, k)),
            N.ptr(o._master_slice(st["exp_avg"], gi, k)), N.ptr(o._master_slice(st["exp_avg_sq"], gi, k)),
            float(group["lr"]), float(beta1), float(beta2), float(group["eps"]), float(group["weight_decay"]),
            int(o.optimizer.adam_w_mode), int(group["bias_correction"]), int(group.get("step", 1)), float(o._loss_scale()),
            float(o.clip_grad), N.ptr(o._sumsq), N.ptr(o._inv_scale), N.ptr(o._skip), N.ptr(self.host_out),
            N.ptr(self.scratch), self.heap.next_epoch(), int(max_ctas), s), "dsb_fused_zero_step_grid")
        K._count()
        self.launches["whole_step" if phases == 7 else "reduce" if phases & PH_REDUCE else "update"] += 1

    def phase_times_ms(self):
        """Phase durations of the LAST fused_zero_step launch from its in-kernel %globaltimer stamps: barrier wait,
        reduce, norm fold + exchange, update (+ all-gather + gradient zeroing), closing barrier."""
        torch.cuda.synchronize()
        t = self.scratch[2560:2560 + 48].view(torch.int64).tolist()
        names = ["entry_barrier", "reduce", "fold_exchange", "update", "closing_barrier"]
        return {n: round((t[i + 1] - t[i]) / 1e6, 4) for i, n in enumerate(names)}

    def _red(self, gi, k):
        """Scratch for the reduced slice of bucket k in the arena dtype (None: the kernel reduces in place in the
        gradient arena; DSB_FUSED_ZERO_INPLACE=1)."""
        if os.environ.get("DSB_FUSED_ZERO_INPLACE", "0") == "1":
            return None
        key = (gi, k)
        if key not in self._reds:
            o = self.opt
            self._reds[key] = torch.empty(o.sub_size[gi], dtype=o.arenas[gi].dtype, device=o._dev)
        return self._reds[key]

    def _adam_step(self, s):
        """Returns (norm, overflow flag) as a 2-element device tensor."""
        o = self.opt
        buckets = [(gi, k) for gi in range(len(o.arenas)) for k in range(o.num_buckets[gi])]
        for group in o.optimizer.param_groups:
            group["step"] = group.get("step", 0) + 1
        # o._sumsq was cleared at the end of the previous step and reduce phases that ran under backward have added to
        # it since; without any of those, clear it here as well (another code path may have used it in between)
        if not self._early:
            o._sumsq.zero_()
        todo = [b for b in buckets if b not in self._early]
        mp = o.mpu is not None and o.mpu.get_model_parallel_world_size() > 1
        if not mp:
            # the norm lives inside the data-parallel group: the last bucket's launch trades it through the heap
            last = buckets[-1]
            for gi, k in todo:
                if (gi, k) != last:
                    self._launch(gi, k, PH_REDUCE, s)
            self._launch(*last, (PH_REDUCE if last in todo else 0) | PH_EXCHANGE | PH_UPDATE, s)
            self._mark("reduce+exchange+update")
            for gi, k in buckets[:-1]:
                self._launch(gi, k, PH_UPDATE, s)
            scalars = self.host_out
        else:
            for gi, k in todo:
                self._launch(gi, k, PH_REDUCE, s)
            self._mark("reduce")
            o._mp_owned_correction(reduced=self._reds or None)      # the arena itself still holds raw local gradients
            norm, bad = o._global_scalars()                    # all-reduce over dp and the model-parallel group
            self._mark("scalars")
            for gi, k in buckets:
                self._launch(gi, k, PH_UPDATE, s)
            self._mark("update")
            scalars = torch.stack([norm, bad.float()])
        self._early = set()
        return scalars

    def step(self):
        o = self.opt
        self._mark("start")
        for a in o.arenas:
            a.absorb_stray_grads()
        s = N.stream_ptr()
        if self.is_lamb:
            scalars = self._lamb_step(s)
        else:
            scalars = self._adam_step(s)
        host = scalars.tolist()                                # the only device -> host read of the step
        if not self.is_lamb:
            o._sumsq.zero_()                                   # ready for reduce phases launched under the next backward
        self._report()
        overflow = bool(host[1])
        o.overflow = overflow
        o._global_grad_norm = -1 if overflow else host[0] / float(o._loss_scale())
        prev = o._loss_scale()
        o._on_step_result(overflow)
        if overflow:   # the kernels were no-ops: undo the step count
            if self.is_lamb:
                for master in o.fp32_masters:
                    st = o.optimizer.state[master]
                    st["step"] = max(st.get("step", 1) - 1, 0)
            else:
                for group in o.optimizer.param_groups:
                    group["step"] = max(group.get("step", 1) - 1, 0)
            logger.info("[deepspeed] OVERFLOW! Rank {} Skipping step. Attempted loss scale: {}, reducing to {}".format(
                dist.get_rank(), prev, o._loss_scale()))
        o._grads_already_zero = not self.is_lamb               # the update phase zeroed every gradient bucket
        o.zero_grad()
        return overflow

    def _lamb_step(self, s):
        o = self.opt
        R, me = o.dp_world, o.dp_rank
        o._sumsq.zero_()
        o._flag.zero_()
        sumsq = o._sumsq.narrow(0, 0, 1)
        reds = {}
        for gi, arena in enumerate(o.arenas):
            sub = o.sub_size[gi]
            for k in range(o.num_buckets[gi]):
                red = self.lamb_red[gi].narrow(0, k * sub, sub)
                reds[(gi, k)] = red
                off = (k * o.partition_count + me) * sub
                N.check(self.lib.dsb_fused_rs_norm(self.grad_ptrs[gi], self.flag_ptrs, R, me, off, sub,
                                                   N.dtype_code(arena.dtype), 1.0 / R, N.ptr(red), N.ptr(sumsq),
                                                   N.ptr(o._flag), N.ptr(self.scratch), self.heap.next_epoch(), s),
                        "dsb_fused_rs_norm")
                K._count()
        self._mark("rs_norm")
        o._sumsq[1] = o._flag[0].float()
        o._mp_owned_correction(reduced=reds)      # the arena still holds raw local gradients here
        norm, bad = o._global_scalars()
        self._mark("scalars")
        self._lamb_update(s)
        self._mark("update_ag")
        return torch.stack([norm, bad.float()])
