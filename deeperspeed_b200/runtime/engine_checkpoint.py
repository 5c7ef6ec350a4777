"""Checkpoint save / load of DeepSpeedEngine (mixin).

On-disk format (reference deepspeed/runtime/engine.py:1462-1915), kept byte-compatible in names and keys:

  <dir>/latest                                                         text file holding the newest tag
  <dir>/<tag>/mp_rank_{mp:02d}_model_states.pt                         module + engine counters (+ optimizer w/o ZeRO)
  <dir>/<tag>/zero_pp_rank_{dp}_mp_rank_{mp:02d}_optim_states.pt       one ZeRO shard per data-parallel rank
  <dir>/<tag>/zero_pp_rank_{dp}_mp_rank_{mp:02d}_model_states.pt       ZeRO-3: the model-states file is per dp rank too
  <dir>/<tag>/zero_to_fp32.py                                          stand-alone recovery script

`CheckpointLayout` owns the naming; the mixin is three explicit phases each way (model states, ZeRO shards, the
`latest` pointer) instead of one routine per file kind threaded through flags.
"""
import hashlib
import os
from collections import OrderedDict
from shutil import copyfile

import torch
import torch.distributed as dist

from ..utils.logging import logger, log_dist

# keys of the model-states file that belong to the engine; everything else is the caller's `client_state`
_ENGINE_KEYS = ("module", "optimizer", "lr_scheduler", "csr_tensor_module_names", "skipped_steps", "global_steps",
                "dp_world_size", "mp_world_size", "amp")


def _torch_load(path, map_location=None):
    return torch.load(path, map_location=map_location, weights_only=False)


class CheckpointLayout:
    """File names of one checkpoint directory."""

    def __init__(self, root, tag):
        self.root, self.tag = root, str(tag)

    @property
    def folder(self):
        return os.path.join(self.root, self.tag)

    def model_states(self, mp_rank, dp_rank=None):
        """dp_rank is given under ZeRO-3, where every data-parallel rank writes its own (placeholder) module file."""
        stem = f"mp_rank_{mp_rank:02d}" if dp_rank is None else f"zero_pp_rank_{dp_rank}_mp_rank_{mp_rank:02d}"
        return os.path.join(self.folder, stem + "_model_states.pt")

    def zero_shard(self, mp_rank, dp_rank):
        return os.path.join(self.folder, f"zero_pp_rank_{dp_rank}_mp_rank_{mp_rank:02d}_optim_states.pt")

    def existing_zero_shard(self, mp_rank, dp_rank):
        """The shard file, also accepting the pre-0.3 spelling without the underscore before 'optim_states'."""
        path = self.zero_shard(mp_rank, dp_rank)
        legacy = path.replace("_optim_states.pt", "optim_states.pt")
        return legacy if (not os.path.exists(path) and os.path.exists(legacy)) else path

    @staticmethod
    def latest_file(root):
        return os.path.join(root, "latest")


class CheckpointMixin:
    # ------------------------------------------------------------------ hooks subclasses override
    def module_state_dict(self, destination=None, prefix="", keep_vars=False):
        return self.module.state_dict(destination=destination, prefix=prefix, keep_vars=keep_vars)

    def load_module_state_dict(self, state_dict, strict=True):
        self.module.load_state_dict(state_dict, strict=strict)

    # ------------------------------------------------------------------ coordinates / names
    def _mp_rank(self):
        return 0 if self.mpu is None else self.mpu.get_model_parallel_rank()

    def _zero_dp_rank(self):
        return dist.get_rank(group=self.optimizer.dp_process_group)

    def _get_rank_zero_ckpt_name(self, checkpoints_path, tag, mp_rank, dp_rank):
        return CheckpointLayout(checkpoints_path, tag).zero_shard(mp_rank, dp_rank)

    def _get_zero_ckpt_name(self, checkpoints_path, tag):
        return CheckpointLayout(checkpoints_path, tag).zero_shard(self._mp_rank(), self._zero_dp_rank())

    def _get_ckpt_name(self, checkpoints_path, tag):
        dp = self._zero_dp_rank() if self.zero_optimization_partition_weights() else None
        return CheckpointLayout(checkpoints_path, tag).model_states(self._mp_rank(), dp)

    # ------------------------------------------------------------------ load
    def load_checkpoint(self, load_dir, tag=None, load_module_strict=True, load_optimizer_states=True,
                        load_lr_scheduler_states=True):
        """Returns (path of the model-states file, client_state); (None, None) when nothing could be loaded."""
        if tag is None:
            pointer = CheckpointLayout.latest_file(load_dir)
            if not os.path.isfile(pointer):
                logger.warning(f"Unable to find latest file at {pointer}, if trying to load latest checkpoint please "
                               "ensure this file exists or pass an explicit checkpoint tag when loading a checkpoint.")
                return None, None
            with open(pointer, "r") as fd:
                tag = fd.read().strip()
        layout = CheckpointLayout(load_dir, tag)
        path, client_state = self._load_model_states(layout, load_module_strict, load_optimizer_states,
                                                     load_lr_scheduler_states)
        if path is not None and self.zero_optimization():
            self._load_zero_shards(layout, load_optimizer_states)
        return path, client_state

    def _load_model_states(self, layout, strict, with_optimizer, with_scheduler):
        path = self._get_ckpt_name(layout.root, layout.tag)
        if not os.path.exists(path) and self.zero_optimization_partition_weights():
            # ZeRO-3 writes one model-states file per data-parallel rank; a run with MORE ranks than the saved one
            # reads rank 0's (counters and client state are the same in all of them, the weights come from the shards)
            path = layout.model_states(self._mp_rank(), 0)
        if not os.path.exists(path):
            logger.warning("Client provided checkpoint load path: {} does not exist ... skip checkpoint load".format(path))
            return None, None
        logger.info(f"rank: {self.global_rank} loading checkpoint: {path}")
        ckpt = _torch_load(path, map_location=lambda storage, loc: storage)
        self._curr_ckpt_path = layout.folder        # pipeline engines read their per-layer files from here
        if self.zero_optimization_partition_weights():
            # the file holds shard placeholders: bring every parameter (the persistent ones too) to that state; the
            # values come from the fp32 shards of the optimizer checkpoint
            self.optimizer.save_checkpoint_prologue()
        self.load_module_state_dict(state_dict=ckpt["module"], strict=strict)

        if self.optimizer is not None and not self.zero_optimization():
            if self.fp16_enabled() and self.precision() != torch.float32:
                self.optimizer.load_state_dict(ckpt["optimizer"], load_optimizer_states=with_optimizer)
            elif with_optimizer:
                self.optimizer.load_state_dict(ckpt["optimizer"])
        if with_scheduler and self.lr_scheduler is not None and ckpt.get("lr_scheduler") is not None:
            self.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])

        if getattr(self, "amp", None) is not None:
            self.amp.load_state_dict(ckpt.get("amp"))      # the GradScaler's scale and growth tracker
        self.csr_tensor_module_names = ckpt["csr_tensor_module_names"]
        self.global_steps = ckpt["global_steps"]
        self.global_samples = ckpt.get("global_samples", self.global_steps * self.train_batch_size())
        self.skipped_steps = ckpt["skipped_steps"]
        self.loaded_checkpoint_mp_world_size = ckpt["mp_world_size"]
        self.loaded_checkpoint_dp_world_size = ckpt["dp_world_size"]
        return path, {k: v for k, v in ckpt.items() if k not in _ENGINE_KEYS}

    def _get_all_zero_checkpoints(self, load_dir, tag):
        """The optimizer state dicts of EVERY data-parallel rank of the saved run (the loader re-partitions them for
        the current world size), or None if a shard is missing."""
        layout = CheckpointLayout(load_dir, tag)
        names = [layout.existing_zero_shard(self._mp_rank(), dp) for dp in range(self.loaded_checkpoint_dp_world_size)]
        missing = [n for n in names if not os.path.exists(n)]
        if missing:
            logger.warning(f"Loading checkpoint failed: {missing} do not exist; ZeRO optimizer state not loaded")
            return None
        shards = [_torch_load(n, map_location="cpu")["optimizer_state_dict"] for n in names]
        print(f"successfully loaded {len(shards)} ZeRO state_dicts for rank {self.global_rank}")
        return shards

    def _load_zero_shards(self, layout, with_optimizer):
        shards = self._get_all_zero_checkpoints(layout.root, layout.tag)
        if shards is None:
            return
        self.optimizer.load_state_dict(state_dict_list=shards, load_optimizer_states=with_optimizer,
                                       load_from_fp32_weights=self.zero_load_from_fp32_weights())
        print(f"loading {len(shards)} zero partition checkpoints for rank {self.global_rank}")

    # ------------------------------------------------------------------ save
    def _checkpoint_tag_validation(self, tag):
        """All ranks must save under the same tag: compare a hash of it with MIN / MAX all-reduces."""
        if not self.checkpoint_tag_validation_enabled():
            return
        digest = torch.ByteTensor([hashlib.sha1(tag.encode()).digest()]).flatten().to(self.device)
        lo, hi = digest.clone(), digest.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        if bool((lo == digest).all()) and bool((hi == digest).all()):
            return
        msg = (f"[rank={dist.get_rank()}] The checkpoint tag name '{tag}' is not consistent across all ranks. "
               "Including rank unique information in checkpoint tag could cause issues when restoring with different "
               "world sizes.")
        if self.checkpoint_tag_validation_fail():
            raise AssertionError(msg)
        logger.warning(msg)

    def save_checkpoint(self, save_dir, tag=None, client_state={}, save_latest=True):
        """Collective: call on ALL ranks (each writes the files it owns). Returns True."""
        zero3 = self.zero_optimization_partition_weights()
        if zero3:
            self.optimizer.save_checkpoint_prologue()
        tag = f"global_step{self.global_steps}" if tag is None else str(tag)
        self._checkpoint_tag_validation(tag)
        layout = CheckpointLayout(save_dir, tag)
        if self.save_non_zero_checkpoint or self.save_zero_checkpoint:
            try:
                os.makedirs(layout.folder, exist_ok=True)
            except Exception:
                logger.error(f"Failed saving model checkpoint to {save_dir} with tag {tag}")
        if self.save_non_zero_checkpoint:
            self._save_model_states(layout, client_state)
        if self.save_zero_checkpoint:
            self._save_zero_shard(layout)
        dist.barrier()
        if save_latest and self.global_rank == 0:
            with open(CheckpointLayout.latest_file(save_dir), "w") as fd:
                fd.write(tag)
        if zero3:
            self.optimizer.save_checkpoint_epilogue()
        return True

    def _save_model_states(self, layout, client_state):
        path = self._get_ckpt_name(layout.root, layout.tag)
        self._curr_ckpt_path = layout.folder        # module_state_dict() of a pipeline engine writes layer files here
        state = {
            "module": self.module_state_dict(),
            "optimizer": self.optimizer.state_dict() if self.optimizer and not self.zero_optimization() else None,
            "lr_scheduler": self.lr_scheduler.state_dict() if self.lr_scheduler is not None else None,
            "csr_tensor_module_names": self.csr_tensor_module_names,
            "skipped_steps": self.skipped_steps,
            "global_steps": self.global_steps,
            "global_samples": self.global_samples,
            "dp_world_size": self.dp_world_size,
            "mp_world_size": self.mp_world_size,
        }
        if getattr(self, "amp", None) is not None:
            state["amp"] = self.amp.state_dict()
        state.update(client_state)
        log_dist(message=f"Saving model checkpoint: {path}", ranks=[0])
        torch.save(state, path)
        self._curr_save_path = None

    def _get_param_shapes(self):
        """name -> full shape (ZeRO-3 parameters report the shape they had before partitioning)."""
        return OrderedDict((name, getattr(p, "ds_shape", p.shape)) for name, p in self.module.named_parameters())

    def _save_zero_shard(self, layout):
        path = self._get_zero_ckpt_name(layout.root, layout.tag)
        torch.save({"optimizer_state_dict": self.optimizer.state_dict(), "param_shapes": self._get_param_shapes()}, path)
        if self.global_rank == 0:
            self._copy_recovery_script(layout.folder)
        logger.info("zero checkpoint saved {}".format(path))

    def _copy_recovery_script(self, folder):
        """Ship zero_to_fp32.py with the checkpoint so fp32 weights can be rebuilt without this package."""
        src = os.path.join(os.path.dirname(os.path.dirname(__file__)), "utils", "zero_to_fp32.py")
        dst = os.path.join(folder, "zero_to_fp32.py")
        try:
            copyfile(src, dst)
            os.chmod(dst, os.stat(dst).st_mode | 0o111)
        except Exception as e:  # pragma: no cover
            logger.warning(f"could not copy zero_to_fp32.py: {e}")

    # ------------------------------------------------------------------ consolidated 16-bit export
    def _zero3_consolidated_fp16_state_dict(self):
        """A regular state_dict on rank 0 (CPU tensors) from ZeRO-3 partitioned weights, gathered one module at a time
        so that at most one module's parameters are whole at any moment; shared parameters are stored once."""
        from .zero.partition_parameters import GatheredParameters
        if not self.zero_optimization_partition_weights():
            raise ValueError("this function requires ZeRO-3 mode")
        on_root = dist.get_rank() == 0
        out = OrderedDict() if on_root else None
        first_key = {}
        stack = [(self.module, "")]
        while stack:
            module, prefix = stack.pop()
            with GatheredParameters(list(module.parameters(recurse=False))):
                if on_root:
                    for name, p in module.named_parameters(recurse=False):
                        if p is None:
                            continue
                        ident = getattr(p, "ds_id", id(p))
                        if ident in first_key:
                            out[prefix + name] = out[first_key[ident]]
                        else:
                            out[prefix + name] = p.detach().cpu()
                            first_key[ident] = prefix + name
                    for name, buf in module.named_buffers(recurse=False):
                        if buf is not None and name not in module._non_persistent_buffers_set:
                            out[prefix + name] = buf.detach().cpu()
            # reversed: the explicit stack must visit children in definition order, like state_dict() does
            stack.extend((child, prefix + name + ".") for name, child in reversed(list(module.named_children()))
                         if child is not None)
        return out

    def save_fp16_model(self, save_dir, save_filename="pytorch_model.bin"):
        """Write the 16-bit weights as a plain state_dict (rank 0); ZeRO-3 shards are consolidated first."""
        path = os.path.join(save_dir, save_filename)
        if not self.zero_optimization_partition_weights():
            state_dict = self.module.state_dict()
        elif self.zero_gather_fp16_weights_on_model_save():
            state_dict = self._zero3_consolidated_fp16_state_dict()
        else:
            logger.info(f"Did not save the model {path} because `stage3_gather_fp16_weights_on_model_save` is False")
            return
        if dist.get_rank() == 0:
            os.makedirs(save_dir, exist_ok=True)
            logger.info(f"Saving model weights to {path}")
            torch.save(state_dict, path)
