"""Checkpoint save / load of DeepSpeedEngine (mixin). On-disk layout is the reference's:

  <dir>/latest
  <dir>/<tag>/mp_rank_{mp:02d}_model_states.pt
  <dir>/<tag>/zero_pp_rank_{dp}_mp_rank_{mp:02d}_optim_states.pt   (+ zero_to_fp32.py recovery script)
  <dir>/<tag>/zero_pp_rank_{dp}_mp_rank_{mp:02d}_model_states.pt   (ZeRO-3)

Reference: deepspeed/runtime/engine.py:1462-1915."""
import glob
import hashlib
import os
from shutil import copyfile

import torch
import torch.distributed as dist

from ..utils.logging import logger, log_dist


def _torch_load(path, map_location=None):
    return torch.load(path, map_location=map_location, weights_only=False)


class CheckpointMixin:
    def module_state_dict(self, destination=None, prefix="", keep_vars=False):
        return self.module.state_dict(destination=destination, prefix=prefix, keep_vars=keep_vars)

    def load_module_state_dict(self, state_dict, strict=True):
        self.module.load_state_dict(state_dict, strict=strict)

    def _get_rank_zero_ckpt_name(self, checkpoints_path, tag, mp_rank, dp_rank):
        filename = "zero_pp_rank_{}".format(dp_rank)
        return os.path.join(checkpoints_path, str(tag), filename + "_mp_rank_{:02d}".format(mp_rank) + "_optim_states.pt")

    def _get_zero_ckpt_name(self, checkpoints_path, tag):
        mp_rank = 0 if self.mpu is None else self.mpu.get_model_parallel_rank()
        pp_rank = dist.get_rank(group=self.optimizer.dp_process_group)
        return self._get_rank_zero_ckpt_name(checkpoints_path, tag, mp_rank, pp_rank)

    def _get_ckpt_name(self, checkpoints_path, tag):
        mp_rank = 0 if self.mpu is None else self.mpu.get_model_parallel_rank()
        if self.zero_optimization_partition_weights():
            filename = "zero_pp_rank_{}".format(dist.get_rank(group=self.optimizer.dp_process_group))
            return os.path.join(checkpoints_path, str(tag), filename + "_mp_rank_{:02d}".format(mp_rank) + "_model_states.pt")
        return os.path.join(checkpoints_path, str(tag), "mp_rank_{:02d}".format(mp_rank) + "_model_states.pt")

    # ------------------------------------------------------------------ load
    def load_checkpoint(self, load_dir, tag=None, load_module_strict=True, load_optimizer_states=True,
                        load_lr_scheduler_states=True):
        """Returns (load_path, client_state); (None, None) when nothing could be loaded."""
        if tag is None:
            latest_path = os.path.join(load_dir, "latest")
            if os.path.isfile(latest_path):
                with open(latest_path, "r") as fd:
                    tag = fd.read().strip()
            else:
                logger.warning(f"Unable to find latest file at {latest_path}, if trying to load latest "
                               "checkpoint please ensure this file exists or pass an explicit checkpoint tag when loading a checkpoint.")
                return None, None
        load_path, client_states = self._load_checkpoint(load_dir, tag, load_module_strict=load_module_strict,
                                                         load_optimizer_states=load_optimizer_states,
                                                         load_lr_scheduler_states=load_lr_scheduler_states)
        if self.zero_optimization() and load_path is not None:
            self._load_zero_checkpoint(load_dir, tag, load_optimizer_states=load_optimizer_states)
        return load_path, client_states

    def _load_checkpoint(self, load_dir, tag, load_module_strict=True, load_optimizer_states=True,
                         load_lr_scheduler_states=True):
        load_path = self._get_ckpt_name(load_dir, tag)
        if not os.path.exists(load_path):
            logger.warning("Client provided checkpoint load path: {} does not exist ... skip checkpoint load".format(load_path))
            return None, None
        logger.info(f"rank: {self.global_rank} loading checkpoint: {load_path}")
        checkpoint = _torch_load(load_path, map_location=lambda storage, loc: storage)
        from .pipe.engine import PipelineEngine
        if isinstance(self, PipelineEngine):
            # pipeline parallelism keeps one file per layer next to the model-states file
            self._curr_ckpt_path = os.path.join(load_dir, tag)
        if self.zero_optimization_partition_weights():
            # the file holds shard placeholders: bring every parameter (also the persistent ones) to that state;
            # the real values come from the fp32 shards in the optimizer checkpoint
            self.optimizer.save_checkpoint_prologue()
        self.load_module_state_dict(state_dict=checkpoint["module"], strict=load_module_strict)
        if self.optimizer is not None and not self.zero_optimization():
            if self.fp16_enabled() and self.precision() != torch.float32:
                self.optimizer.load_state_dict(checkpoint["optimizer"], load_optimizer_states=load_optimizer_states)
            elif load_optimizer_states:
                self.optimizer.load_state_dict(checkpoint["optimizer"])
        if load_lr_scheduler_states and self.lr_scheduler is not None and checkpoint.get("lr_scheduler") is not None:
            self.lr_scheduler.load_state_dict(checkpoint["lr_scheduler"])
        self.csr_tensor_module_names = checkpoint["csr_tensor_module_names"]
        self.global_steps = checkpoint["global_steps"]
        self.global_samples = checkpoint.get("global_samples", self.global_steps * self.train_batch_size())
        self.skipped_steps = checkpoint["skipped_steps"]
        self.loaded_checkpoint_mp_world_size = checkpoint["mp_world_size"]
        self.loaded_checkpoint_dp_world_size = checkpoint["dp_world_size"]
        deepspeed_states = ["module", "optimizer", "lr_scheduler", "csr_tensor_module_names", "skipped_steps",
                            "global_steps", "dp_world_size", "mp_world_size"]
        client_state = {key: value for key, value in checkpoint.items() if key not in deepspeed_states}
        return load_path, client_state

    def _load_zero_checkpoint(self, load_dir, tag, load_optimizer_states=True):
        zero_sd_list = self._get_all_zero_checkpoints(load_dir, tag)
        if zero_sd_list is None:
            return
        self.optimizer.load_state_dict(state_dict_list=zero_sd_list, load_optimizer_states=load_optimizer_states,
                                       load_from_fp32_weights=self.zero_load_from_fp32_weights())
        print(f"loading {len(zero_sd_list)} zero partition checkpoints for rank {self.global_rank}")

    def _get_mp_rank_zero_checkpoint_names(self, load_dir, tag, mp_rank, dp_world_size):
        names = []
        for dp_rank in range(dp_world_size):
            ckpt_name = self._get_rank_zero_ckpt_name(checkpoints_path=load_dir, tag=tag, mp_rank=mp_rank, dp_rank=dp_rank)
            if not os.path.exists(ckpt_name):
                # tolerate the legacy name without the underscore before "optim_states"
                legacy = ckpt_name.replace("_optim_states.pt", "optim_states.pt")
                if os.path.exists(legacy):
                    ckpt_name = legacy
            names.append(ckpt_name)
        return names

    def _get_all_zero_checkpoint_names(self, load_dir, tag, mp_world_size, dp_world_size):
        return [self._get_mp_rank_zero_checkpoint_names(load_dir, tag, mp_rank, dp_world_size)
                for mp_rank in range(mp_world_size)]

    def _get_all_zero_checkpoints(self, load_dir, tag):
        mp_rank = 0 if self.mpu is None else self.mpu.get_model_parallel_rank()
        zero_ckpt_names = self._get_mp_rank_zero_checkpoint_names(load_dir=load_dir, tag=tag, mp_rank=mp_rank,
                                                                  dp_world_size=self.loaded_checkpoint_dp_world_size)
        invalid = [n for n in zero_ckpt_names if not os.path.exists(n)]
        if invalid:
            logger.warning(f"Loading checkpoint failed: {invalid} do not exist; ZeRO optimizer state not loaded")
            return None
        zero_sd_list = []
        for ckpt_name in zero_ckpt_names:
            zero_sd_list.append(_torch_load(ckpt_name, map_location="cpu"))
        zero_optimizer_sd = [sd["optimizer_state_dict"] for sd in zero_sd_list]
        print(f"successfully loaded {len(zero_optimizer_sd)} ZeRO state_dicts for rank {self.global_rank}")
        return zero_optimizer_sd

    # ------------------------------------------------------------------ save
    def _checkpoint_tag_validation(self, tag):
        if self.checkpoint_tag_validation_enabled():
            s_hash = hashlib.sha1(tag.encode())
            bhash = torch.ByteTensor([s_hash.digest()]).flatten().to(self.device)
            max_bhash = bhash.clone()
            min_bhash = bhash.clone()
            dist.all_reduce(max_bhash, op=torch.distributed.ReduceOp.MAX)
            dist.all_reduce(min_bhash, op=torch.distributed.ReduceOp.MIN)
            valid = all(min_bhash == bhash) and all(max_bhash == bhash)
            msg = (f"[rank={dist.get_rank()}] The checkpoint tag name '{tag}' is not consistent across "
                   "all ranks. Including rank unique information in checkpoint tag could cause issues when "
                   "restoring with different world sizes.")
            if self.checkpoint_tag_validation_fail():
                assert valid, msg
            elif not valid:
                logger.warning(msg)

    def save_checkpoint(self, save_dir, tag=None, client_state={}, save_latest=True):
        """Must be called on ALL ranks. Returns True."""
        if self.zero_optimization_partition_weights():
            self.optimizer.save_checkpoint_prologue()
        if tag is None:
            tag = f"global_step{self.global_steps}"
        tag = str(tag)
        self._checkpoint_tag_validation(tag)
        if self.save_non_zero_checkpoint:
            self._create_checkpoint_file(save_dir, tag, False)
            self._save_checkpoint(save_dir, tag, client_state=client_state)
        if self.save_zero_checkpoint:
            self._create_zero_checkpoint_files(save_dir, tag)
            self._save_zero_checkpoint(save_dir, tag)
        dist.barrier()
        if save_latest and self.global_rank == 0:
            with open(os.path.join(save_dir, "latest"), "w") as fd:
                fd.write(tag)
        if self.zero_optimization_partition_weights():
            self.optimizer.save_checkpoint_epilogue()
        return True

    def _create_checkpoint_file(self, save_dir, tag, zero_checkpoint):
        name_function = self._get_zero_ckpt_name if zero_checkpoint else self._get_ckpt_name
        try:
            checkpoint_name = name_function(save_dir, tag)
            os.makedirs(os.path.dirname(checkpoint_name), exist_ok=True)
        except Exception:
            logger.error(f"Failed saving model checkpoint to {save_dir} with tag {tag}")
            return False
        return True

    def _create_zero_checkpoint_files(self, save_dir, tag):
        success = True
        for rank in range(self.world_size):
            if rank == self.global_rank:
                success = self._create_checkpoint_file(save_dir, tag, True)
            dist.barrier()
        return success

    def _save_checkpoint(self, save_dir, tag, client_state={}):
        save_path = self._get_ckpt_name(save_dir, tag)
        # module_state_dict() of the pipeline engine writes per-layer files relative to this directory
        self._curr_ckpt_path = os.path.join(save_dir, tag)
        state = dict(
            module=self.module_state_dict(),
            optimizer=self.optimizer.state_dict() if self.optimizer and not self.zero_optimization() else None,
            lr_scheduler=self.lr_scheduler.state_dict() if self.lr_scheduler is not None else None,
            csr_tensor_module_names=self.csr_tensor_module_names,
            skipped_steps=self.skipped_steps,
            global_steps=self.global_steps,
            global_samples=self.global_samples,
            dp_world_size=self.dp_world_size,
            mp_world_size=self.mp_world_size,
        )
        state.update(client_state)
        log_dist(message=f"Saving model checkpoint: {save_path}", ranks=[0])
        torch.save(state, save_path)
        self._curr_save_path = None

    def _get_param_shapes(self):
        from collections import OrderedDict
        param_shapes = OrderedDict()
        for name, param in self.module.named_parameters():
            param_shapes[name] = param.ds_shape if hasattr(param, "ds_shape") else param.shape
        return param_shapes

    def _copy_recovery_script(self, save_path):
        base_dir = os.path.dirname(os.path.dirname(__file__))
        script = "zero_to_fp32.py"
        src = os.path.join(base_dir, "utils", script)
        dst = os.path.join(save_path, script)
        try:
            copyfile(src, dst)
            os.chmod(dst, os.stat(dst).st_mode | 0o111)
        except Exception as e:  # pragma: no cover
            logger.warning(f"could not copy {script}: {e}")

    def _save_zero_checkpoint(self, save_path, tag):
        zero_checkpoint_name = self._get_zero_ckpt_name(save_path, tag)
        zero_sd = dict(optimizer_state_dict=self.optimizer.state_dict(), param_shapes=self._get_param_shapes())
        torch.save(zero_sd, zero_checkpoint_name)
        if self.global_rank == 0:
            self._copy_recovery_script(os.path.join(save_path, str(tag)))
        logger.info("zero checkpoint saved {}".format(zero_checkpoint_name))

    # ------------------------------------------------------------------ consolidated fp16 export
    def _zero3_consolidated_fp16_state_dict(self):
        """Gather ZeRO-3 partitioned weights layer by layer onto rank 0 (CPU) and return a regular state_dict."""
        from collections import OrderedDict
        from .zero.partition_parameters import GatheredParameters
        if not self.zero_optimization_partition_weights():
            raise ValueError("this function requires ZeRO-3 mode")
        state_dict = OrderedDict() if dist.get_rank() == 0 else None
        shared_weights = {}

        def get_layer_state_dict(module, prefix=""):
            with GatheredParameters(list(module.parameters(recurse=False))):
                if dist.get_rank() == 0:
                    for name, param in module.named_parameters(recurse=False):
                        if param is None:
                            continue
                        key = prefix + name
                        if getattr(param, "ds_id", None) in shared_weights:
                            state_dict[key] = state_dict[shared_weights[param.ds_id]]
                        else:
                            state_dict[key] = param.detach().cpu()
                            shared_weights[getattr(param, "ds_id", id(param))] = key
                    for name, buf in module.named_buffers(recurse=False):
                        if buf is not None and name not in module._non_persistent_buffers_set:
                            state_dict[prefix + name] = buf.detach().cpu()
            for name, child in module.named_children():
                if child is not None:
                    get_layer_state_dict(child, prefix + name + ".")

        get_layer_state_dict(self.module, prefix="")
        return state_dict

    def save_fp16_model(self, save_dir, save_filename="pytorch_model.bin"):
        """Save low-precision weights loadable with plain torch.load; consolidates ZeRO-3 shards on rank 0."""
        path = os.path.join(save_dir, save_filename)
        if self.zero_optimization_partition_weights():
            if self.zero_gather_fp16_weights_on_model_save():
                state_dict = self._zero3_consolidated_fp16_state_dict()
            else:
                logger.info(f"Did not save the model {path} because `stage3_gather_fp16_weights_on_model_save` is False")
                return
        else:
            state_dict = self.module.state_dict()
        if dist.get_rank() == 0:
            os.makedirs(save_dir, exist_ok=True)
            logger.info(f"Saving model weights to {path}")
            torch.save(state_dict, path)
