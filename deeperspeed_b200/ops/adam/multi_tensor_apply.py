"""`MultiTensorApply` under the reference's import path (deepspeed/ops/adam/multi_tensor_apply.py:10-15): calls a
multi-tensor op with (chunk_size, noop_flag, tensor_lists, *args). Our fused optimizers work on flat arenas and on
pointer tables built natively (ops/adam/fused_adam.py), so this thin caller exists for code written against apex's
calling convention."""


class MultiTensorApply(object):
    def __init__(self, chunk_size):
        self.chunk_size = chunk_size

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
