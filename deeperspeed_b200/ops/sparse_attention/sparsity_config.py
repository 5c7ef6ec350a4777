"""Block-sparsity layouts: Dense, Fixed, Variable, BigBird, BSLongformer, LocalSlidingWindow.

A layout is an int64 tensor [num_heads, seq/block, seq/block] of 0/1. Patterns and constructor arguments follow
reference deepspeed/ops/sparse_attention/sparsity_config.py:29-740; layouts are assembled with tensor
slicing instead of per-element Python loops."""
import random

import torch


class SparsityConfig:
    """Abstract configuration: number of heads, block size and whether heads get different layouts."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False):
        self.num_heads = num_heads
        self.block = block
        self.different_layout_per_head = different_layout_per_head
        self.num_layout_heads = num_heads if different_layout_per_head else 1

    def setup_layout(self, seq_len):
        if seq_len % self.block != 0:
            raise ValueError(f"Sequence Length, {seq_len}, needs to be dividable by Block size {self.block}!")
        num_blocks = seq_len // self.block
        return torch.zeros((self.num_heads, num_blocks, num_blocks), dtype=torch.int64)

    def check_and_propagate_first_head_layout(self, layout):
        if not self.different_layout_per_head:
            layout[1:self.num_heads, :, :] = layout[0, :, :]
        return layout


class DenseSparsityConfig(SparsityConfig):
    """All blocks present (for comparison / debugging)."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False):
        super().__init__(num_heads, block, different_layout_per_head)

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        layout[:, :, :] = 1
        return layout


def _check_attention(attention):
    if attention not in ("unidirectional", "bidirectional"):
        raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')


class FixedSparsityConfig(SparsityConfig):
    """"Fixed" pattern of Sparse Transformers: dense local windows of `num_local_blocks` blocks plus global columns
    taken from the tail of every window (optionally rotated per head and mirrored horizontally)."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_local_blocks=4, num_global_blocks=1,
                 attention="bidirectional", horizontal_global_attention=False, num_different_global_patterns=1):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_local_blocks = num_local_blocks
        if num_local_blocks % num_global_blocks != 0:
            raise ValueError(f"Number of blocks in a local window, {num_local_blocks}, must be dividable by number of "
                             f"global blocks, {num_global_blocks}!")
        self.num_global_blocks = num_global_blocks
        _check_attention(attention)
        self.attention = attention
        if attention != "bidirectional" and horizontal_global_attention:
            raise ValueError('only "bi-directional" attentions can support horizontal global attention!')
        self.horizontal_global_attention = horizontal_global_attention
        if num_different_global_patterns > 1 and not different_layout_per_head:
            raise ValueError("Number of different layouts cannot be more than one when you have set a single layout "
                             "for all heads! Set different_layout_per_head to True.")
        if num_different_global_patterns > (num_local_blocks // num_global_blocks):
            raise ValueError(f"Number of layout versions (num_different_global_patterns), "
                             f"{num_different_global_patterns}, cannot be larger than number of local window blocks "
                             f"divided by number of global blocks, {num_local_blocks} / {num_global_blocks} = "
                             f"{num_local_blocks // num_global_blocks}!")
        self.num_different_global_patterns = num_different_global_patterns

    def set_local_layout(self, h, layout):
        num_blocks = layout.shape[1]
        for start in range(0, num_blocks, self.num_local_blocks):
            end = min(start + self.num_local_blocks, num_blocks)
            win = torch.ones(end - start, end - start, dtype=layout.dtype)
            if self.attention == "unidirectional":
                win = torch.tril(win)
            layout[h, start:end, start:end] = win
        return layout

    def set_global_layout(self, h, layout):
        num_blocks = layout.shape[1]
        first_global_block_idx = (self.num_local_blocks - (1 + h % self.num_different_global_patterns) * self.num_global_blocks)
        end = num_blocks - (num_blocks % self.num_local_blocks)
        for i in range(first_global_block_idx, end, self.num_local_blocks):
            # vertical: every (later, if causal) row sees the window's global blocks
            first_row = 0 if self.attention == "bidirectional" else i
            layout[h, first_row:, i:i + self.num_global_blocks] = 1
            if self.horizontal_global_attention:
                layout[h, i:i + self.num_global_blocks, :] = 1
        # a trailing short window contributes its own tail as global blocks
        if end < num_blocks:
            start = min(end + first_global_block_idx, num_blocks - self.num_global_blocks)
            stop = start + self.num_global_blocks
            first_row = 0 if self.attention == "bidirectional" else start
            layout[h, first_row:, start:stop] = 1
            if self.horizontal_global_attention:
                layout[h, start:stop, :] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(0, self.num_layout_heads):
            layout = self.set_local_layout(h, layout)
            layout = self.set_global_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class VariableSparsityConfig(SparsityConfig):
    """Random blocks + variable-size local windows (last size repeats) + explicit global block indices / ranges."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=0,
                 local_window_blocks=[4], global_block_indices=[0], global_block_end_indices=None,
                 attention="bidirectional", horizontal_global_attention=False):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_random_blocks = num_random_blocks
        self.local_window_blocks = local_window_blocks
        self.global_block_indices = global_block_indices
        if global_block_end_indices is not None:
            if len(global_block_indices) != len(global_block_end_indices):
                raise ValueError(f"Global block start indices length, {len(global_block_indices)}, must be same as "
                                 f"global block end indices length, {len(global_block_end_indices)}!")
            for start_idx, end_idx in zip(global_block_indices, global_block_end_indices):
                if start_idx >= end_idx:
                    raise ValueError(f"Global block start index, {start_idx}, must be smaller than global block end "
                                     f"index, {end_idx}!")
        self.global_block_end_indices = global_block_end_indices
        _check_attention(attention)
        self.attention = attention
        if attention != "bidirectional" and horizontal_global_attention:
            raise ValueError('only "bi-directional" attentions can support horizontal global attention!')
        self.horizontal_global_attention = horizontal_global_attention

    def set_random_layout(self, h, layout):
        num_blocks = layout.shape[1]
        if num_blocks < self.num_random_blocks:
            raise ValueError(f"Number of random blocks, {self.num_random_blocks}, must be smaller than overal number "
                             f"of blocks in a row, {num_blocks}!")
        for row in range(0, num_blocks):
            rnd_cols = random.sample(range(0, num_blocks), self.num_random_blocks)
            layout[h, row, rnd_cols] = 1
        return layout

    def set_local_layout(self, h, layout):
        num_blocks = layout.shape[1]
        start = 0
        sizes = list(self.local_window_blocks)
        idx = 0
        while start < num_blocks:
            size = sizes[min(idx, len(sizes) - 1)]
            end = min(start + size, num_blocks)
            win = torch.ones(end - start, end - start, dtype=layout.dtype)
            if self.attention == "unidirectional":
                win = torch.tril(win)
            layout[h, start:end, start:end] = win
            start = end
            idx += 1
        return layout

    def set_global_layout(self, h, layout):
        num_blocks = layout.shape[1]
        if self.global_block_end_indices is None:
            ranges = [(i, i + 1) for i in self.global_block_indices]
        else:
            ranges = list(zip(self.global_block_indices, self.global_block_end_indices))
        for start_idx, end_idx in ranges:
            if start_idx < num_blocks:
                end_idx = min(end_idx, num_blocks)
                if self.horizontal_global_attention:
                    layout[h, start_idx:end_idx, :] = 1
                first_row = 0 if self.attention == "bidirectional" else start_idx
                layout[h, first_row:, start_idx:end_idx] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(0, self.num_layout_heads):
            layout = self.set_random_layout(h, layout)
            layout = self.set_local_layout(h, layout)
            layout = self.set_global_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class BigBirdSparsityConfig(SparsityConfig):
    """BigBird: random blocks + sliding window + global first blocks (rows and columns)."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=1,
                 num_sliding_window_blocks=3, num_global_blocks=1, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_random_blocks = num_random_blocks
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.num_global_blocks = num_global_blocks
        _check_attention(attention)
        self.attention = attention

    def set_random_layout(self, h, layout):
        num_blocks = layout.shape[1]
        if num_blocks < self.num_random_blocks:
            raise ValueError(f"Number of random blocks, {self.num_random_blocks}, must be smaller than overal number "
                             f"of blocks in a row, {num_blocks}!")
        for row in range(0, num_blocks):
            sample_range = range(0, num_blocks) if self.attention == "bidirectional" else range(0, row + 1)
            rnd_cols = random.sample(sample_range, min(self.num_random_blocks, len(sample_range)))
            layout[h, row, rnd_cols] = 1
        return layout

    def set_sliding_window_layout(self, h, layout):
        num_blocks = layout.shape[1]
        if num_blocks < self.num_sliding_window_blocks:
            raise ValueError(f"Number of sliding window blocks, {self.num_sliding_window_blocks}, must be smaller "
                             f"than overal number of blocks in a row, {num_blocks}!")
        w = self.num_sliding_window_blocks // 2
        idx = torch.arange(num_blocks)
        band = (idx[:, None] - idx[None, :]).abs() <= w
        if self.attention == "unidirectional":
            band = band & (idx[None, :] <= idx[:, None])
        layout[h][band] = 1
        return layout

    def set_global_layout_itc(self, h, layout):
        num_blocks = layout.shape[1]
        if num_blocks < self.num_global_blocks:
            raise ValueError(f"Number of global blocks, {self.num_global_blocks}, must be smaller than overal number "
                             f"of blocks in a row, {num_blocks}!")
        layout[h, 0:self.num_global_blocks, :] = 1
        layout[h, :, 0:self.num_global_blocks] = 1
        if self.attention == "unidirectional":
            layout[h] = torch.tril(layout[h])
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(0, self.num_layout_heads):
            layout = self.set_random_layout(h, layout)
            layout = self.set_sliding_window_layout(h, layout)
            layout = self.set_global_layout_itc(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class BSLongformerSparsityConfig(SparsityConfig):
    """Block-sparse Longformer: sliding window + explicit global rows/columns."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_sliding_window_blocks=3,
                 global_block_indices=[0], global_block_end_indices=None, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.global_block_indices = global_block_indices
        self.attention = attention
        if global_block_end_indices is not None:
            if len(global_block_indices) != len(global_block_end_indices):
                raise ValueError(f"Global block start indices length, {len(global_block_indices)}, must be same as "
                                 f"global block end indices length, {len(global_block_end_indices)}!")
            for start_idx, end_idx in zip(global_block_indices, global_block_end_indices):
                if start_idx >= end_idx:
                    raise ValueError(f"Global block start index, {start_idx}, must be smaller than global block end "
                                     f"index, {end_idx}!")
        self.global_block_end_indices = global_block_end_indices

    def set_sliding_window_layout(self, h, layout):
        num_blocks = layout.shape[1]
        if num_blocks < self.num_sliding_window_blocks:
            raise ValueError(f"Number of sliding window blocks, {self.num_sliding_window_blocks}, must be smaller "
                             f"than overal number of blocks in a row, {num_blocks}!")
        w = self.num_sliding_window_blocks // 2
        idx = torch.arange(num_blocks)
        layout[h][(idx[:, None] - idx[None, :]).abs() <= w] = 1
        return layout

    def set_global_layout(self, h, layout):
        num_blocks = layout.shape[1]
        if self.global_block_end_indices is None:
            ranges = [(i, i + 1) for i in self.global_block_indices]
        else:
            ranges = list(zip(self.global_block_indices, self.global_block_end_indices))
        for start_idx, end_idx in ranges:
            if start_idx < num_blocks:
                end_idx = min(end_idx, num_blocks)
                layout[h, start_idx:end_idx, :] = 1
                layout[h, :, start_idx:end_idx] = 1
        if self.attention == "unidirectional":
            layout[h] = torch.tril(layout[h])
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(0, self.num_layout_heads):
            layout = self.set_sliding_window_layout(h, layout)
            layout = self.set_global_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class LocalSlidingWindowSparsityConfig(SparsityConfig):
    """Pure local sliding window (DeeperSpeed addition)."""

    def __init__(self, num_heads, block=16, num_sliding_window_blocks=3, attention="unidirectional"):
        super().__init__(num_heads, block)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.attention = attention

    def set_sliding_window_layout(self, h, layout):
        num_blocks = layout.shape[1]
        if num_blocks < self.num_sliding_window_blocks:
            raise ValueError(f"Number of sliding window blocks, {self.num_sliding_window_blocks}, must be smaller "
                             f"than overal number of blocks in a row, {num_blocks}!")
        w = self.num_sliding_window_blocks // 2
        idx = torch.arange(num_blocks)
        diff = idx[:, None] - idx[None, :]
        band = (diff >= 0) & (diff <= w) if self.attention == "unidirectional" else diff.abs() <= w
        layout[h][band] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(0, self.num_layout_heads):
            layout = self.set_sliding_window_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)
