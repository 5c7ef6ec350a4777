"""Block-sparsity layouts: Dense, Fixed, Variable, BigBird, BSLongformer, LocalSlidingWindow.

A layout is an int64 tensor [num_heads, seq/block, seq/block] of 0/1. Class names, constructor arguments, method names
and the patterns themselves follow reference deepspeed/ops/sparse_attention/sparsity_config.py. Here a
pattern is a composition of a few tensor-level primitives (diagonal windows, bands, global stripes, random picks)
written once below; each class lists the pattern methods it applies per head in `_patterns`, and the methods are
one-liners over the primitives.
"""
import random

import torch


# ------------------------------------------------------------------------------------------------ primitives
def _need_blocks(what, count, num_blocks):
    if num_blocks < count:
        raise ValueError(f"Number of {what} blocks, {count}, must be smaller than overal number of blocks in a row, "
                         f"{num_blocks}!")


def _check_attention(attention):
    if attention not in ("unidirectional", "bidirectional"):
        raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')


def _check_horizontal(attention, horizontal):
    if attention != "bidirectional" and horizontal:
        raise ValueError('only "bi-directional" attentions can support horizontal global attention!')


def _check_global_ranges(starts, ends):
    if ends is None:
        return
    if len(starts) != len(ends):
        raise ValueError(f"Global block start indices length, {len(starts)}, must be same as global block end indices "
                         f"length, {len(ends)}!")
    for lo, hi in zip(starts, ends):
        if lo >= hi:
            raise ValueError(f"Global block start index, {lo}, must be smaller than global block end index, {hi}!")


def _global_ranges(starts, ends):
    return [(i, i + 1) for i in starts] if ends is None else list(zip(starts, ends))


def _diagonal_windows(head, sizes, causal):
    """Dense square windows along the diagonal; `sizes` yields window widths (the last one repeats)."""
    n, at, k = head.shape[0], 0, 0
    while at < n:
        stop = min(at + sizes[min(k, len(sizes) - 1)], n)
        window = torch.ones(stop - at, stop - at, dtype=head.dtype)
        head[at:stop, at:stop] = torch.tril(window) if causal else window
        at, k = stop, k + 1


def _band(head, half_width, mode):
    """|row - col| <= half_width; mode "both", "causal" (only col <= row) ."""
    idx = torch.arange(head.shape[0])
    diff = idx[:, None] - idx[None, :]
    head[(diff >= 0) & (diff <= half_width) if mode == "causal" else diff.abs() <= half_width] = 1


def _column_stripe(head, lo, hi, causal, horizontal=False):
    """Blocks [lo, hi) are global: every row (causal: every row from `lo` on) sees them; optionally they see everyone."""
    head[(lo if causal else 0):, lo:hi] = 1
    if horizontal:
        head[lo:hi, :] = 1


def _random_blocks(head, per_row, causal):
    n = head.shape[0]
    _need_blocks("random", per_row, n)
    for row in range(n):
        pool = range(0, row + 1) if causal else range(0, n)
        head[row, random.sample(pool, min(per_row, len(pool)))] = 1


# ------------------------------------------------------------------------------------------------ configurations
class SparsityConfig:
    """Abstract configuration: number of heads, block size and whether heads get different layouts."""

    _patterns = ()      # names of the methods (h, layout) -> layout applied per head, in order

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

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            for name in self._patterns:
                layout = getattr(self, name)(h, layout)
        return self.check_and_propagate_first_head_layout(layout)

    def _causal(self):
        return getattr(self, "attention", "bidirectional") == "unidirectional"


class DenseSparsityConfig(SparsityConfig):
    """All blocks present (for comparison / debugging)."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False):
        super().__init__(num_heads, block, different_layout_per_head)

    def make_layout(self, seq_len):
        return self.setup_layout(seq_len).fill_(1)


class FixedSparsityConfig(SparsityConfig):
    """"Fixed" pattern of Sparse Transformers: dense local windows of `num_local_blocks` blocks plus global columns
    taken from the tail of every window (optionally rotated per head and mirrored horizontally)."""

    _patterns = ("set_local_layout", "set_global_layout")

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_local_blocks=4, num_global_blocks=1,
                 attention="bidirectional", horizontal_global_attention=False, num_different_global_patterns=1):
        super().__init__(num_heads, block, different_layout_per_head)
        if num_local_blocks % num_global_blocks != 0:
            raise ValueError(f"Number of blocks in a local window, {num_local_blocks}, must be dividable by number of "
                             f"global blocks, {num_global_blocks}!")
        _check_attention(attention)
        _check_horizontal(attention, horizontal_global_attention)
        if num_different_global_patterns > 1 and not different_layout_per_head:
            raise ValueError("Number of different layouts cannot be more than one when you have set a single layout "
                             "for all heads! Set different_layout_per_head to True.")
        if num_different_global_patterns > (num_local_blocks // num_global_blocks):
            raise ValueError(f"Number of layout versions (num_different_global_patterns), "
                             f"{num_different_global_patterns}, cannot be larger than number of local window blocks "
                             f"divided by number of global blocks, {num_local_blocks} / {num_global_blocks} = "
                             f"{num_local_blocks // num_global_blocks}!")
        self.num_local_blocks, self.num_global_blocks = num_local_blocks, num_global_blocks
        self.attention, self.horizontal_global_attention = attention, horizontal_global_attention
        self.num_different_global_patterns = num_different_global_patterns

    def set_local_layout(self, h, layout):
        _diagonal_windows(layout[h], [self.num_local_blocks], self._causal())
        return layout

    def set_global_layout(self, h, layout):
        n, width, window = layout.shape[1], self.num_global_blocks, self.num_local_blocks
        # the global blocks of a window: its last `width` blocks, moved one step earlier per head pattern
        offset = window - (1 + h % self.num_different_global_patterns) * width
        whole = n - n % window
        stripes = [(lo, lo + width) for lo in range(offset, whole, window)]
        if whole < n:      # a trailing short window contributes its own tail
            lo = min(whole + offset, n - width)
            stripes.append((lo, lo + width))
        for lo, hi in stripes:
            _column_stripe(layout[h], lo, hi, self._causal(), self.horizontal_global_attention)
        return layout


class VariableSparsityConfig(SparsityConfig):
    """Random blocks + variable-size local windows (last size repeats) + explicit global block indices / ranges."""

    _patterns = ("set_random_layout", "set_local_layout", "set_global_layout")

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=0,
                 local_window_blocks=[4], global_block_indices=[0], global_block_end_indices=None,
                 attention="bidirectional", horizontal_global_attention=False):
        super().__init__(num_heads, block, different_layout_per_head)
        _check_global_ranges(global_block_indices, global_block_end_indices)
        _check_attention(attention)
        _check_horizontal(attention, horizontal_global_attention)
        self.num_random_blocks, self.local_window_blocks = num_random_blocks, local_window_blocks
        self.global_block_indices, self.global_block_end_indices = global_block_indices, global_block_end_indices
        self.attention, self.horizontal_global_attention = attention, horizontal_global_attention

    def set_random_layout(self, h, layout):
        _random_blocks(layout[h], self.num_random_blocks, causal=False)
        return layout

    def set_local_layout(self, h, layout):
        _diagonal_windows(layout[h], list(self.local_window_blocks), self._causal())
        return layout

    def set_global_layout(self, h, layout):
        n = layout.shape[1]
        for lo, hi in _global_ranges(self.global_block_indices, self.global_block_end_indices):
            if lo < n:
                _column_stripe(layout[h], lo, min(hi, n), self._causal(), self.horizontal_global_attention)
        return layout


class BigBirdSparsityConfig(SparsityConfig):
    """BigBird: random blocks + sliding window + global first blocks (rows and columns)."""

    _patterns = ("set_random_layout", "set_sliding_window_layout", "set_global_layout_itc")

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=1,
                 num_sliding_window_blocks=3, num_global_blocks=1, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        _check_attention(attention)
        self.num_random_blocks, self.num_sliding_window_blocks = num_random_blocks, num_sliding_window_blocks
        self.num_global_blocks, self.attention = num_global_blocks, attention

    def set_random_layout(self, h, layout):
        _random_blocks(layout[h], self.num_random_blocks, self._causal())
        return layout

    def set_sliding_window_layout(self, h, layout):
        _need_blocks("sliding window", self.num_sliding_window_blocks, layout.shape[1])
        _band(layout[h], self.num_sliding_window_blocks // 2, "causal" if self._causal() else "both")
        return layout

    def set_global_layout_itc(self, h, layout):
        _need_blocks("global", self.num_global_blocks, layout.shape[1])
        _column_stripe(layout[h], 0, self.num_global_blocks, causal=False, horizontal=True)
        if self._causal():
            layout[h] = torch.tril(layout[h])
        return layout


class BSLongformerSparsityConfig(SparsityConfig):
    """Block-sparse Longformer: sliding window + explicit global rows/columns."""

    _patterns = ("set_sliding_window_layout", "set_global_layout")

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_sliding_window_blocks=3,
                 global_block_indices=[0], global_block_end_indices=None, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        _check_global_ranges(global_block_indices, global_block_end_indices)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.global_block_indices, self.global_block_end_indices = global_block_indices, global_block_end_indices
        self.attention = attention

    def set_sliding_window_layout(self, h, layout):
        _need_blocks("sliding window", self.num_sliding_window_blocks, layout.shape[1])
        _band(layout[h], self.num_sliding_window_blocks // 2, "both")
        return layout

    def set_global_layout(self, h, layout):
        n = layout.shape[1]
        for lo, hi in _global_ranges(self.global_block_indices, self.global_block_end_indices):
            if lo < n:
                _column_stripe(layout[h], lo, min(hi, n), causal=False, horizontal=True)
        if self._causal():
            layout[h] = torch.tril(layout[h])
        return layout


class LocalSlidingWindowSparsityConfig(SparsityConfig):
    """Pure local sliding window (DeeperSpeed addition)."""

    _patterns = ("set_sliding_window_layout",)

    def __init__(self, num_heads, block=16, num_sliding_window_blocks=3, attention="unidirectional"):
        super().__init__(num_heads, block)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.attention = attention

    def set_sliding_window_layout(self, h, layout):
        _need_blocks("sliding window", self.num_sliding_window_blocks, layout.shape[1])
        _band(layout[h], self.num_sliding_window_blocks // 2, "causal" if self._causal() else "both")
        return layout
