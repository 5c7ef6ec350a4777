from .sparsity_config import (SparsityConfig, DenseSparsityConfig, FixedSparsityConfig, VariableSparsityConfig,
                              BigBirdSparsityConfig, BSLongformerSparsityConfig, LocalSlidingWindowSparsityConfig)
from .softmax import Softmax
from .matmul import MatMul
from .sparse_self_attention import SparseSelfAttention
from .bert_sparse_self_attention import BertSparseSelfAttention
from .sparse_attention_utils import SparseAttentionUtils
