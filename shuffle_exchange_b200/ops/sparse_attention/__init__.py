"""Block-sparse self-attention (reference ``deepspeed/ops/sparse_attention``)."""
from .sparsity_config import (SparsityConfig, DenseSparsityConfig, FixedSparsityConfig, VariableSparsityConfig,
                              BigBirdSparsityConfig, BSLongformerSparsityConfig, LocalSlidingWindowSparsityConfig)
from .matmul import MatMul
from .softmax import Softmax
from .sparse_self_attention import SparseSelfAttention
from .bert_sparse_self_attention import BertSparseSelfAttention
from .sparse_attention_utils import SparseAttentionUtils
