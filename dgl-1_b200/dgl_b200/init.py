"""Feature initializers (python/dgl/init.py:5-61): what fills rows of a column that a partial
write does not touch (Frame.add_column / _append, python/dgl/frame.py:317-400) and the rows of
zero-degree nodes on the degree-bucketing path (NewDictExecutor, runtime/ir/executor.py:765-778).
Same signature as the reference: initializer(shape, dtype, ctx, id_range)."""
import torch

__all__ = ["base_initializer", "zero_initializer"]


def base_initializer(shape, dtype, ctx, id_range):  # pylint: disable=unused-argument
    raise NotImplementedError


def zero_initializer(shape, dtype, ctx, id_range):  # pylint: disable=unused-argument
    return torch.zeros(shape, dtype=dtype, device=ctx)
