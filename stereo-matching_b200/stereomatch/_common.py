"""Output-buffer reuse rules (API of reference stereomatch/_common.py:9-53): a caller-
supplied tensor is reused iff shape, dtype and device all match; otherwise a new one is
allocated."""
from typing import Optional

import torch


def _reusable(t: Optional[torch.Tensor], sizes, dtype, device) -> bool:
    return (t is not None and tuple(t.size()) == tuple(sizes) and t.dtype == dtype
            and t.device == torch.device(device) and t.is_contiguous())


def empty_tensor(*sizes, dtype: torch.dtype = torch.float32,
                 reuse_tensor: Optional[torch.Tensor] = None,
                 device: torch.device = torch.device("cpu")) -> torch.Tensor:
    """`reuse_tensor` if compatible, else a fresh uninitialised tensor."""
    if _reusable(reuse_tensor, sizes, dtype, device):
        return reuse_tensor
    return torch.empty(sizes, dtype=dtype, device=device)


def zeros_tensor_like(base_tensor: torch.Tensor,
                      reuse_tensor: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`reuse_tensor` zero-filled if compatible with `base_tensor`, else torch.zeros_like."""
    if _reusable(reuse_tensor, base_tensor.size(), base_tensor.dtype, base_tensor.device):
        return reuse_tensor.zero_()
    return torch.zeros_like(base_tensor, memory_format=torch.contiguous_format)
