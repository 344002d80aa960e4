"""Shared helpers of the -m gpu tests."""
import numpy as np
import torch


def dev(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def host(t: torch.Tensor) -> np.ndarray:
    return t.detach().float().cpu().numpy()


def unaligned(t: torch.Tensor, shift: int = 1) -> torch.Tensor:
    """Same values at an address that is NOT 16-byte aligned (forces the scalar path)."""
    buf = torch.empty(t.numel() + 8, dtype=t.dtype, device=t.device)
    assert buf.data_ptr() % 16 == 0
    v = buf[shift:shift + t.numel()]
    v.copy_(t)
    assert v.data_ptr() % 16 != 0
    return v
