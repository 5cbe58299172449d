"""Delay-pattern apply / revert, closed form (test infrastructure; see oracle/__init__.py).

Reference: zonos/codebook_pattern.py:5-7 (apply: right-pad each row by n_q with mask, roll row k by k+1)
and :10-12 (revert: row k is the slice [k+1 : T_p - n_q + k + 1]).
"""
import numpy as np
import torch


def apply_delay_pattern_np(codes: np.ndarray, mask_token: int) -> np.ndarray:
    """codes int64 [B, Q, T] -> [B, Q, T+Q];  out[b,k,t] = codes[b,k,t-(k+1)] if 0 <= t-(k+1) < T else mask."""
    B, Q, T = codes.shape
    out = np.full((B, Q, T + Q), mask_token, dtype=codes.dtype)
    for k in range(Q):
        out[:, k, k + 1:k + 1 + T] = codes[:, k, :]
    return out


def revert_delay_pattern_np(delayed: np.ndarray) -> np.ndarray:
    """delayed [B, Q, Tp] -> [B, Q, Tp-Q];  out[b,k,t] = delayed[b,k,t+k+1]."""
    B, Q, Tp = delayed.shape
    out = np.empty((B, Q, Tp - Q), dtype=delayed.dtype)
    for k in range(Q):
        out[:, k, :] = delayed[:, k, k + 1:k + 1 + Tp - Q]
    return out


def apply_delay_pattern(codes: torch.Tensor, mask_token: int) -> torch.Tensor:
    return torch.from_numpy(apply_delay_pattern_np(codes.cpu().numpy(), mask_token)).to(codes.device)


def revert_delay_pattern(delayed: torch.Tensor) -> torch.Tensor:
    return torch.from_numpy(revert_delay_pattern_np(delayed.cpu().numpy())).to(delayed.device)
