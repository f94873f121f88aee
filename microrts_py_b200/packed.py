"""The packed form of the step outputs (include/mrts_b200.h: mrts_step_packed) and its unpacking in plain torch.

  obs_words    (N, H*W) int32         bit k of a cell's word = plane k of the one-hot observation
                                      (GameState.getVectorObservation + _encode_obs, gym_microrts/envs/vec_env.py:311-321)
  mask_entries (N, 1 + cap, 4) int32  row 0 = {n, 0, 0, 0}; rows 1..n = {cell, bits 0-31, 32-63, 64-77} of the mask rows
                                      that are not all zero (JNIGridnetClient.getMasks, vec_env.py:1091-1101)

A policy that consumes the packed form directly reads ~1.4 KB per 16x16 view instead of 110 KB.  ``unpack_obs`` /
``unpack_mask`` give the reference's tensors (they are what the tests compare ``mrts_expand`` and the int32 outputs with)."""
import torch


def unpack_obs(words: torch.Tensor, height: int, width: int, planes: int, dtype=torch.int32) -> torch.Tensor:
    n = words.shape[0]
    k = torch.arange(planes, device=words.device, dtype=torch.int32)
    bits = (words.to(torch.int32).reshape(n, height * width, 1) >> k) & 1
    return bits.to(dtype).reshape(n, height, width, planes)


def unpack_mask(entries: torch.Tensor, cells: int, dtype=torch.int32) -> torch.Tensor:
    n, rows, _ = entries.shape
    e = entries.to(torch.int32)
    count = e[:, 0, 0]
    body = e[:, 1:, :]
    valid = torch.arange(rows - 1, device=e.device).unsqueeze(0) < count.unsqueeze(1)
    k = torch.arange(32, device=e.device, dtype=torch.int32)
    bits = ((body[:, :, 1:4].unsqueeze(-1) >> k) & 1).reshape(n, rows - 1, 96)[:, :, :78]
    out = torch.zeros((n, cells, 78), dtype=dtype, device=e.device)
    env = torch.arange(n, device=e.device).unsqueeze(1).expand(n, rows - 1)
    out[env[valid], body[:, :, 0][valid].long()] = bits[valid].to(dtype)
    return out
