"""All-gather fused into the first consuming GEMM (BASELINE north star "Parameter refresh").

In the synchronous engine every rank owns a contiguous slice of each flat bucket.  For ONE designated Dense kernel
`W[K, N]` (the first GEMM of the forward pass) the bucket kernel does not push the refreshed slice to the peers;
instead the forward GEMM itself (`dfb_ag_gemm_bf16`, csrc/gemm_sm100.cu) pulls the remote slices over NVLink with two
extra warps per CTA while its k-loop starts on the locally owned rows, gated per 64-row k-block by a readiness counter.
This module computes the static description of what to pull ("pieces").
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

import torch

KBLOCK_ROWS = 64
CHUNK = 512


@dataclass
class AgPlan:
    pieces: torch.Tensor        # int64 [P, 4] device: src address, dst address, bytes, k-block
    chunk0: torch.Tensor        # int32 [P + 1] device
    expected: torch.Tensor      # int32 [k-blocks] device (uint32 semantics)
    ready: torch.Tensor         # int32 [k-blocks + 1] device
    num_pieces: int
    rot_kb: int
    remote_bytes: int


def build_plan(K: int, N: int, elem_size: int, w_elem_off: int, shard_elems: int, rank: int, world: int,
               peer_param_bases: Sequence[int], device) -> AgPlan:
    """`w_elem_off`: offset of W inside the bucket (elements); `peer_param_bases[q]`: address of the bucket's parameter
    buffer on rank q (peer-mapped); rank q owns bucket elements [q*shard, (q+1)*shard)."""
    numel = K * N
    kb_bytes = KBLOCK_ROWS * N * elem_size
    nkb = -(-K // KBLOCK_ROWS)
    pieces: List[Tuple[int, int, int, int]] = []
    expected = [0] * nkb
    local_bytes = [0] * nkb
    for q in range(world):
        lo = max(w_elem_off, q * shard_elems)
        hi = min(w_elem_off + numel, (q + 1) * shard_elems)
        if hi <= lo:
            continue
        b0, b1 = (lo - w_elem_off) * elem_size, (hi - w_elem_off) * elem_size       # byte range inside W
        while b0 < b1:
            kb = b0 // kb_bytes
            end = min(b1, (kb + 1) * kb_bytes)
            if q == rank:
                local_bytes[kb] += end - b0
            else:
                off = w_elem_off * elem_size + b0
                pieces.append((peer_param_bases[q] + off, peer_param_bases[rank] + off, end - b0, kb))
                expected[kb] += end - b0
            b0 = end
    rot = next((kb for kb in range(nkb) if local_bytes[kb] > 0 and expected[kb] == 0), None)
    if rot is None:
        rot = next((kb for kb in range(nkb) if local_bytes[kb] > 0), 0)
    chunk0 = [0]
    for p in pieces:
        chunk0.append(chunk0[-1] + -(-p[2] // CHUNK))
    pt = torch.tensor(pieces if pieces else [(0, 0, 0, 0)], dtype=torch.int64, device=device)
    return AgPlan(pt, torch.tensor(chunk0, dtype=torch.int32, device=device),
                  torch.tensor(expected, dtype=torch.int32, device=device),
                  torch.zeros(nkb + 1, dtype=torch.int32, device=device), len(pieces), rot, sum(expected))
