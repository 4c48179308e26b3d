"""Symmetric (peer-mapped, NVSwitch-multicast) memory for the bucket arena.

Plumbing only: torch.distributed._symmetric_memory allocates the buffer with the CUDA VMM API,
exchanges the handles between the ranks of one NVLink domain and maps every rank's copy (and, where
the switch supports it, a multicast alias) into this process; it also provides the cross-GPU barrier.
The bytes are moved by csrc/fused_allreduce.cu."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from . import _lib


class SymmetricArena:
    """A flat buffer that exists at the same offset on every rank of `group`, plus a symmetric
    array of per-chunk partial sums, and the peer map (include/ddp_b200.h: ddpb200_peer_map)."""

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, group, max_chunks: int = 0):
        import torch.distributed._symmetric_memory as symm_mem
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > _lib.MAX_PEERS:
            raise RuntimeError(f"ddp_b200 fused allreduce supports up to {_lib.MAX_PEERS} ranks per NVLink domain")
        self.tensor = symm_mem.empty(max(numel, 1), dtype=dtype, device=device)
        self.tensor.zero_()
        self.hdl = symm_mem.rendezvous(self.tensor, group)
        self.partials = self.phdl = None
        if max_chunks > 0:              # parameter arenas carry no partial sums
            self.partials = symm_mem.empty(max_chunks, dtype=torch.float32, device=device)
            self.partials.zero_()
            self.phdl = symm_mem.rendezvous(self.partials, group)
        self.multicast = int(self.hdl.multicast_ptr or 0)
        mode = os.environ.get("DDPB200_FUSED_MODE", "auto")
        # NVLink bytes per GPU and direction: multicast (1+1/n)S, peer 2(n-1)/n S -> multicast from 4 ranks up
        self.use_multicast = bool(self.multicast) and (mode == "mc" or (mode == "auto" and self.world >= 4))
        self._maps = {}
        self._phase = 0

    def peer_map(self, chunk_off: int = 0) -> "_lib.PeerMap":
        pm = self._maps.get(chunk_off)
        if pm is None:
            pm = _lib.PeerMap()
            for p in range(self.world):
                pm.arena[p] = int(self.hdl.buffer_ptrs[p])
                pm.partials[p] = (int(self.phdl.buffer_ptrs[p]) + 4 * chunk_off) if self.phdl is not None else None
            pm.arena_multicast = self.multicast or None
            pm.rank, pm.world = self.rank, self.world
            self._maps[chunk_off] = pm
        return pm

    def barrier(self) -> None:
        """Cross-GPU barrier on the current stream (a small kernel on every rank)."""
        self.hdl.barrier(channel=self._phase)
        self._phase ^= 1
