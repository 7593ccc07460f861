"""Symmetric-memory arena: one VMM allocation per rank, mapped into every peer's address space (and, where the
fabric supports NVLS, into a multicast address), obtained through `torch.distributed._symmetric_memory`.
Kernels receive raw peer pointers and issue plain ld/st (or multimem.*) over NVLink — no NCCL on those paths."""
from __future__ import annotations

import ctypes
from typing import Dict, List, Tuple

import torch
import torch.distributed as dist


def _al(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


class SymmArena:
    def __init__(self, regions: List[Tuple[str, int, torch.dtype]], device: torch.device, group=None):
        """regions: (name, numel, dtype). All ranks must pass identical region lists (collective constructor)."""
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.offsets: Dict[str, Tuple[int, int, torch.dtype]] = {}
        off = 0
        for name, numel, dtype in regions:
            nbytes = numel * torch.empty((), dtype=dtype).element_size()
            self.offsets[name] = (off, numel, dtype)
            off = _al(off + nbytes)
        self.nbytes = max(off, 256)
        self.buf = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        try:
            self.hdl = symm_mem.rendezvous(self.buf, self.group)
        except Exception:
            symm_mem.enable_symm_mem_for_group(self.group.group_name)
            self.hdl = symm_mem.rendezvous(self.buf, self.group.group_name)
        self.rank, self.world = self.hdl.rank, self.hdl.world_size
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        try:
            self.mc_ptr = int(self.hdl.multicast_ptr or 0)
        except Exception:
            self.mc_ptr = 0
        dist.barrier(group=self.group)

    def view(self, name: str) -> torch.Tensor:
        off, numel, dtype = self.offsets[name]
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        return self.buf[off:off + nbytes].view(dtype)

    def peer_view(self, name: str, peer: int) -> torch.Tensor:
        """A tensor aliasing region `name` in rank `peer`'s arena (P2P-mapped)."""
        off, numel, dtype = self.offsets[name]
        esz = torch.empty((), dtype=dtype).element_size()
        return self.hdl.get_buffer(peer, (numel,), dtype, off // esz)

    def ptr_table(self, name: str):
        off = self.offsets[name][0]
        return (ctypes.c_void_p * self.world)(*[p + off for p in self.ptrs])

    def mc(self, name: str) -> int:
        return self.mc_ptr + self.offsets[name][0] if self.mc_ptr else 0
