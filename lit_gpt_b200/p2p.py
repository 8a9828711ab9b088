"""Peer-memory communicator for the tensor-parallel decode path (one process per GPU, one node).

The reference sums the row-parallel partial outputs with ``torch.distributed.all_reduce`` from a forward hook
(generate/tp.py:73-74): two NCCL launches per block, 8-16 KB each, pure latency at decode.  Here the exchange is part
of the GEMV kernels (``b200_tp_handoff`` in include/b200gpt.h): the producer's epilogue writes its partial into every
rank's slot over NVLink, the consumer's prologue waits for the arrival flags and adds the slots.  This module only owns
the symmetric buffers: allocation through the library (cudaMalloc + CUDA IPC), handle exchange over the existing
``torch.distributed`` group, and the per-launch ``TpHandoff`` descriptors.

Buffer layout per rank:  [2 hand-off buffers][world slots][slot_bytes]  then, per buffer, 128 bytes of
``uint32 seq | .. | recv_done (+32)``.  A slot holds one 32-bit packet ``tag << 16 | bf16`` per element (the tag makes the
packet self-validating: no fence and no separate arrival flag).  Consecutive hand-offs alternate between the two buffers
(attention branch: 0, MLP branch: 1): a rank can only overwrite a peer's slot of buffer b after that peer has produced
the next hand-off on the other buffer, i.e. after it has finished reading buffer b.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from . import _cabi as C

FLAG_BLOCK = 128


class PeerComm:
    def __init__(self, group, device: torch.device, slot_bytes: int, gather_slot_bytes: int = 0) -> None:
        import torch.distributed as dist

        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if not 2 <= self.world <= 8:
            raise C.B200Error(f"PeerComm supports 2..8 ranks on one node (got {self.world})")
        self.device = device
        self.slot_bytes = (int(slot_bytes) + 255) & ~255
        self.data_bytes = 2 * self.world * self.slot_bytes
        # two-stage exchange (b200_tp_handoff::reduced): every consumer CTA sums only its slice of the vector over the ranks and
        # republishes it in a local buffer that all CTAs then read.  OFF by default: bit-identical, but the extra dependent round
        # trip costs more than the world x fewer L2 reads save (TP-8 Llama-2-7B 703 vs 851 tok/s, 70B 205 vs 214).
        # B200_TP_TWO_STAGE_MIN=<world> switches it on from that many ranks.
        import os

        self.two_stage = self.world >= int(os.environ.get("B200_TP_TWO_STAGE_MIN", "99"))
        # third buffer with its own counters: all-gather of a row-sharded lm_head (b200_tp_gather); 0 = none
        self.gather_slot_bytes = (int(gather_slot_bytes) + 255) & ~255
        self.gather_off = self.data_bytes + 2 * FLAG_BLOCK + 2 * self.slot_bytes
        total = self.gather_off + self.world * self.gather_slot_bytes + FLAG_BLOCK
        lib = C.load()
        self._local, self._opened, self.base = 0, [], []
        # every stage is followed by a collective agreement, so that a failure on one rank fails all of them the same way
        mine: Optional[bytes] = None
        err = ""
        try:
            with torch.cuda.device(device):
                ptr = ctypes.c_void_p()
                handle = (ctypes.c_ubyte * 64)()
                C.check(lib.b200_p2p_alloc(total, ctypes.byref(ptr), handle), "p2p_alloc")
            self._local, mine = ptr.value, bytes(handle)
        except Exception as e:  # noqa: BLE001
            err = str(e)
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, mine, group=group)
        if any(h is None for h in handles):
            self.close()
            raise C.B200Error(f"peer buffer allocation failed on a rank ({err or 'another rank'})")
        try:
            with torch.cuda.device(device):
                for r in range(self.world):
                    if r == self.rank:
                        self.base.append(self._local)
                        continue
                    p = ctypes.c_void_p()
                    buf = (ctypes.c_ubyte * 64).from_buffer_copy(handles[r])
                    C.check(lib.b200_p2p_open(buf, ctypes.byref(p)), "p2p_open")
                    self.base.append(p.value)
                    self._opened.append(p.value)
        except Exception as e:  # noqa: BLE001
            err = str(e)
        oks: List[Optional[bool]] = [None] * self.world
        dist.all_gather_object(oks, err == "", group=group)
        if not all(oks):
            self.close()
            raise C.B200Error(f"mapping a peer buffer failed on a rank ({err or 'another rank'})")

    # -- addresses ------------------------------------------------------------------------------------------
    def _slot(self, base: int, buf: int, r: int) -> int:
        return base + (buf * self.world + r) * self.slot_bytes

    def _flags(self, base: int, buf: int) -> int:
        return base + self.data_bytes + buf * FLAG_BLOCK

    def handoff(self, send: Optional[int] = None, recv: Optional[int] = None) -> "C.TpHandoff":
        """Descriptor for one GEMV launch: ``send`` = buffer this launch's partial goes to (producer), ``recv`` =
        buffer whose slots replace ``delta`` (consumer)."""
        h = C.TpHandoff()
        h.world, h.rank = self.world, self.rank
        if send is not None:
            for r in range(self.world):
                h.peer_y[r] = self._slot(self.base[r], send, self.rank)
            h.send_seq = self._flags(self._local, send)
        if recv is not None:
            h.slots = self._slot(self._local, recv, 0)
            h.slot_stride = self.slot_bytes
            h.recv_seq = self._flags(self._local, recv)
            h.recv_done = self._flags(self._local, recv) + 32
            if self.two_stage:
                h.reduced = self._local + self.data_bytes + 2 * FLAG_BLOCK + recv * self.slot_bytes
        return h

    def gather_send(self) -> "C.TpHandoff":
        """Producer descriptor of the all-gather: this rank's outputs go to slot `rank` of every rank's gather buffer."""
        h = C.TpHandoff()
        h.world, h.rank = self.world, self.rank
        for r in range(self.world):
            h.peer_y[r] = self.base[r] + self.gather_off + self.rank * self.gather_slot_bytes
        h.send_seq = self._local + self.gather_off + self.world * self.gather_slot_bytes
        return h

    def gather_recv(self) -> "C.TpHandoff":
        """Consumer descriptor of the all-gather (b200_tp_gather)."""
        h = C.TpHandoff()
        h.world, h.rank = self.world, self.rank
        h.slots = self._local + self.gather_off
        h.slot_stride = self.gather_slot_bytes
        h.recv_seq = self._local + self.gather_off + self.world * self.gather_slot_bytes
        h.recv_done = h.recv_seq + 32
        return h

    def close(self) -> None:
        lib = C.load()
        for p in self._opened:
            lib.b200_p2p_close(p)
        self._opened = []
        if self._local:
            lib.b200_p2p_free(self._local)
            self._local = 0
