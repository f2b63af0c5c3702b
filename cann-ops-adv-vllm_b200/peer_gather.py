"""Output all-gather of the KV-head tensor-parallel decode step, fused into the attention launch.

`kvhead_tp.HeadGather` is the NCCL baseline (one all-gather after the kernel, ~15 us of launch + protocol
latency for a 16-64 KiB payload).  Here the combine epilogue of the chain kernel stores every output row
straight into the gathered buffer [B, Nq, D] of every rank (NVLink peer stores) and the launch ends with a
system-scope flag exchange (include/spa_b200.h: spaExecutorSetPeerOutput), so when the launch has
completed on a rank's stream that rank holds the full attention_out: no NCCL call, no extra launch.

torch.distributed is used once, at construction, to exchange the CUDA IPC handles of the buffers
(all_gather_object).  The reference has no distributed code on this path.
"""
from typing import List, Optional

import torch
import torch.distributed as dist

import custom_ops_lib
from custom_ops import FusedDecodePlan
from kvhead_tp import HeadShard

_SIGNAL_BYTES = 256


class PeerGatheredOutput:
    """`buffers` gathered-output images [B, Nq, D] + one signal block in ONE peer-mapped allocation per rank.

    Two buffers, used alternately, make back-to-back launches safe without any further synchronisation:
    a rank cannot finish launch t before every peer has finished launch t, hence buffer t % 2 is not
    overwritten while a peer still reads what launch t-2 left there."""

    def __init__(self, batch: int, num_heads: int, head_dim: int, dtype: torch.dtype, shard: HeadShard,
                 device: torch.device, group: Optional[dist.ProcessGroup] = None, buffers: int = 2):
        self.shard, self.device, self.dtype = shard, torch.device(device), dtype
        self.shape = (batch, num_heads, head_dim)
        elem = torch.empty(0, dtype=dtype).element_size()
        self.buf_bytes = (batch * num_heads * head_dim * elem + 255) // 256 * 256
        self.buffers = buffers
        self.total = self.buf_bytes * buffers + _SIGNAL_BYTES
        with torch.cuda.device(self.device):
            self._local_ptr, handle = custom_ops_lib.peer_buffer_create(self.total)
            handles: List[Optional[bytes]] = [None] * shard.world
            if shard.world > 1:
                dist.all_gather_object(handles, handle, group=group)
            self._ptrs = []
            for r in range(shard.world):
                self._ptrs.append(self._local_ptr if r == shard.rank else custom_ops_lib.peer_buffer_open(handles[r]))
            if shard.world > 1:
                dist.barrier(group=group)          # every rank has mapped every buffer before anyone launches
        self._mem = custom_ops_lib.peer_buffer_tensor(self._local_ptr, self.total, self.device.index or 0)
        self.signals = self._mem[self.buf_bytes * buffers:].view(torch.int32)

    def gathered(self, i: int) -> torch.Tensor:
        """Local image of gathered buffer i: [B, Nq, D]."""
        B, N, D = self.shape
        raw = self._mem[i * self.buf_bytes:(i + 1) * self.buf_bytes].view(self.dtype)
        return raw[:B * N * D].view(B, N, D)

    def shard_view(self, i: int) -> torch.Tensor:
        """This rank's rows of gathered buffer i: [B, Nq/G, D] (strided) -- the kernel's attention_out."""
        return self.gathered(i)[:, self.shard.q_lo:self.shard.q_hi, :]

    def attach(self, plan: FusedDecodePlan, i: int) -> None:
        """Bind a plan whose attention_out is shard_view(i) to the peers' images of buffer i."""
        sig_off = self.buf_bytes * self.buffers
        plan._plan.set_peer_output(self.shard.rank, self.shard.world, [p + i * self.buf_bytes for p in self._ptrs],
                                   [p + sig_off for p in self._ptrs])

    def timed_out(self) -> bool:
        """True if a flag exchange gave up (a peer did not launch within 2 s).  Synchronises."""
        return bool(self.signals[9].item() != 0)

    def exchange_ns(self) -> int:
        """Total ns this rank's launches have spent in the end-of-launch flag exchange (fence, flag stores,
        waiting for the slowest peer).  Synchronises."""
        return int(self.signals[10:12].view(torch.int64).item())

    def exchange_send_ns(self) -> int:
        """The part of exchange_ns() spent before the wait (system fence + flag stores)."""
        return int(self.signals[12:14].view(torch.int64).item())

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        for r, p in enumerate(self._ptrs):
            if r != self.shard.rank and p:
                custom_ops_lib.peer_buffer_close(p)
        self._ptrs = []
        if self._local_ptr:
            self._mem = self.signals = None
            custom_ops_lib.peer_buffer_destroy(self._local_ptr)
            self._local_ptr = 0


class TPFusedDecode:
    """KV-head tensor-parallel fused decode step of one rank: SparsePagedFusionAttention over this rank's KV
    heads whose launch also delivers the output rows to every peer.  `launch()` returns the gathered
    [B, Nq, D] tensor (valid once the launch has completed on the current stream)."""

    def __init__(self, shard: HeadShard, num_heads_total: int, query, key, value, blocktable, l1_cent, block_ids,
                 total_seq_len, block_position, page_position_length, max_page_position_length, scale_value,
                 block_size=128, *, k=64, antiquant_scale=None, antiquant_offset=None, group=None):
        B, nq, D = query.shape[0], query.shape[1], query.shape[-1]
        assert nq == shard.num_heads, "query must hold this rank's heads only"
        self.out = PeerGatheredOutput(B, num_heads_total, D, query.dtype, shard, query.device, group)
        self.plans = []
        for i in range(2):
            plan = FusedDecodePlan(query, key, value, blocktable, l1_cent, block_ids, total_seq_len, block_position,
                                   page_position_length, max_page_position_length, self.out.shard_view(i), nq,
                                   scale_value, shard.num_kv_heads, block_size, k=k, antiquant_scale=antiquant_scale,
                                   antiquant_offset=antiquant_offset)
            self.out.attach(plan, i)
            self.plans.append(plan)
        self.step = 0
        self.graphs = None

    @property
    def num_launches(self) -> int:
        return self.plans[0].num_launches

    def launch(self) -> torch.Tensor:
        i = self.step & 1
        self.step += 1
        self.plans[i].launch()
        return self.out.gathered(i)

    def capture(self) -> None:
        """Capture one CUDA graph per gathered buffer (the launch count of the flag exchange lives in the signal
        block on the device, so a captured launch replays correctly); `replay()` then alternates between them
        exactly as `launch()` alternates between the plans.  Collective in the sense that EVERY rank must call it at
        the same point: both plans are launched once eagerly first (workspace clear outside the capture)."""
        if self.graphs is not None:
            return
        dev = self.out.device
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            self.launch()
            self.launch()
        cur.wait_stream(side)
        graphs = []
        for i in range(2):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.plans[i].launch()
            graphs.append(g)
        self.graphs = graphs

    def replay(self) -> torch.Tensor:
        """`launch()` from the captured graphs."""
        if self.graphs is None:
            self.capture()
        i = self.step & 1
        self.step += 1
        self.graphs[i].replay()
        return self.out.gathered(i)

    def close(self) -> None:
        self.graphs = None
        self.plans = []
        self.out.close()
