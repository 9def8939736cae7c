"""Multi-GPU host logic (no reference counterpart: the reference's NRC path is single-device, Device.cpp:2436).

One process per GPU.  Cache queries are independent -> contiguous tile-aligned shards, replicated weights, NO
collective.  Training is data parallel: every rank runs the fused forward/backward kernel on its shard of the
batch with the loss normalised by the GLOBAL batch size, the fp32 weight-gradient vectors are summed with one
all-reduce (NCCL over NVLink on the GPU box; gloo in the CPU tests), then every rank applies the identical
Adam+EMA step, so the replicas stay bit-identical without a weight broadcast.
"""
import torch
import torch.distributed as dist

TILE = 128   # queries per kernel tile


def shard_range(n, rank, world, tile=TILE):
    """Contiguous [lo, hi) of rank's share of n units; boundaries are multiples of `tile` (last shard ragged)."""
    tiles = (n + tile - 1) // tile
    lo = min(n, (tiles * rank // world) * tile)
    hi = min(n, (tiles * (rank + 1) // world) * tile)
    return lo, hi


class _DevPtr:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


class DataParallelTrainer:
    """Drives `net` (nrc_b200.Network, or any object with train_grad / train_apply / n_params) data-parallel.

    mode "nccl":  backward kernel -> torch.distributed all-reduce (NCCL / gloo) -> Adam+EMA kernel.
    mode "fused": ONE kernel per rank and step; the all-reduce happens inside it through NVLink peer memory
                  (nrc_b200_dp_export / _connect / _train_dp; CUDA IPC handles are all-gathered here once)."""

    def __init__(self, net, group=None, mode="nccl"):
        self.net, self.group, self.mode = net, group, mode
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._views = {}
        if mode == "fused" and self.world > 1:
            mine = net.dp_export()
            handles = [None] * self.world
            dist.all_gather_object(handles, mine, group=group)
            net.dp_connect(self.rank, self.world, b"".join(handles))
            dist.barrier(group=group)      # every rank has mapped every buffer before the first push

    def _grad_tensor(self, g):
        if isinstance(g, torch.Tensor):
            return g
        if g not in self._views:   # raw device pointer from the C ABI -> zero-copy torch view
            self._views[g] = torch.as_tensor(_DevPtr(g, self.net.n_params()[0]), device="cuda")
        return self._views[g]

    def train(self, inputs, targets, n_local, stream=None, want_loss=False):
        """One optimizer step on a global batch of world*n_local samples; `inputs`/`targets` are this rank's shard."""
        n_global = n_local * self.world
        if self.mode == "fused" and self.world > 1:
            return self.net.train_dp(inputs, targets, n_local, n_global, stream, want_loss=want_loss)
        g = self._grad_tensor(self.net.train_grad(inputs, targets, n_local, n_global, stream))
        if self.world > 1:
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=self.group)
        loss = self.net.train_apply(stream, want_loss=want_loss)
        if want_loss and self.world > 1:   # each rank holds the sum over its shard of loss terms / (3*n_global)
            t = torch.tensor([loss], dtype=torch.float64, device=g.device)
            dist.all_reduce(t, group=self.group)
            loss = float(t.item())
        return loss
