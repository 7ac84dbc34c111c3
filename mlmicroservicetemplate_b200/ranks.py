"""One-process-per-GPU plumbing for the replica-parallel bench: rank discovery, a barrier, and the
max-over-ranks of a device-measured time.  The predict path shards by request — whole batches go to
independent replicas — so there is NO data-path collective; torch.distributed carries only these two
control operations (NCCL on GPUs, gloo in the CPU tests)."""
from __future__ import annotations

import os
from dataclasses import dataclass


@dataclass
class Ranks:
    rank: int = 0
    local_rank: int = 0
    world: int = 1
    backend: str = ""
    _cpu_group: object = None

    @classmethod
    def from_env(cls, backend: str = "nccl", device=None) -> "Ranks":
        world = int(os.environ.get("WORLD_SIZE", "1"))
        r = cls(int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), world, backend)
        if world > 1:
            import torch.distributed as dist
            kw = {"device_id": device} if (device is not None and backend == "nccl") else {}
            dist.init_process_group(backend, **kw)
            # a host-only group: waiting on it launches nothing on the GPUs (an NCCL barrier parks a spinning
            # kernel on every rank's device, which would sit on SMs another process is measuring)
            r._cpu_group = dist.new_group(backend="gloo") if backend == "nccl" else None
        return r

    def host_barrier(self) -> None:
        """Barrier that touches no GPU (bench.py: the other ranks idle here while rank 0 drives all GPUs)."""
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self._cpu_group) if self._cpu_group is not None else dist.barrier()

    def _tensor(self, v: float):
        import torch
        dev = "cuda" if self.backend == "nccl" else "cpu"
        return torch.tensor([v], dtype=torch.float64, device=dev)

    def barrier(self) -> None:
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        if self.backend == "nccl":
            import torch
            torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return float(v)
        import torch.distributed as dist
        t = self._tensor(v)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def same_on_all_ranks(self, digest: bytes) -> bool:
        """True when every rank holds the same digest (the same batch through the replica on any GPU must
        give bit-identical logits).  48 bits of it travel as an exact float64 through a MAX and a MIN reduction."""
        v = float(int.from_bytes(digest[:6], "big"))
        if self.world == 1:
            return True
        import torch.distributed as dist
        hi, lo = self._tensor(v), self._tensor(v)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        return float(hi.item()) == float(lo.item())

    def job_throughput(self, units_per_rank: int, elapsed_ms: float) -> float:
        """Whole-job units/s: every rank processed `units_per_rank`; the job took the slowest rank's time."""
        return self.world * units_per_rank / (self.max_over_ranks(elapsed_ms) * 1e-3)

    def shard(self, n_items: int):
        """Contiguous share of n_items independent work units for this rank (no exchange afterwards)."""
        base, extra = divmod(n_items, self.world)
        lo = self.rank * base + min(self.rank, extra)
        return range(lo, lo + base + (1 if self.rank < extra else 0))

    def close(self) -> None:
        if self.world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()
