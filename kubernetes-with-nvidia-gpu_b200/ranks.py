"""One process per GPU: the only cross-rank plumbing the workload needs.

The vectors shard with no communication (BASELINE.json: "NCCL is not used, because the vectors
shard with no communication"), so there is no data-path collective.  What the ranks do share is
bookkeeping: a start barrier, the max over ranks of the device-timed durations, the AND of the
verify bits and the sum of the shard checksums.  torch.distributed carries those few scalars
(NCCL on GPUs, gloo in the CPU tests); rendezvous comes from the torchrun environment.
"""
import os


class Ranks:
    def __init__(self, backend=None, device_sync=None):
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.device_sync = device_sync
        self.dist = None
        self.torch = None
        self.device = "cpu"
        self.host_group = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            self.torch, self.dist = torch, dist
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            if backend == "nccl":
                torch.cuda.set_device(self.local_rank)
                self.device = f"cuda:{self.local_rank}"
                dist.init_process_group("nccl", device_id=torch.device(self.device))
                # a second, CPU-only group for waits that must not put a kernel on the GPUs (an NCCL barrier
                # parks a spinning kernel on every waiting rank's device).  One node only, so the loopback
                # interface will do when the container's hostname does not resolve; if gloo cannot come up
                # at all, host_barrier() falls back to the NCCL barrier.
                os.environ.setdefault("GLOO_SOCKET_IFNAME", "lo")
                try:
                    self.host_group = dist.new_group(backend="gloo")
                except Exception:
                    self.host_group = None
            else:
                dist.init_process_group(backend)

    # -- the four reductions ---------------------------------------------------------------
    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        if self.device_sync is not None:
            self.device_sync()

    def host_barrier(self):
        """Wait for every rank on the host only: no GPU work is enqueued anywhere while ranks wait.
        Used while rank 0 runs the product's launcher binaries on all the GPUs."""
        if self.dist is not None:
            self.dist.barrier(group=self.host_group) if self.host_group is not None else self.dist.barrier()

    def _reduce(self, values, op, dtype):
        if self.dist is None:
            return list(values)
        t = self.torch.tensor(list(values), dtype=dtype, device=self.device)
        self.dist.all_reduce(t, op=op)
        return t.tolist()

    def max(self, x):
        """Max over ranks of a device-timed duration (ms)."""
        if self.dist is None:
            return float(x)
        return float(self._reduce([x], self.dist.ReduceOp.MAX, self.torch.float64)[0])

    def all_ok(self, ok):
        if self.dist is None:
            return bool(ok)
        return bool(self._reduce([1 if ok else 0], self.dist.ReduceOp.MIN, self.torch.int64)[0])

    def sum_checksums(self, pair):
        """Sum of the shards' (plain, weighted) checksums mod 2^64 = the whole vector's checksum.
        Sent as 16-bit limbs in int64 so the sum over up to 2^47 ranks cannot overflow."""
        if self.dist is None:
            return int(pair[0]), int(pair[1])
        limbs = []
        for v in pair:
            limbs += [(int(v) >> (16 * k)) & 0xFFFF for k in range(4)]
        out = self._reduce(limbs, self.dist.ReduceOp.SUM, self.torch.int64)
        res = []
        for h in range(2):
            v = sum(int(out[4 * h + k]) << (16 * k) for k in range(4))
            res.append(v & ((1 << 64) - 1))
        return res[0], res[1]

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()
            self.dist = None
