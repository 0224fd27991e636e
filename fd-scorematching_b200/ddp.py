"""Data-parallel plumbing: one process per GPU, batch sharded by rank, exactly one all-reduce of
the flat fp32 gradient bucket per step.  Replaces the reference's single-process
``nn.DataParallel`` (DeepEBM/train_ssm.py:33-34, ncsn/runners/anneal_runner.py:116,
ncsnv2/runners/ncsn_esm_runner.py:67), which re-broadcasts parameters and gathers outputs to GPU 0
every step.  Works with the NCCL backend on GPUs and with gloo on CPU (host-logic tests).
"""
import torch
import torch.distributed as dist


class FlatParams:
    """Re-homes a module's parameters (and their gradients) into two flat fp32 buffers so that the
    gradient all-reduce, the global-norm clip and the optimizer each touch ONE contiguous tensor."""

    def __init__(self, module):
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise ValueError("module has no trainable parameters")
        dev, dt = params[0].device, params[0].dtype
        # 16-byte align every tensor inside the bucket (128-bit loads in the kernels)
        offs, n = [], 0
        for p in params:
            offs.append(n)
            n += (p.numel() + 3) // 4 * 4
        self.flat = torch.zeros(n, device=dev, dtype=dt)
        self.grad = torch.zeros(n, device=dev, dtype=dt)
        self.params, self.offsets = params, offs
        for p, o in zip(params, offs):
            self.flat[o:o + p.numel()].copy_(p.data.reshape(-1))
            p.data = self.flat[o:o + p.numel()].view_as(p)
            p.grad = self.grad[o:o + p.numel()].view_as(p)
        self.numel = n

    def zero_grad(self):
        self.grad.zero_()
        for p, o in zip(self.params, self.offsets):  # autograd may have re-pointed .grad
            if p.grad is None or p.grad.data_ptr() != self.grad.data_ptr() + 4 * o:
                p.grad = self.grad[o:o + p.numel()].view_as(p)


def is_distributed():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def shard_range(global_batch, rank=None, world=None):
    """Rank r owns samples [r*b, (r+1)*b) of the global batch; all rows of a sample stay on one rank."""
    initialised = dist.is_available() and dist.is_initialized()     # a single process owns the whole batch
    rank = (dist.get_rank() if initialised else 0) if rank is None else rank
    world = (dist.get_world_size() if initialised else 1) if world is None else world
    if global_batch % world:
        raise ValueError(f"global batch {global_batch} not divisible by world size {world}")
    b = global_batch // world
    return rank * b, (rank + 1) * b


def shard_sigmas(sigmas, rank=None, world=None):
    """DeepEBM's sigma schedule is positional within the GLOBAL batch (train_dsm.py:50-57): a shard takes
    its slice, it does not re-linspace (SURVEY.md §7 hard part 7)."""
    lo, hi = shard_range(sigmas.shape[0], rank, world)
    return sigmas[lo:hi]


def allreduce_mean_(flat_grad):
    """The step's single collective: sum over ranks, pre-scaled by 1/world (loss is a per-rank mean)."""
    if not is_distributed():
        return None
    flat_grad.mul_(1.0 / dist.get_world_size())
    return dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, async_op=False)


def broadcast_params_(flat):
    """Once, at step 0.  Replicas then stay bit-identical because they apply identical averaged grads."""
    if is_distributed():
        dist.broadcast(flat, src=0)


class PeerAllReduce:
    """The gradient exchange as ONE kernel over NVLink peer memory (`fdsm_peer_allreduce`, csrc/fd_peer.cu): every rank
    pushes its flat gradient into the peers' receive slots, then sums the slots in rank order -- fused with the
    squared-norm partials of the clip and the optimizer's step counter.  Graph-capturable, so a distributed training step
    is a single CUDA-graph replay.  The receive buffers are torch symmetric memory (peer-mapped over NVLink)."""

    fused_sumsq = True

    def __init__(self, n, device, group=None):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        self._C, self._lib = C, _lib.load()
        group = dist.group.WORLD if group is None else group
        self.world, self.rank, self.n = dist.get_world_size(group), dist.get_rank(group), int(n)
        nbytes = self._lib.fdsm_peer_buffer_bytes(self.n, self.world)
        if nbytes < 0:
            raise ValueError(f"world size {self.world} not supported by fdsm_peer_allreduce")
        self.buf = symm.empty(nbytes // 4, dtype=torch.float32, device=device)
        self.hdl = symm.rendezvous(self.buf, group)
        self.buf.zero_()
        self.scratch = torch.zeros(256, dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        dist.barrier(group)                       # nobody pushes before every receive buffer is zeroed
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self._ptrs = (C.c_void_p * self.world)(*ptrs)

    def __call__(self, grad, sumsq=None, step=None, out=None):
        C = self._C
        out = grad if out is None else out
        if grad.numel() != self.n or grad.dtype != torch.float32 or not grad.is_contiguous():
            raise ValueError("PeerAllReduce: gradient must be the contiguous fp32 vector it was sized for")
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        rc = self._lib.fdsm_peer_allreduce(p(grad), p(out), self.n, self._ptrs, self.rank, self.world, 1.0 / self.world,
                                           p(sumsq), p(step), p(self.scratch),
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc:
            from ._lib import check
            check(rc, "fdsm_peer_allreduce")
        return out

    def phase_stamps_us(self):
        """Debug: block 0's phase durations of the LAST launch (us): push, wait + reduce, tail."""
        t = self.scratch[192:200].view(torch.int64).cpu().tolist()
        return [(t[i + 1] - t[i]) / 1e3 for i in range(3)]

    def timed_out(self):
        """True if any launch gave up waiting for a peer (host sync)."""
        return bool(int(self.scratch[2]))


def make_allreduce(n, device):
    """The exchange step for a flat fp32 gradient of n elements: the peer-memory kernel when the ranks share an NVLink
    domain (FDSM_ALLREDUCE=nccl forces the NCCL library collective), else NCCL.  The choice is collective: the ranks
    all-reduce a success flag after trying to set the peer path up, so one rank without P2P / symmetric memory moves
    EVERY rank to NCCL (mixed backends would deadlock in the first step)."""
    import os
    import sys
    if not is_distributed():
        return allreduce_mean_
    peer, want = None, os.environ.get("FDSM_ALLREDUCE", "peer") != "nccl" and torch.device(device).type == "cuda"
    if want:
        try:
            peer = PeerAllReduce(n, device)
        except Exception as e:   # no P2P / symmetric memory on this box: library collective
            print(f"[fdsm] peer all-reduce unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
    ok = torch.tensor([1 if (peer is not None or not want) else 0], device=device, dtype=torch.int32)
    if want:                     # FDSM_ALLREDUCE must be set identically on every rank (it decides whether this collective runs)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if peer is not None and int(ok) == 1:
        return peer
    return allreduce_mean_
