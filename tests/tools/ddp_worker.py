"""Worker of tests/test_ddp_gpu.py (one process per GPU, launched by torch.distributed.run): the peer-memory gradient
exchange (csrc/fd_peer.cu) against the library collective, under CUDA-graph replay, with skewed ranks, and its timeout path.
Prints `DDP_WORKER_OK` from rank 0 when every check passed on every rank."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import fdsm_b200  # noqa: E402,F401
from fdsm_b200 import ddp  # noqa: E402


def rank_order_mean(x, world):
    parts = [torch.empty_like(x) for _ in range(world)]
    dist.all_gather(parts, x)
    s = parts[0].clone()
    for p in parts[1:]:
        s += p
    return s * (1.0 / world)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    replays = int(os.environ.get("FDSM_DDP_REPLAYS", "10000"))
    fails = []

    # (1) bit-for-bit against the sum in rank order (what the kernel promises) and against ncclAllReduce (bit-for-bit at
    #     world 2, where the sum has one order; 1 ulp-level otherwise), odd sizes included
    for n in (240005, 7, 1, 4096):
        ar = ddp.PeerAllReduce(n, dev)
        sumsq = torch.zeros(32, device=dev)
        step = torch.zeros((), dtype=torch.int64, device=dev)
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        for it in range(4):
            x = torch.randn(n, device=dev, generator=g)
            want = rank_order_mean(x, world)
            nccl = x * (1.0 / world)
            dist.all_reduce(nccl)
            got = ar(x.clone(), sumsq=sumsq, step=step)
            torch.cuda.synchronize()
            if not torch.equal(got, want):
                fails.append(f"n={n} it={it}: differs from the rank-order sum")
            if world == 2 and not torch.equal(got, nccl):
                fails.append(f"n={n} it={it}: differs from ncclAllReduce")
            if world > 2 and float((got - nccl).abs().max()) > 1e-6:
                fails.append(f"n={n} it={it}: far from ncclAllReduce")
            ss = float((got.double() ** 2).sum())
            if abs(float(sumsq.sum()) - ss) > 1e-5 * max(ss, 1e-30) or int(step) != it + 1:
                fails.append(f"n={n} it={it}: squared norm / step counter")
            # the partials set the clip coefficient: replicas stay bit-identical only if these are (both protocols;
            # FDSM_PEER_TWO_SHOT=1 runs the two-shot kernel at world 2)
            sq = [torch.empty_like(sumsq) for _ in range(world)]
            dist.all_gather(sq, sumsq)
            if not all(torch.equal(sq[0], q) for q in sq[1:]):
                fails.append(f"n={n} it={it}: squared-norm partials differ across ranks")

    # (2) stress: `replays` back-to-back graph replays, no host sync in between, the two ranks alternately delayed so that
    #     either may run a full step ahead (the parity double buffer).  Data: small integers derived in-graph from the device
    #     step counter, so every sum is exact and the expected result is computed in-graph as well; mismatches accumulate in
    #     a device counter.
    n = 240005
    ar = ddp.PeerAllReduce(n, dev)
    sumsq = torch.zeros(32, device=dev)
    step = torch.zeros((), dtype=torch.int64, device=dev)
    idx = torch.arange(n, device=dev, dtype=torch.float32)
    bad = torch.zeros((), dtype=torch.int64, device=dev)
    out = torch.empty(n, device=dev)

    def body(slow_rank):
        k = step.to(torch.float32)
        x = torch.remainder(idx * 7.0 + k * 13.0 + rank * 101.0, 1024.0)
        want = sum(torch.remainder(idx * 7.0 + k * 13.0 + r * 101.0, 1024.0) for r in range(world)) * (1.0 / world)
        if rank == slow_rank:
            torch.cuda._sleep(40000)
        ar(x, sumsq=sumsq, step=step, out=out)
        bad.add_((out != want).sum())

    graphs = []
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for slow in range(world):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                body(slow)
            graphs.append(gr)
    torch.cuda.current_stream().wait_stream(s)
    dist.barrier()
    for it in range(replays):
        graphs[(it // 3) % world].replay()
    torch.cuda.synchronize()
    if int(bad) != 0 or int(step) != replays or ar.timed_out():
        fails.append(f"stress: {int(bad)} mismatching elements over {replays} replays, step counter {int(step)}, timed_out {ar.timed_out()}")

    # (3) a rank that never arrives: the launch gives up after FDSM_PEER_TIMEOUT_MS, poisons the squared norm, leaves the
    #     step counter alone, and the status word stays set
    if os.environ.get("FDSM_PEER_TIMEOUT_MS"):
        ar = ddp.PeerAllReduce(4096, dev)
        sumsq = torch.zeros(32, device=dev)
        step = torch.zeros((), dtype=torch.int64, device=dev)
        dist.barrier()
        if rank == 0:
            ar(torch.ones(4096, device=dev), sumsq=sumsq, step=step)
            torch.cuda.synchronize()
            if not ar.timed_out() or not bool(torch.isnan(sumsq).all()) or int(step) != 0:
                fails.append(f"timeout path: timed_out {ar.timed_out()} sumsq {sumsq[:2].tolist()} step {int(step)}")
        dist.barrier()

    flag = torch.tensor([len(fails)], device=dev)
    dist.all_reduce(flag)
    for f in fails:
        print(f"[rank {rank}] {f}", flush=True)
    if rank == 0 and int(flag) == 0:
        print("DDP_WORKER_OK", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
