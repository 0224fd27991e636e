"""torchrun --nproc-per-node N scripts/peer_allreduce_check.py
Checks fdsm_peer_allreduce (NVLink peer-memory all-reduce fused with the clip's squared norm) against NCCL on the same
vectors: eager and CUDA-graph replays, odd n (scalar tail), many back-to-back epochs, bit-identity across ranks; then
times both.  Prints one line per check from rank 0; exits non-zero on mismatch."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200  # noqa: E402,F401
from fdsm_b200 import ddp  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for n in (238003, 7, 4096, 1 << 20):
        ar = ddp.PeerAllReduce(n, dev)
        sumsq = torch.zeros(32, device=dev)
        step = torch.zeros((), dtype=torch.int64, device=dev)
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        for it in range(6):
            x = torch.randn(n, device=dev, generator=g)
            want = x.clone()
            dist.all_reduce(want)
            want /= world
            got = ar(x.clone(), sumsq=sumsq, step=step)
            torch.cuda.synchronize()
            err = float((got - want).abs().max())
            ss_err = abs(float(sumsq.sum()) - float((got.double() ** 2).sum())) / max(float((got.double() ** 2).sum()), 1e-30)
            # bit-identical across ranks
            ref = got.clone()
            dist.broadcast(ref, src=0)
            same = bool(torch.equal(ref, got))
            if err > 1e-6 or ss_err > 1e-5 or not same or int(step) != it + 1:
                ok = False
                print(f"[rank {rank}] n={n} it={it}: max err {err:.3e} sumsq rel {ss_err:.2e} identical={same} step={int(step)}")
        # graph replay: the epoch lives in device memory
        buf = torch.randn(n, device=dev, generator=g)
        src = buf.clone()
        gr = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s), torch.cuda.graph(gr):
            buf.copy_(src)
            ar(buf, sumsq=sumsq, step=step)
        torch.cuda.current_stream().wait_stream(s)
        want = src.clone()
        dist.all_reduce(want)
        want /= world
        for it in range(5):
            gr.replay()
            torch.cuda.synchronize()
            if float((buf - want).abs().max()) > 1e-6:
                ok = False
                print(f"[rank {rank}] n={n} graph replay {it}: max err {float((buf - want).abs().max()):.3e}")
        if ar.timed_out():
            ok = False
            print(f"[rank {rank}] n={n}: a launch timed out waiting for a peer")
        # timing: 200 back-to-back launches (graph of 20) vs NCCL
        if n == 238003:
            g20 = torch.cuda.CUDAGraph()
            with torch.cuda.stream(s), torch.cuda.graph(g20):
                for _ in range(20):
                    ar(buf, sumsq=sumsq, step=step)
            torch.cuda.current_stream().wait_stream(s)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for _ in range(3):
                g20.replay()
            dist.barrier(); torch.cuda.synchronize()
            e0.record()
            for _ in range(10):
                g20.replay()
            e1.record(); torch.cuda.synchronize()
            t_peer = e0.elapsed_time(e1) / 200 * 1e3
            for _ in range(20):
                dist.all_reduce(buf)
            dist.barrier(); torch.cuda.synchronize()
            e0.record()
            for _ in range(200):
                buf.mul_(1.0 / world)
                dist.all_reduce(buf)
            e1.record(); torch.cuda.synchronize()
            t_nccl = e0.elapsed_time(e1) / 200 * 1e3
            ph = ar.phase_stamps_us()
            if rank == 0:
                print("block-0 phases of the last launch (us): push %.1f  wait+reduce %.1f  tail %.1f" % tuple(ph))
                print(f"n={n} ({4 * n / 1e6:.2f} MB) world={world}: peer kernel {t_peer:.1f} us/launch (incl. sumsq + step), "
                      f"NCCL mul_+all_reduce {t_nccl:.1f} us/iter (eager launches)")
        if rank == 0:
            print(f"n={n}: {'ok' if ok else 'MISMATCH'}")
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.destroy_process_group()
    sys.exit(1 if int(flag) else 0)


if __name__ == "__main__":
    main()
