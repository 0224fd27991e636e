"""World-size-2 `gloo` tests of the data-parallel host logic: batch sharding by rank, the
positional sigma slice, the flat gradient bucket and the step's single all-reduce.  CPU only; the loss
inside each rank is the oracle's FD-SSM restatement (checker role: the thing under test is ddp.py)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make(seed=3, D=12, H=16):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(D, H), torch.nn.Softplus(), torch.nn.Linear(H, 1)).double()   # fp64: the FD loss cancels


def _fd_loss(net, x, u, eps=0.1):
    from oracle import fd_losses
    v = fd_losses.sphere_direction(u, eps)
    return fd_losses.fd_ssm(lambda z: net(z).view(-1), x, v, x.shape[1] / eps ** 2)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import fdsm_b200  # noqa: F401
        from fdsm_b200 import ddp
        g = torch.Generator().manual_seed(11)
        X, U = torch.randn(8, 12, generator=g).double(), torch.randn(8, 12, generator=g).double()
        net = _make(seed=3 + rank)                 # ranks start different: the broadcast must fix that
        flat = ddp.FlatParams(net)
        ddp.broadcast_params_(flat.flat)
        lo, hi = ddp.shard_range(8)
        flat.zero_grad()
        _fd_loss(net, X[lo:hi], U[lo:hi]).backward()
        ddp.allreduce_mean_(flat.grad)
        sig = ddp.shard_sigmas(torch.linspace(0.1, 0.8, 8))
        # numpy arrays travel by value: a torch tensor in an mp queue is a shared-memory handle that dies with this process
        q.put((rank, flat.flat.detach().numpy().copy(), flat.grad.detach().numpy().copy(), (lo, hi), sig.numpy().copy(),
               ddp.is_distributed()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_gloo_step_matches_single_process():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=100) for _ in range(2)), key=lambda t: t[0])
    got = [tuple(torch.from_numpy(a) if hasattr(a, "dtype") else a for a in t) for t in got]
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    # single-process reference: full batch, rank 0's initial parameters
    from fdsm_b200 import ddp
    g = torch.Generator().manual_seed(11)
    X, U = torch.randn(8, 12, generator=g).double(), torch.randn(8, 12, generator=g).double()
    net = _make(seed=3)
    flat = ddp.FlatParams(net)
    flat.zero_grad()
    _fd_loss(net, X, U).backward()
    (r0, p0, g0, s0, sig0, d0), (r1, p1, g1, s1, sig1, d1) = got
    assert d0 and d1
    assert s0 == (0, 4) and s1 == (4, 8)
    assert torch.equal(p0, p1) and torch.equal(p0, flat.flat)          # broadcast from rank 0
    assert torch.equal(g0, g1)                                         # replicas see identical averaged grads
    torch.testing.assert_close(g0, flat.grad, rtol=1e-9, atol=1e-11)    # mean of shard means == full-batch mean
    assert torch.equal(sig0, torch.linspace(0.1, 0.8, 8)[:4]) and torch.equal(sig1, torch.linspace(0.1, 0.8, 8)[4:])


def test_shard_range_rejects_ragged():
    from fdsm_b200 import ddp
    with pytest.raises(ValueError):
        ddp.shard_range(10, rank=0, world=4)
    assert ddp.shard_range(12, rank=2, world=4) == (6, 9)


def test_flat_params_alignment_and_views():
    from fdsm_b200 import ddp
    net = _make(D=11, H=30)
    before = [p.detach().clone() for p in net.parameters()]
    flat = ddp.FlatParams(net)
    assert all(o % 4 == 0 for o in flat.offsets)
    for p, b in zip(net.parameters(), before):
        assert torch.equal(p, b)
    flat.flat.mul_(2.0)
    for p, b in zip(net.parameters(), before):
        assert torch.equal(p, 2 * b)          # parameters are views into the bucket


def test_numa_binding_is_optional_and_never_raises(fdsm, monkeypatch):
    """bind_host_thread_to_gpu_node: no NVML (this container) or FDSM_NUMA_BIND=0 -> returns None, affinity untouched."""
    import os
    from fdsm_b200.train import bind_host_thread_to_gpu_node
    before = os.sched_getaffinity(0)
    assert bind_host_thread_to_gpu_node(0) is None or os.sched_getaffinity(0) <= before
    os.sched_setaffinity(0, before)
    monkeypatch.setenv("FDSM_NUMA_BIND", "0")
    assert bind_host_thread_to_gpu_node(0) is None
    assert os.sched_getaffinity(0) == before


def test_trainer_rejects_unknown_objectives(fdsm):
    import pytest
    from fdsm_b200.train import FDSSMTrainer
    with pytest.raises(ValueError):
        FDSSMTrainer(object(), loss="nope")
    with pytest.raises(ValueError):
        FDSSMTrainer(object(), loss="mdsm_fd")      # needs the positional noise schedule


def test_numa_binding_decodes_the_nvml_mask(fdsm, monkeypatch):
    """With a stub NVML reporting CPUs {0, 2} as GPU-local, the calling thread ends up on exactly those of them it may use."""
    import os
    import sys
    import types
    from fdsm_b200.train import bind_host_thread_to_gpu_node
    before = os.sched_getaffinity(0)
    if not {0, 2} <= before or before == {0, 2}:
        import pytest
        pytest.skip("this process may not run on CPUs 0 and 2 (or on nothing else)")
    stub = types.ModuleType("pynvml")
    stub.nvmlInit = lambda: None
    stub.nvmlDeviceGetHandleByUUID = lambda u: (_ for _ in ()).throw(RuntimeError("no such uuid"))
    stub.nvmlDeviceGetHandleByIndex = lambda i: ("handle", i)
    stub.nvmlDeviceGetCpuAffinity = lambda h, n: [0b0101] + [0] * (n - 1)
    monkeypatch.setitem(sys.modules, "pynvml", stub)
    monkeypatch.delenv("FDSM_NUMA_BIND", raising=False)
    try:
        assert bind_host_thread_to_gpu_node(0) == {0, 2}
        assert os.sched_getaffinity(0) == {0, 2}
        # GPU-local CPUs outside the cpuset: nothing changes
        stub.nvmlDeviceGetCpuAffinity = lambda h, n: [0] * (n - 1) + [1 << 63]
        os.sched_setaffinity(0, before)
        assert bind_host_thread_to_gpu_node(0) is None and os.sched_getaffinity(0) == before
    finally:
        os.sched_setaffinity(0, before)


def test_single_process_owns_the_whole_batch():
    """No process group: shard_range / shard_sigmas give the full range (FDSSMTrainer at N = 1 goes through them)."""
    from fdsm_b200 import ddp
    assert ddp.shard_range(12) == (0, 12)
    s = torch.linspace(0.1, 0.8, 8)
    assert torch.equal(ddp.shard_sigmas(s), s)


def test_trainer_hands_the_dsm_step_its_sigma_slice(fdsm):
    """FDSSMTrainer(loss='mdsm_fd') host logic with a stand-in for the fused network: the step receives the batch, this
    process's slice of the positional sigma schedule and the DeepEBM defaults."""
    from fdsm_b200.train import FDSSMTrainer

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.flat = torch.nn.Parameter(torch.zeros(8))
            self.calls = []

        def flat_params(self):
            return self

        def configure_optimizer(self, **kw):
            self.opt = kw

        def fd_ssm_step(self, x, eps, allreduce=None, distributed=False):
            self.calls.append(("ssm", x.shape, eps, distributed))
            return torch.zeros(())

        def mdsm_fd_step(self, x, sigmas, sigma02, eps, loss_coe, allreduce=None, distributed=False):
            self.calls.append(("dsm", x.shape, sigmas.clone(), sigma02, eps, loss_coe, distributed))
            return torch.zeros(())

    sig = torch.linspace(0.05, 1.2, 6).view(-1, 1)
    net = Net()
    tr = FDSSMTrainer(net, eps=0.05, loss="mdsm_fd", sigmas=sig, sigma02=0.01, loss_coe=1e-4)
    assert tr.host_cpus is None                       # CPU parameters: no GPU to be local to
    tr.step(torch.rand(6, 1, 4, 4))
    kind, shape, s, sigma02, eps, coe, distributed = net.calls[0]
    assert kind == "dsm" and shape == (6, 1, 4, 4) and torch.equal(s, sig.view(-1)) and not distributed
    assert (sigma02, eps, coe) == (0.01, 0.05, 1e-4)
    with pytest.raises(RuntimeError):                 # schedule and batch disagree
        tr.step(torch.rand(5, 1, 4, 4))
    tr2 = FDSSMTrainer(net, eps=0.05)
    tr2.step(torch.rand(6, 16))
    assert net.calls[-1][0] == "ssm"
