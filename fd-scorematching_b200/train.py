"""Training-step driver for the FD-SSM / FD-DSM hot loop (DeepEBM/train_ssm.py:60-70):
loss -> zero_grad -> backward -> clip_grad_norm_ -> Adam step, data-parallel over ranks.

This is the call a user makes (`bench.py`'s e2e leg goes through `step_from_host`)."""
import os
import sys

import torch

from . import ddp, losses, ops


def bind_host_thread_to_gpu_node(device_index):
    """Pin the calling host thread to the CPUs NVML reports as local to the GPU (its NUMA node), so that the pinned
    staging buffers of `step_from_host` -- allocated by this thread -- and the launch path sit on the socket the GPU's
    PCIe root hangs off: a remote-socket staging buffer costs H2D bandwidth.  Returns the CPU set applied, or None when
    nothing was changed (NVML missing, FDSM_NUMA_BIND=0, or the local CPUs are outside this process's cpuset).  Never
    raises: binding is an optimisation of the host side, not a requirement."""
    if os.environ.get("FDSM_NUMA_BIND", "1") == "0" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            handle = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
        except Exception:
            handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (ncpu + 63) // 64)
        local = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus = local & allowed
        if not cpus or cpus == allowed:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception as e:   # noqa: BLE001 -- see docstring
        if os.environ.get("FDSM_DEBUG"):
            print(f"fdsm_b200: host thread not bound to the GPU's NUMA node ({type(e).__name__}: {e})", file=sys.stderr)
        return None


def read_loss_to_host(loss, loss_host):
    """Device scalar -> `loss_host`, ordered on the current stream.  A pinned fp32 destination is written by a kernel
    (ops.publish_scalars: no copy-engine operation between two graph replays; FDSM_E2E_LOSS_MEMCPY=1 keeps the memcpy);
    anything else takes torch's non-blocking copy."""
    if (loss.is_cuda and loss.dtype == torch.float32 and loss_host.dtype == torch.float32 and loss_host.is_pinned()
            and loss_host.is_contiguous() and loss.numel() == loss_host.numel()
            and os.environ.get("FDSM_E2E_LOSS_MEMCPY") != "1"):
        return ops.publish_scalars(loss.reshape(-1), loss_host)
    loss_host.copy_(loss, non_blocking=True)
    return loss_host


class FDSSMTrainer:
    """loss = "ssm_fd" (train_ssm.py) or "mdsm_fd" (train_dsm.py: pass the (B_global, 1) per-batch-position noise
    schedule of train_dsm.py:50-57 as `sigmas`; each rank uses its own slice of it)."""

    def __init__(self, net, *, eps=0.05, lr=1e-5, betas=(0.9, 0.95), clip=0.1, loss="ssm_fd", seed=1234, sigmas=None,
                 sigma02=0.01, loss_coe=1e-4, scale=None):
        if loss not in ("ssm_fd", "mdsm_fd"):
            raise ValueError("loss must be 'ssm_fd' or 'mdsm_fd'")
        if loss == "mdsm_fd" and sigmas is None:
            raise ValueError("mdsm_fd needs the per-batch-position noise schedule `sigmas`")
        self.net = net
        dev = next(net.parameters()).device
        self.host_cpus = bind_host_thread_to_gpu_node(dev.index if dev.index is not None else 0) if dev.type == "cuda" else None
        self.eps, self.clip, self.loss_name = eps, clip, loss
        self.sigmas, self.sigma02, self.loss_coe = sigmas, sigma02, loss_coe
        self.scale = scale      # None: 1/eps^2 (loss_train.py:232); efficient_sm.py:71 uses D/eps^2
        self.fused = hasattr(net, "fd_ssm_step")       # mlp.FusedEnergyMLP: kernels (1)-(4) + Adam in CUDA
        rank = torch.distributed.get_rank() if ddp.is_distributed() else 0
        if self.fused:
            self.flat = net.flat_params()
            net.configure_optimizer(lr=lr, betas=betas, clip=clip, seed=seed, rank=rank)
        else:
            self.flat = ddp.FlatParams(net)
            self.opt = torch.optim.Adam([self.flat.flat], lr=lr, betas=betas, fused=self.flat.flat.is_cuda)
            self.flat.flat.grad = self.flat.grad
        ddp.broadcast_params_(self.flat.flat.data)
        self._allreduce = ddp.make_allreduce(self.flat.flat.numel(), self.flat.flat.device) if self.fused else ddp.allreduce_mean_
        ops.manual_seed(seed)
        self._rank = rank
        self._copy_stream = None
        self._staged = None
        self._nstep = 0
        self.peer_check_every = 256     # steps between host reads of the peer exchange's status word (one D2H sync each)

    def check_exchange(self):
        """Raises if the peer-memory all-reduce ever gave up waiting for a rank (fd_peer.cu: the launch then publishes NaN
        squared norms and stops advancing the step counter).  Called every `peer_check_every` steps by step(); call it
        before saving a checkpoint."""
        timed_out = getattr(self._allreduce, "timed_out", None)
        if timed_out is not None and timed_out():
            raise RuntimeError("fdsm_b200: the gradient exchange timed out waiting for a peer rank; the parameters of this "
                               "rank are no longer trustworthy -- restart from the last checkpoint")

    # -- one optimisation step on a device-resident batch ------------------------------------------
    def step(self, x, borrow=False, x_noise=0.0):
        self._nstep += 1
        if self._nstep % self.peer_check_every == 0:
            self.check_exchange()
        B = x.shape[0]
        ops.set_row_offset(self._rank * B)
        sig = None
        if self.loss_name == "mdsm_fd":      # this rank's slice of the positional schedule
            sig = ddp.shard_sigmas(self.sigmas.reshape(-1))
            if sig.numel() != B:
                raise RuntimeError(f"the sigma schedule gives this rank {sig.numel()} entries for a local batch of {B}")
        if self.fused:
            if sig is not None:
                return self.net.mdsm_fd_step(x, sig, self.sigma02, self.eps, self.loss_coe, allreduce=self._allreduce,
                                             distributed=ddp.is_distributed())
            kw = {} if self.scale is None else {"scale": self.scale}
            if borrow:
                kw["borrow"] = True
            if x.dtype == torch.uint8:
                kw["x_noise"] = x_noise
            return self.net.fd_ssm_step(x, eps=self.eps, allreduce=self._allreduce, distributed=ddp.is_distributed(), **kw)
        self.flat.zero_grad()
        if sig is not None:
            loss = losses.mdsm_fd(self.net, x, sig, self.sigma02, int(x[0].numel()), eps=self.eps, batch_size=B,
                                  loss_coe=self.loss_coe)
        else:
            loss = losses.ssm_fd(self.net, x, eps=self.eps, batch_size=B)
            if self.scale is not None:
                loss = loss * (self.scale * self.eps ** 2)
        loss.backward()
        ddp.allreduce_mean_(self.flat.grad)
        if self.clip is not None:
            # clip_grad_norm_ on the flat bucket: one norm kernel, no host sync
            nrm = torch.linalg.vector_norm(self.flat.grad)
            self.flat.grad.mul_(torch.clamp(self.clip / (nrm + 1e-6), max=1.0))
        self.opt.step()
        return loss.detach()

    # -- end-to-end: host batch in pinned memory -> loss on the host --------------------------------
    def step_from_host_u8(self, x_host_u8, loss_host, noise_std):
        """The same end-to-end step for a pinned host batch of uint8 PIXELS (what the reference's loaders hold before
        ToTensor): 1 byte per pixel over the link instead of 4; x / 255 + noise_std * N(0,1) (DeepEBM/train_ssm.py:61-62)
        is built on the device inside the step's input kernel.  ssm_fd on the fused difference-propagating net only."""
        if x_host_u8.dtype != torch.uint8:
            raise ValueError("step_from_host_u8 takes a uint8 batch")
        return self.step_from_host(x_host_u8, loss_host, x_noise=noise_std)

    def step_from_host(self, x_host, loss_host, x_noise=0.0):
        """H2D copy of this step's batch (pinned -> device, one copy stream per staging buffer, double-buffered against the
        previous step's compute), the step, and the loss into `loss_host` (pinned: stored by a kernel).  After the first two
        calls for a batch shape (which capture the step's graph for each staging buffer) a call is two library calls and
        one graph replay: `fdsm_stage_h2d` (wait for the buffer, copy, record, make the step wait), the replay,
        `fdsm_publish_scalars_record` -- the host loop stays well under the 100 us of a device step also on a busy host."""
        import ctypes as C
        cur = torch.cuda.current_stream()
        if self._copy_stream is None:
            # one copy stream per staging buffer: batch n + 1 is handed to the copy engines while batch n is still in
            # flight, so the link never idles between two copies (FDSM_E2E_COPY_STREAMS=1: a single stream)
            n = 1 if os.environ.get("FDSM_E2E_COPY_STREAMS") == "1" else 2
            self._copy_stream = [torch.cuda.Stream() for _ in range(n)]
            self._staging = {}
        key = (tuple(x_host.shape), x_host.dtype, float(x_noise))
        sg = self._staging.get(key)
        if sg is None:      # one double buffer per batch shape (the last batch of an epoch may be shorter)
            sg = self._staging[key] = dict(bufs=[torch.empty_like(x_host, device="cuda") for _ in range(2)],
                                           evt=[torch.cuda.Event() for _ in range(2)], free=[torch.cuda.Event() for _ in range(2)], i=0,
                                           lean=[None, None])
            for e in sg["evt"] + sg["free"]:
                e.record(cur)                        # creates the cudaEvent_t handles the lean path passes to the library
        i = sg["i"]
        sg["i"] ^= 1
        lean = sg["lean"][i]
        if lean is not None and x_host.is_pinned() and x_host.is_contiguous():
            lib, g1, nlaunch, dst, nbytes, cs, ev_free, ev_ready, loss_dev, loss_ptr = lean
            if loss_host.data_ptr() == loss_ptr:
                main = C.c_void_p(cur.cuda_stream)
                rc = lib.fdsm_stage_h2d(dst, C.c_void_p(x_host.data_ptr()), nbytes, cs, ev_free, ev_ready, main)
                if rc == 0:
                    g1.replay()
                    rc = lib.fdsm_publish_scalars_record(loss_dev, C.c_void_p(loss_ptr), 1, main, ev_free)
                if rc:
                    from ._lib import check
                    check(rc, "step_from_host")
                ops._count(nlaunch + 1)
                self._nstep += 1
                if self._nstep % self.peer_check_every == 0:
                    self.check_exchange()
                return loss_host
        self._bufs, self._evt, self._free = sg["bufs"], sg["evt"], sg["free"]
        cs = self._copy_stream[i % len(self._copy_stream)]
        with torch.cuda.stream(cs):
            cs.wait_event(self._free[i])      # previous consumer of this buffer is done
            self._bufs[i].copy_(x_host, non_blocking=True)
            self._evt[i].record(cs)
        cur.wait_event(self._evt[i])
        # the fused ssm_fd step reads the staging buffer where it lies (one captured graph per staging buffer): no D2D copy
        fast = self.fused and self.loss_name == "ssm_fd"
        loss = self.step(self._bufs[i], borrow=fast, x_noise=x_noise)
        self._free[i].record(cur)
        # remember what a later call for this buffer replays: the single graph of the step just taken (none if the step ran
        # eagerly or needs a library collective between two graphs), the staging pointers and event handles
        graphs = getattr(self.net, "_last_graphs", None) if fast else None
        if (graphs is not None and graphs[1] is None and loss_host.is_pinned() and loss_host.dtype == torch.float32
                and loss_host.numel() == 1 and loss.dtype == torch.float32 and os.environ.get("FDSM_E2E_LEAN", "1") != "0"
                and os.environ.get("FDSM_E2E_LOSS_MEMCPY") != "1"):
            from . import _lib
            sg["lean"][i] = (_lib.load(), graphs[0], graphs[2], C.c_void_p(self._bufs[i].data_ptr()),
                             self._bufs[i].numel() * self._bufs[i].element_size(), C.c_void_p(cs.cuda_stream),
                             C.c_void_p(self._free[i].cuda_event), C.c_void_p(self._evt[i].cuda_event),
                             C.c_void_p(loss.data_ptr()), loss_host.data_ptr())
        return read_loss_to_host(loss, loss_host)


class GenericFDTrainer:
    """The same hot loop for ANY network in generic mode (configs 3-5: Res18Quadratic, CondRefineNetDilated, NCSNv2 --
    or the caller's own module): `loss_fn(net, x) -> scalar` is one of the adapters in `losses` (kernels (1)+(2) and (4)
    of ours around the caller's torch network), then zero_grad / backward / ONE all-reduce of the flat gradient / optional
    global-norm clip / Adam / optional EMA (ncsnv2/runners/ncsn_esm_runner.py:146-151) -- the last three each ONE launch on
    the flat parameter vector (`ddp.FlatParams`)."""

    def __init__(self, net, loss_fn, *, lr=1e-4, betas=(0.9, 0.999), clip=None, ema_rate=None, seed=1234):
        self.net, self.loss_fn, self.clip = net, loss_fn, clip
        self.flat = ddp.FlatParams(net)
        self.opt = torch.optim.Adam([self.flat.flat], lr=lr, betas=betas, fused=self.flat.flat.is_cuda)
        self.flat.flat.grad = self.flat.grad
        ddp.broadcast_params_(self.flat.flat.data)
        self.ema_rate = ema_rate
        self.ema_shadow = self.flat.flat.data.clone() if ema_rate else None
        self._rank = torch.distributed.get_rank() if ddp.is_distributed() else 0
        ops.manual_seed(seed)
        self._h2d = None

    def step(self, x):
        ops.set_row_offset(self._rank * x.shape[0])
        self.flat.zero_grad()
        loss = self.loss_fn(self.net, x)
        loss.backward()
        ddp.allreduce_mean_(self.flat.grad)
        if self.clip:
            nrm = torch.linalg.vector_norm(self.flat.grad)
            self.flat.grad.mul_(torch.clamp(self.clip / (nrm + 1e-6), max=1.0))
        self.opt.step()
        if self.ema_shadow is not None:
            ops.ema_update(self.ema_shadow, self.flat.flat.data, self.ema_rate)
        return loss.detach()

    def step_from_host(self, x_host, loss_host):
        """Pinned host batch -> device (copy stream, double-buffered) -> step -> async D2H of the loss."""
        if self._h2d is None:
            self._h2d = HostStager()
        x = self._h2d.stage(x_host)
        loss = self.step(x)
        self._h2d.release()
        return read_loss_to_host(loss, loss_host)


class HostStager:
    """Double-buffered pinned-host -> device staging on a copy stream (one pair of device buffers per batch shape)."""

    def __init__(self):
        self.stream = torch.cuda.Stream()
        self.slots = {}
        self._last = None

    def stage(self, x_host):
        cur = torch.cuda.current_stream()
        key = (tuple(x_host.shape), x_host.dtype)
        if key not in self.slots:
            self.slots[key] = dict(bufs=[torch.empty_like(x_host, device="cuda") for _ in range(2)],
                                   ready=[torch.cuda.Event() for _ in range(2)], free=[torch.cuda.Event() for _ in range(2)], i=0)
        sg = self.slots[key]
        i = sg["i"]
        sg["i"] ^= 1
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(sg["free"][i])
            sg["bufs"][i].copy_(x_host, non_blocking=True)
            sg["ready"][i].record(self.stream)
        cur.wait_event(sg["ready"][i])
        self._last = (sg, i)
        return sg["bufs"][i]

    def release(self):
        sg, i = self._last
        sg["free"][i].record(torch.cuda.current_stream())
