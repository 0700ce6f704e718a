"""Data-parallel training over the GPUs of one NVSwitch box: one process per GPU, the clip batch sharded
contiguously across ranks.  The reference is single-process; N-rank semantics are pinned to
"N ranks == one process on the concatenated batch":

  * gradients: ONE all-reduce (sum) per optimizer step over the flat gradient arena (optim.FlatParams.grad; no
    bucketing copies, NVLink/NVSwitch makes it bandwidth- not latency-bound), averaged by Adam's grad_scale = 1/W;
  * train-mode BatchNorm: the per-channel partial sums are all-reduced between the two halves of the BN kernels
    (nn_ops.BNActFn, `sync_bn=True`), so statistics are those of the global batch;
  * parameters / buffers are broadcast from rank 0 at start.

torch.distributed (NCCL on GPUs, gloo in the CPU tests) is the plumbing; nothing here touches tensor values on the host.
"""
import os

import torch
import torch.distributed as dist

from . import nn_ops

_GROUP = None
_WORLD = 1
_COMM = None        # maavss_comm* (ctypes.c_void_p): the library's own NCCL communicator (csrc/comm.cu)
_PEER = None        # maavss_peer* : NVLink peer-memory exchange of the BatchNorm partial sums (csrc/peer_exchange.cu)
_PEER_CAP = 0
PEER_CHANNELS, PEER_MAX_DOUBLES = 1, 8192


def init(backend=None, sync_bn=True):
    """Initialise from the torchrun environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).  Returns (rank, world)."""
    global _GROUP, _WORLD
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    _WORLD = world
    _GROUP = dist.group.WORLD if world > 1 else None
    nn_ops.SYNC_BN_GROUP = _GROUP if (sync_bn and world > 1) else None
    if world > 1 and dist.get_backend() == "nccl" and os.environ.get("MAAVSS_TORCH_COLLECTIVES") != "1":
        _init_native_comm(rank, world)
        if sync_bn and world <= 16 and os.environ.get("MAAVSS_PEER_EXCHANGE", "1") != "0":
            _init_peer_exchange(rank, world)
    return rank, world


def _init_peer_exchange(rank, world):
    """One box, one process per GPU: every rank allocates its exchange buffer, the 64-byte CUDA IPC handles travel through
    torch.distributed once, every rank maps the others' buffers.  Falls back to NCCL when the GPUs cannot map each other."""
    global _PEER, _PEER_CAP
    import ctypes
    from ._lib import check, lib
    L = lib()
    peer = ctypes.c_void_p()
    handle = (ctypes.c_ubyte * 64)()

    def all_ok(flag):
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)          # every rank takes the same decision: all ranks or none
        return int(t.item()) == 1

    try:
        check(L.maavss_peer_create(ctypes.byref(peer), rank, world, PEER_CHANNELS, PEER_MAX_DOUBLES, handle))
        created = True
    except RuntimeError:
        created = False
    if not all_ok(created):
        return
    mine = torch.tensor(list(handle), dtype=torch.uint8, device="cuda")
    every = torch.empty(world * 64, dtype=torch.uint8, device="cuda")
    dist.all_gather_into_tensor(every, mine)
    buf = (ctypes.c_ubyte * (64 * world))(*every.cpu().tolist())
    try:
        check(L.maavss_peer_connect(peer, buf))
        connected = True
    except RuntimeError:
        connected = False
    if all_ok(connected):
        _PEER, _PEER_CAP = peer, int(L.maavss_peer_capacity(peer))


def peer_exchange():
    return _PEER


def peer_capacity():
    return _PEER_CAP


_EXCHANGE_STREAMS = {}


def exchange_stream(device):
    """The one stream per device on which every peer-memory exchange runs (total order = host issue order)."""
    s = _EXCHANGE_STREAMS.get(device)
    if s is None:
        s = _EXCHANGE_STREAMS[device] = torch.cuda.Stream(device=device)
    return s


def peer_allreduce_sum_(t, channel):
    """In-place sum of a small fp64 tensor over the ranks through NVLink peer memory; False when it does not qualify."""
    if _PEER is None or not (t.is_cuda and t.is_contiguous() and t.dtype == torch.float64 and 0 < t.numel() <= _PEER_CAP):
        return False
    import ctypes
    from ._lib import check, lib, stream_ptr
    check(lib().maavss_peer_allreduce_sum_f64(_PEER, int(channel), ctypes.c_void_p(t.data_ptr()), t.numel(), stream_ptr()))
    return True


def peer_status():
    """0 = fine; non-zero = a rank gave up waiting ~30 s for a peer (see csrc/peer_exchange.cu)."""
    if _PEER is None:
        return 0
    import ctypes
    from ._lib import check, lib
    err = ctypes.c_int(0)
    check(lib().maavss_peer_status(_PEER, ctypes.byref(err)))
    return err.value


def _init_native_comm(rank, world):
    """The hot-path collectives (gradient arena, BatchNorm partial sums, phasegram maxima) go through the library's C ABI
    (maavss_comm_init / maavss_grad_allreduce / ...); torch.distributed only carries the 128-byte ncclUniqueId to the ranks
    and stays the fallback (MAAVSS_TORCH_COLLECTIVES=1, and the gloo backend of the CPU tests)."""
    global _COMM
    import ctypes
    from ._lib import check, lib
    L = lib()
    ident = (ctypes.c_ubyte * 128)()
    if rank == 0:
        check(L.maavss_comm_unique_id(ident))
    t = torch.tensor(list(ident), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, src=0)
    ident = (ctypes.c_ubyte * 128)(*t.cpu().tolist())
    comm = ctypes.c_void_p()
    check(L.maavss_comm_init(ctypes.byref(comm), rank, world, ident))
    _COMM = comm


def native_comm():
    return _COMM


def allreduce_sum_(t):
    """In-place sum over ranks on the current stream (fp32 gradient slices, fp64 BatchNorm partial sums)."""
    if _GROUP is None:
        return t
    if _COMM is not None and t.is_cuda and t.is_contiguous() and t.dtype in (torch.float32, torch.float64):
        import ctypes
        from ._lib import check, lib, stream_ptr
        fn = lib().maavss_grad_allreduce if t.dtype == torch.float32 else lib().maavss_allreduce_sum_f64
        check(fn(_COMM, ctypes.c_void_p(t.data_ptr()), t.numel(), stream_ptr()))
        return t
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=_GROUP)
    return t


def world_size():
    return _WORLD


def group():
    """The data-parallel process group (None on one process)."""
    return _GROUP


def rolled_clip_indices(n_global, rank, world):
    """Global indices of the clips whose ATTENTION frames rank `rank` loads for the train.py path.

    utilities.video_phasegram fft-shifts every axis (utilities.py:210), so on one process the phasegram that meets the
    audio of clip i is that of clip (i - B//2) mod B of the WHOLE batch (reproduced, not fixed).  A rank that
    holds the audio of its contiguous shard [lo, hi) therefore reads the attention frames of
    (i - n_global//2) mod n_global for i in [lo, hi); its kernels then roll nothing (batch_roll=0) and the window maxima
    are all-reduced (utilities.py:226-227 normalises by the max of the whole batch)."""
    lo, hi = shard_bounds(n_global, rank, world)
    return [(i - n_global // 2) % n_global for i in range(lo, hi)]


def shard_bounds(n_items, rank, world, require_even=False):
    """Contiguous shard [lo, hi) of n_items for `rank`; the first n_items % world ranks get one extra item.
    require_even: raise unless every rank gets the same count -- the train steps need that (sync-BN partial shapes,
    the 1/W gradient average and the per-window MSE means all assume equal shards)."""
    base, rem = divmod(n_items, world)
    if require_even and rem:
        raise ValueError(f"{n_items} clips do not split evenly over {world} ranks: the data-parallel train steps need equal shards")
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def broadcast_module(module, src=0):
    if _GROUP is None:
        return
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src, group=_GROUP)


def allreduce_gradients(optimizer):
    """Sum the flat gradient arena over ranks; the 1/W average is folded into the Adam kernel (grad_scale)."""
    if _GROUP is None:
        return
    allreduce_sum_(optimizer.flat.grad)
    optimizer.grad_scale = 1.0 / _WORLD


# CTAs of the early Adam pass (0 = the kernel's own grid): 3 per SM measured best beside the LSTM backward (2.140 ms/step; 2.156 at 2 per SM, 2.25 unbounded)
EARLY_ADAM_CTAS = int(os.environ.get("MAAVSS_EARLY_ADAM_CTAS", "444"))
EARLY_ADAM = os.environ.get("MAAVSS_EARLY_ADAM", "1") != "0"


class EarlyGradAllReduce:
    """Take the tail of the optimizer step off the critical path.

    The flat arenas of the fusion model are [encoders | lstm | fc1 fc2 (decoders: unused) a_fc1 v_fc1].  Two hooks of the
    backward pass hand pieces of the arena to a side stream:

    * `start_heads()`, when the gradient of the LSTM OUTPUT exists: the gradients from `head_offset` (first fc1 weight) on
      -- 90 % of all parameters -- are final and no later kernel reads those weights.  They are all-reduced (N > 1) and,
      with `early_adam`, Adam-updated and zeroed right away: the 1.1 GB optimizer stream then runs beside the LSTM
      backward, which is 8 small latency-bound launches per direction and leaves HBM idle.  (Placed any later it runs
      beside the encoder backward, whose im2col matrices live in L2: measured slower than not overlapping at all.)
    * `start()`, when the gradient of the LSTM INPUT exists: the LSTM weights' gradients [offset, head_offset) are
      all-reduced (N > 1) while the encoders back-propagate.

    `finish()` is the rest of the optimizer step: exchange of what is left, Adam + zero_grad of what is left, joined with
    the side stream.  Everything is captured when the step is recorded into a CUDA graph."""

    def __init__(self, optimizer, offset, early_adam=True, head_offset=None, out_offset=None):
        # the early Adam pays only on one GPU: behind the heads' all-reduce (N > 1) it would land in the encoder backward, whose
        # im2col matrices live in L2 (measured at N = 2: 2.54 ms/step with it, 2.46 without)
        self.opt, self.offset, self.early_adam = optimizer, int(offset), early_adam and EARLY_ADAM and _GROUP is None
        self.head_offset = int(head_offset) if head_offset is not None else None
        # [out_offset, end): the output heads a_fc1 / v_fc1 (70 % of all parameters), final even earlier -- when the gradient
        # of their common input exists (start_outputs)
        self.out_offset = int(out_offset) if (out_offset is not None and head_offset is not None and out_offset > head_offset) else None
        self.side = torch.cuda.Stream()
        self.started = self.heads_started = self.outputs_started = False
        if _GROUP is not None:
            optimizer.grad_scale = 1.0 / _WORLD

    def _fork(self):
        self.side.wait_stream(torch.cuda.current_stream())
        nn_ops.join_wgrad_streams(target=self.side, clear=False)      # this part's weight gradients may still be in flight

    def _tail(self, lo, hi, prepare):
        self._fork()
        with torch.cuda.stream(self.side):
            if _GROUP is not None:
                allreduce_sum_(self.opt.flat.grad[lo:hi])
            if self.early_adam:
                if prepare:
                    self.opt.prepare()
                self.opt.apply_range(lo, hi, zero_grad=True, max_ctas=EARLY_ADAM_CTAS)

    def start_outputs(self):
        if self.outputs_started or self.heads_started or self.out_offset is None or (_GROUP is None and not self.early_adam):
            return
        self.outputs_started = True
        self._tail(self.out_offset, None, True)

    def start_heads(self):
        if self.heads_started or self.head_offset is None or (_GROUP is None and not self.early_adam):
            return
        self.heads_started = True
        self._tail(self.head_offset, self.out_offset if self.outputs_started else None, not self.outputs_started)

    def start(self):
        if self.started or _GROUP is None:
            return
        self.started = True
        self._fork()
        with torch.cuda.stream(self.side):
            hi = self.head_offset if self.heads_started else (self.out_offset if self.outputs_started else None)
            allreduce_sum_(self.opt.flat.grad[self.offset:hi])

    def finish(self):
        """Everything of optimizer.step(); optimizer.zero_grad() (and the gradient exchange) that the hooks have not done."""
        opt, g = self.opt, self.opt.flat.grad
        if _GROUP is not None:
            left = self.offset if self.started else (self.head_offset if self.heads_started else
                                                    (self.out_offset if self.outputs_started else g.numel()))
            if left > 0:
                allreduce_sum_(g[:left])
        if self.started or self.heads_started or self.outputs_started:
            torch.cuda.current_stream().wait_stream(self.side)
        if self.heads_started and self.early_adam:
            opt.apply_range(0, self.head_offset, zero_grad=True)
        elif self.outputs_started and self.early_adam:
            opt.apply_range(0, self.out_offset, zero_grad=True)
        else:
            opt.step_and_zero_grad()
        self.started = self.heads_started = self.outputs_started = False


def allreduce_max_(t):
    """In-place MAX over ranks; int32 / uint32 words (float bits of non-negative values) go through the C ABI."""
    if _GROUP is None:
        return t
    if _COMM is not None and t.is_cuda and t.is_contiguous() and t.dtype == torch.int32:
        import ctypes
        from ._lib import check, lib, stream_ptr
        check(lib().maavss_allreduce_max_u32(_COMM, ctypes.c_void_p(t.data_ptr()), t.numel(), stream_ptr()))
        return t
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=_GROUP)
    return t
