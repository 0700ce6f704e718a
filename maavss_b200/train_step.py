"""One optimizer step of the reference training loops, for a batch that is already resident in HBM.

`FusionTrainStep`  = train.py:132-162            (avse_model.AV_Fusion_Model / avse_model_conv.AV_Fusion_Model):
    for j in range(num_seq): phasegram of 8 attention frames, 64 STFT frames, forward, MSE(a) + loss_coeff * MSE(v),
    / num_seq, backward (gradients accumulate in the flat arena); then Adam, zero_grad.
`FramesTrainStep`  = train_avse_frames.py:150-181 (avse_model_final.AV_Fusion_Model_Frames): raw attention frames in,
    the middle frame's STFT hops and attention frame as targets.

B200 execution model: the whole step (4 forward/backward micro-steps, the data-parallel gradient all-reduce and the
Adam kernel) is captured ONCE into a CUDA graph over static input buffers and replayed per step -- no tracing compiler,
no per-op Python on the replay path.  `graph=False` runs the same code eagerly (what an unchanged train.py does).
"""
import os

import torch

from . import layers, nn_ops, parallel, phasegram_ops


REPORT_STREAM = os.environ.get("MAAVSS_REPORT_STREAM", "1") != "0"      # benchmarks only: 0 keeps the reported losses on the main stream


class _TrainStepBase:
    def __init__(self, model, optimizer, graph=True, warmup=2, async_wgrad=True):
        self.model, self.opt = model, optimizer
        self.async_wgrad = async_wgrad
        self.use_graph, self.warmup = graph, warmup
        self._graph, self._static_in, self._static_out = None, None, None
        self.launches_per_step = None

    def micro_steps(self, *inputs):
        raise NotImplementedError

    early = None      # parallel.EarlyGradAllReduce, set by subclasses that know where the arena can be split

    _running = False

    def _eager(self, *inputs):
        self._running = True          # the LSTM hook of another step object built on the same model stays quiet
        prev_async = nn_ops.ASYNC_WGRAD
        nn_ops.ASYNC_WGRAD = self.async_wgrad
        try:
            out = self.micro_steps(*inputs)
        finally:
            self._running = False
            nn_ops.ASYNC_WGRAD = prev_async
        nn_ops.join_wgrad_streams()   # weight gradients ran on companion streams: all of them before the exchange / Adam
        if self.early is not None:
            self.early.finish()           # exchange + Adam + zero_grad of whatever the early hook has not done yet
        else:
            parallel.allreduce_gradients(self.opt)
            self.opt.step_and_zero_grad()
        return out

    def _capture(self, inputs):
        self._static_in = [torch.empty_like(t).copy_(t) for t in inputs]
        # Warm-up steps (allocator / lazy plan state / NCCL channels) run on a side stream; the training state they touch
        # -- parameters, Adam moments and step counter, BatchNorm running statistics -- is snapshotted and put back, so
        # the first replay is the first optimizer step, as in the reference loop.
        opt = self.opt
        keep = [t.clone() for t in (opt.flat.data, opt.exp_avg, opt.exp_avg_sq, opt.dev_state)]
        bufs = [(b, b.clone()) for b in self.model.buffers()]
        count = opt.step_count
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup):
                self._eager(*self._static_in)
        torch.cuda.current_stream().wait_stream(side)
        with torch.no_grad():
            for dst, src in zip((opt.flat.data, opt.exp_avg, opt.exp_avg_sq, opt.dev_state), keep):
                dst.copy_(src)
            for b, v in bufs:
                b.copy_(v)
        opt.step_count = count
        torch.cuda.synchronize()
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._static_out = self._eager(*self._static_in)
        # the capture itself did not execute: host mirror of the step counter must not count it
        self.opt.step_count = count

    def __call__(self, *inputs):
        if not self.use_graph:
            return self._eager(*inputs)
        if self._graph is None:
            self._capture(inputs)
        for s, t in zip(self._static_in, inputs):
            if s.data_ptr() != t.data_ptr():
                s.copy_(t, non_blocking=True)
        self._graph.replay()
        self.opt.step_count += 1
        return self._static_out

    def static_inputs(self):
        """The graph's input buffers: a producer (H2D copy, front-end kernels) may write them directly."""
        return self._static_in


class FusionTrainStep(_TrainStepBase):
    """batch_windows=True (default): the num_seq accumulation windows are stacked along the batch axis and run as ONE
    forward/backward pass -- the weights are constant between them (the optimizer steps only after the last one), so
    the result is the reference's sum of per-window gradients; train-mode BatchNorm keeps per-window statistics
    (nn_ops.BN_GROUPS) and applies its running-statistics updates in window order.  Every weight matrix is then
    streamed from HBM once per optimizer step instead of num_seq times and the launch count drops by num_seq."""

    def __init__(self, model, optimizer, num_seq=4, num_frames=8, hops_per_frame=8, p_size=64, loss_coeff=0.001,
                 batch_windows=True, overlap_allreduce=True, global_phasegram=True, **kw):
        super().__init__(model, optimizer, **kw)
        self.global_phasegram = global_phasegram
        self.num_seq, self.num_frames, self.hpf, self.p_size, self.loss_coeff = num_seq, num_frames, hops_per_frame, p_size, loss_coeff
        self.batch_windows = batch_windows
        self._setup_early_allreduce(overlap_allreduce)

    def _setup_early_allreduce(self, enabled):
        """avse_model.AV_Fusion_Model.forward trains [encoders | lstm | fc1 fc2 (decoders: unused) a_fc1 v_fc1]: the fc /
        head part is final (gradients written, weights no longer read) once the LSTM output's gradient exists, the LSTM
        part once the LSTM input's gradient exists; parallel.EarlyGradAllReduce takes them from there."""
        m = self.model
        if not enabled or getattr(m, "VARIANT", None) != "fc" or not self.batch_windows:
            return
        first = next(m.lstm.parameters())
        head = next(m.fc1.parameters())
        flat = self.opt.flat
        if id(first) not in flat.index or id(head) not in flat.index:
            return
        off, head_off = flat.offsets[flat.index[id(first)]], flat.offsets[flat.index[id(head)]]
        # the claim above, checked: every parameter behind fc1 in the arena belongs to fc1 / fc2 / the heads / the unused decoders
        tail_ok = {id(p) for mod in (m.fc1, m.fc2, m.a_fc1, m.v_fc1, m.stft_decoder, m.phasegram_decoder) for p in mod.parameters()}
        if head_off < off or any(id(p) not in tail_ok for p, o in zip(flat.params, flat.offsets) if o >= head_off):
            head_off = None
        out_off = None
        if head_off is not None:
            outs = {id(p) for mod in (m.a_fc1, m.v_fc1) for p in mod.parameters()}
            first_out = min((o for p, o in zip(flat.params, flat.offsets) if id(p) in outs), default=None)
            # the output heads must be the END of the arena for their range [first_out, numel) to hold nothing else
            if first_out is not None and all(id(p) in outs for p, o in zip(flat.params, flat.offsets) if o >= first_out):
                out_off = first_out
        self.early = parallel.EarlyGradAllReduce(self.opt, off, head_offset=head_off, out_offset=out_off)

        def pre_hook(mod, args):
            x = args[0]
            if torch.is_tensor(x) and x.requires_grad:
                x.register_hook(lambda g: (self.early.start() if self._running else None, None)[1])

        def post_hook(mod, args, out):
            y = out[0] if isinstance(out, tuple) else out
            if torch.is_tensor(y) and y.requires_grad:
                y.register_hook(lambda g: (self.early.start_heads() if self._running else None, None)[1])
        m.lstm.register_forward_pre_hook(pre_hook)
        m.lstm.register_forward_hook(post_hook)

    def _pg_kw(self):
        """Data parallel with global_phasegram: `attn` holds the frames of parallel.rolled_clip_indices (the loader applied
        the whole-batch roll of utilities.py:210), the kernels roll nothing and the window maxima are all-reduced."""
        if self.global_phasegram and parallel.world_size() > 1:
            return dict(batch_roll=0, max_group=parallel.group())
        return {}

    def _windows_batched(self, x_stft, y_stft, attn):
        G, B = self.num_seq, x_stft.shape[0]
        span = self.hpf * self.num_frames
        def phasegrams():
            # the 12 distinct frames of a clip are transformed once, the 4 overlapping windows are index arithmetic
            return phasegram_ops.video_phasegram_windows(attn, G, self.num_frames, resize=(self.p_size, self.p_size), **self._pg_kw())
        if layers.PARALLEL_BRANCHES and attn.is_cuda:
            # the phasegrams feed the video encoder only: they start the side-stream branch (layers.run_pair continues it),
            # so the audio encoder does not wait for them
            cur, side = torch.cuda.current_stream(attn.device), layers._side_stream(attn.device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                pgram = phasegrams()
            pgram.record_stream(cur)
        else:
            pgram = phasegrams()
        xa = nn_ops.stack_windows(x_stft, G, self.hpf, span)
        ya = nn_ops.stack_windows(y_stft, G, self.hpf, span)
        nn_ops.BN_GROUPS = G
        try:
            yh_a, yh_v, fused = self.model(xa, pgram)
            if self.early is not None and fused.requires_grad:
                # both output heads have back-propagated when the gradient of their common input exists
                fused.register_hook(lambda g: (self.early.start_outputs() if self._running else None, None)[1])
            # What the sequential loop reports -- the last window's losses -- is a side-stream job: the backward pass is
            # seeded with d loss / d yh directly (same kernel and seeds as MSELoss.backward under loss.backward()), so no
            # loss reduction sits between the forward and the backward pass.
            cur = torch.cuda.current_stream(xa.device)
            rep = layers._report_stream(xa.device) if REPORT_STREAM else cur
            rep.wait_stream(cur)
            with torch.cuda.stream(rep), torch.no_grad():
                a_loss = nn_ops.mse_loss(yh_a[-B:], ya[-B:])
                v_loss = nn_ops.mse_loss(yh_v[-B:], pgram[-B:])
                last = (a_loss + self.loss_coeff * v_loss) / G
            # equal window sizes: the MSE over the stack is the mean of the per-window MSEs = sum_j (.)_j / num_seq
            ga = nn_ops.mse_grad(yh_a, ya, nn_ops.const_scalar(1.0, xa.device))
            gv = nn_ops.mse_grad(yh_v, pgram, nn_ops.const_scalar(self.loss_coeff, xa.device))
            torch.autograd.backward([yh_a, yh_v], [ga, gv])
            cur.wait_stream(rep)
            for t in (last, a_loss, v_loss):
                t.record_stream(cur)
        finally:
            nn_ops.BN_GROUPS = 1
        return last, a_loss, v_loss

    def micro_steps(self, x_stft, y_stft, attn):
        """x_stft, y_stft (B, 2, hpf*(num_frames+num_seq), F); attn (B, 1, num_frames+num_seq, H, W)."""
        if self.batch_windows:
            return self._windows_batched(x_stft, y_stft, attn)
        loss = a_loss = v_loss = None
        for j in range(self.num_seq):
            pgram = phasegram_ops.video_phasegram(attn[:, :, j:j + self.num_frames], resize=(self.p_size, self.p_size), **self._pg_kw())
            s0 = self.hpf * j
            s1 = s0 + self.hpf * self.num_frames
            yh_a, yh_v, _ = self.model(x_stft[:, :, s0:s1, :], pgram)
            a_loss = nn_ops.mse_loss(yh_a, y_stft[:, :, s0:s1, :])
            v_loss = nn_ops.mse_loss(yh_v, pgram)
            loss = (a_loss + self.loss_coeff * v_loss) / self.num_seq
            loss.backward()
        return loss.detach(), a_loss.detach(), v_loss.detach()


class FramesTrainStep(_TrainStepBase):
    """tensor_core=True runs the stride-1 Conv3d layers 2-4 of the visual encoder (94 % of the model's FLOPs) on the
    tcgen05 kernels (bf16 operands, fp32 accumulate: ~3e-3 relative error on those layers' outputs / gradients)."""

    def __init__(self, model, optimizer, num_seq=4, num_frames=8, hops_per_frame=8, loss_coeff=0.001, tensor_core=False, **kw):
        super().__init__(model, optimizer, **kw)
        self.num_seq, self.num_frames, self.hpf, self.loss_coeff = num_seq, num_frames, hops_per_frame, loss_coeff
        self.tensor_core = tensor_core

    def micro_steps(self, x_stft, y_stft, attn):
        prev = nn_ops.TENSOR_CORE_CONV
        nn_ops.TENSOR_CORE_CONV = bool(self.tensor_core) or prev
        try:
            return self._micro_steps(x_stft, y_stft, attn)
        finally:
            nn_ops.TENSOR_CORE_CONV = prev

    def _micro_steps(self, x_stft, y_stft, attn):
        loss = a_loss = v_loss = None
        mid = self.num_frames // 2
        for j in range(self.num_seq):
            frames = attn[:, :, j:j + self.num_frames]
            s0 = self.hpf * j
            s1 = s0 + self.hpf * self.num_frames
            m0 = s0 + self.hpf * mid                                   # train_avse_frames.py:158-162
            yh_a, yh_v, _ = self.model(x_stft[:, :, s0:s1, :], frames)
            a_loss = nn_ops.mse_loss(yh_a, y_stft[:, :, m0:m0 + self.hpf, :])
            v_loss = nn_ops.mse_loss(yh_v, attn[:, :, j + mid])
            loss = (a_loss + self.loss_coeff * v_loss) / self.num_seq
            loss.backward()
        return loss.detach(), a_loss.detach(), v_loss.detach()


class InputPrefetcher:
    """Double-buffered host -> device input pipeline for the train steps (generator.py feeds one clip batch per step):
    `put()` copies a batch of pinned host tensors on a copy stream, `get()` hands the oldest pending batch to the
    current stream (which waits for that copy only), `release()` returns its device buffers.  With one batch in flight
    the H2D copy of step i+1 (52 MB for 16 clips) overlaps the kernels of step i instead of preceding them."""

    def __init__(self, example, device, slots=2):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.bufs = [tuple(torch.empty_like(t, device=self.device) for t in example) for _ in range(slots)]
        self.ready = [torch.cuda.Event() for _ in range(slots)]
        self.free = [None] * slots           # event recorded when the consumer finished with the slot
        self.pending, self.head, self.held = [], 0, []

    def put(self, host_tensors):
        slot = self.head
        self.head = (self.head + 1) % len(self.bufs)
        assert slot not in self.pending and slot not in self.held, "InputPrefetcher: all slots in flight"
        with torch.cuda.stream(self.stream):
            if self.free[slot] is not None:
                self.stream.wait_event(self.free[slot])
            for d, h in zip(self.bufs[slot], host_tensors):
                d.copy_(h, non_blocking=True)
            self.ready[slot].record(self.stream)
        self.pending.append(slot)

    def get(self):
        slot = self.pending.pop(0)
        torch.cuda.current_stream(self.device).wait_event(self.ready[slot])
        self.held.append(slot)
        return self.bufs[slot]

    def release(self):
        slot = self.held.pop(0)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.free[slot] = ev
