"""The training step of the reference's Trainer (explainn/selene/__init__.py:283-295) fed from HOST batches.

    inputs, targets -> device; predictions = model(inputs); loss = criterion(...); zero_grad;
    loss.backward(); optimizer.step(); loss.item()

The reference does these strictly in sequence: a blocking fp32 H2D of the one-hot batch (16*L bytes per
sequence), the step, then `loss.item()` which drains the device every step.  `StepPipeline` keeps the same
per-step work -- every batch is copied from pinned host memory, every loss is read back to the host -- but
  (i)   batch i+1 is copied on a copy stream into the other of two device slots while step i computes;
  (ii)  loss i is copied to pinned host memory asynchronously and handed back one call later, so the device never
        idles on the host;
  (iii) the host batch may be the reference's float one-hot (B,4,L), the raw ASCII bases uint8 (B,L) (what
        train.py:266-284 reads before `one_hot_encode`; 16x fewer bytes over PCIe, encoded on the device by
        xnn_pack_ascii) or already packed planes int32 (B,3,ceil(L/32)) (56x fewer);
  (iv)  with graph=True the whole step -- forward, loss, backward, gradient all-reduce, Adam -- is captured ONCE into
        a CUDA graph over static buffers and replayed: one graph launch per step (plus the encode of the incoming
        batch).  Adam's step count and the dropout draw index live in device memory for that (optim.FusedAdam,
        UnitBank.enable_device_draws).
Gradient all-reduce over ranks goes through parallel.GradBucket (flat arena, overlapped with the backward).
"""
import torch

from . import ops
from .architectures import PackedDNA


class StepPipeline:
    def __init__(self, model, criterion, optimizer, bucket=None, device=None, graph=False, graph_warmup=2):
        self.model, self.criterion, self.optimizer, self.bucket = model, criterion, optimizer, bucket
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        if self.device.type != "cuda":
            raise RuntimeError("StepPipeline runs on a CUDA device only (explainn_b200 has no CPU fallback)")
        self.copy_stream = torch.cuda.Stream(self.device)
        self.capture_stream = torch.cuda.Stream(self.device)
        self.slots = [None, None]            # (x_dev, y_dev)
        self.copied = [torch.cuda.Event(), torch.cuda.Event()]
        self.consumed = [None, None]         # event: the step that read slot s has finished with it
        self.loss_host = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(2)]   # (loss, bad-input flag)
        self.loss_ready = [None, None]
        self.i = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.use_graph, self.graph_warmup = bool(graph), max(1, int(graph_warmup))   # optimizer state exists before capture
        self.graph = None
        self.static = None                   # (planes, y, loss) the captured graph reads / writes
        self.graph_kernels = 0               # kernel launches captured into the graph (ops' own count)
        self.length = None
        model.linears.input_check = "lazy"   # non-one-hot float batches are reported with the loss, without a sync
        model.linears._bad_flag(self.device)
        if hasattr(optimizer, "enable_device_step") and self.use_graph:
            optimizer.enable_device_step(self.device)
        if self.use_graph:
            model.linears.enable_device_draws(self.device)
            if getattr(model, "_arena_buf", None) is None:
                model.use_grad_arena()       # graph replays need the gradients at stable addresses

    # ---- host -> device ------------------------------------------------------------------------------------
    def _slot_buffers(self, s, x_host, y_host):
        cur = self.slots[s]
        if cur is None or cur[0].shape != x_host.shape or cur[0].dtype != x_host.dtype or cur[1].shape != y_host.shape:
            # a new block may still be in use by queued compute-stream kernels of earlier steps (the host runs ahead
            # of the device): allocate on the copy stream only after it has caught up with the compute stream
            self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.copy_stream):
                cur = (torch.empty(x_host.shape, dtype=x_host.dtype, device=self.device),
                       torch.empty(y_host.shape, dtype=y_host.dtype, device=self.device))
            for t in cur:
                t.record_stream(torch.cuda.current_stream(self.device))
            self.slots[s] = cur
            self.consumed[s] = None
        return cur

    def _as_input(self, x_dev, L=None):
        """device batch -> what the model consumes: float one-hot as is; ASCII / planes as PackedDNA"""
        if x_dev.dtype == torch.float32:
            return x_dev
        if x_dev.dtype == torch.uint8:
            return PackedDNA(ops.pack_ascii(x_dev), x_dev.shape[1])
        if x_dev.dtype == torch.int32:
            if L is None:
                raise ValueError("packed planes need sequence_length: StepPipeline.step(planes, y, length=L)")
            return PackedDNA(x_dev, L)
        raise TypeError("host batch: float32 one-hot (B,4,L), uint8 ASCII (B,L) or int32 planes (B,3,W)")

    # ---- the step itself ---------------------------------------------------------------------------------
    def _step_body(self, x_in, y_dev):
        pred = self.model(x_in)
        loss = self.criterion(pred, y_dev)
        self.optimizer.zero_grad()
        loss.backward()
        if self.bucket is not None and self.bucket.world > 1:
            if self.bucket.model is not None:
                self.bucket.allreduce_sum()
                self.optimizer.step(grad_scale=self.bucket.scale)
            else:
                self.bucket.allreduce_mean()
                self.optimizer.step()
        else:
            self.optimizer.step()
        return loss

    def _to_planes(self, x_dev, L, out=None):
        """ASCII / float / planes slot -> planes (into `out` when given: the graph's static input)"""
        if x_dev.dtype == torch.uint8:
            return ops.pack_ascii(x_dev, out=out)
        if x_dev.dtype == torch.float32:
            bad = self.model.linears._bad_flag(self.device)
            return ops.pack_onehot(x_dev, bad=bad, out=out)
        if out is None:
            return x_dev
        out.copy_(x_dev, non_blocking=True)
        return out

    def _capture(self, x_dev, y_dev, L):
        B = x_dev.shape[0]
        planes = torch.empty((B, 3, ops.planes_words(L)), dtype=torch.int32, device=self.device)
        y_static = torch.empty_like(y_dev)
        self._to_planes(x_dev, L, out=planes)
        y_static.copy_(y_dev, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        g = torch.cuda.CUDAGraph()
        k0 = ops.kernel_launches
        with torch.cuda.graph(g, stream=self.capture_stream):
            loss = self._step_body(PackedDNA(planes, L), y_static).detach()
        self.graph_kernels = ops.kernel_launches - k0
        if hasattr(self.optimizer, "note_replays"):
            self.optimizer.note_replays(-1)         # the capture counted one host step without executing it
        self.graph, self.static, self.length = g, (planes, y_static, loss), L
        self.static_shape = (tuple(x_dev.shape), x_dev.dtype, tuple(y_dev.shape))

    def step(self, x_host, y_host, length=None):
        """Enqueue one training step on a pinned host batch.  Returns the loss of the PREVIOUS step
        (None on the first call); `flush()` returns the last one."""
        if not (x_host.is_pinned() and y_host.is_pinned()):
            x_host, y_host = x_host.pin_memory(), y_host.pin_memory()
        s = self.i & 1
        compute = torch.cuda.current_stream(self.device)
        x_dev, y_dev = self._slot_buffers(s, x_host, y_host)
        with torch.cuda.stream(self.copy_stream):
            if self.consumed[s] is not None:
                self.copy_stream.wait_event(self.consumed[s])
            x_dev.copy_(x_host, non_blocking=True)
            y_dev.copy_(y_host, non_blocking=True)
            self.copied[s].record(self.copy_stream)
        self.h2d_bytes += x_host.numel() * x_host.element_size() + y_host.numel() * y_host.element_size()
        compute.wait_event(self.copied[s])
        return self._run(s, x_dev, y_dev, length, compute)

    def step_device(self, x_dev, y_dev, length=None):
        """The same step on a batch that is ALREADY in device memory (float one-hot, ASCII or planes): no H2D copy;
        the loss still comes back one call late.  bench.py times `value` through this."""
        return self._run(self.i & 1, x_dev, y_dev, length, torch.cuda.current_stream(self.device))

    def _run(self, s, x_dev, y_dev, length, compute):
        L = length if x_dev.dtype == torch.int32 else x_dev.shape[-1]
        if L is None:
            raise ValueError("packed planes need sequence_length: StepPipeline.step(planes, y, length=L)")
        if self.use_graph and self.i >= self.graph_warmup:
            shape = (tuple(x_dev.shape), x_dev.dtype, tuple(y_dev.shape))
            if self.graph is None or self.static_shape != shape or self.length != L:
                self._capture(x_dev, y_dev, L)      # (re)capture on a new batch shape, e.g. the short last batch
            else:
                self._to_planes(x_dev, L, out=self.static[0])
                self.static[1].copy_(y_dev, non_blocking=True)
            self.graph.replay()
            if hasattr(self.optimizer, "note_replays"):
                self.optimizer.note_replays(1)
            loss = self.static[2]
        elif self.use_graph:
            # warm-up steps of a pipeline that will capture: run them on the capture stream, so that the autograd
            # nodes that outlive an iteration (AccumulateGrad) belong to the stream the graph is captured on
            cs = self.capture_stream
            cs.wait_stream(compute)
            with torch.cuda.stream(cs):
                loss = self._step_body(self._as_input(x_dev, L), y_dev).detach()
            compute.wait_stream(cs)
        else:
            loss = self._step_body(self._as_input(x_dev, L), y_dev).detach()
        lh = self.loss_host[s]
        lh[0:1].copy_(loss.reshape(1), non_blocking=True)
        lh[1:2].copy_(self.model.linears._bad_flag(self.device).view(torch.float32), non_blocking=True)
        self.d2h_bytes += 8
        ev = torch.cuda.Event()
        ev.record(compute)
        self.consumed[s] = ev
        self.loss_ready[s] = ev
        self.i += 1
        return self._previous(s ^ 1)

    def _previous(self, s):
        ev = self.loss_ready[s]
        if ev is None:
            return None
        ev.synchronize()
        self.loss_ready[s] = None
        if self.loss_host[s][1:2].view(torch.int32).item() != 0:
            self.model.linears._bad_flag(self.device).zero_()
            raise ValueError("input is not one-hot: a float batch had columns that are neither one-hot nor all-zero "
                             "(explainn_b200 computes on the packed 2-bit layout)")
        return float(self.loss_host[s][0])

    def flush(self):
        """Loss of the most recent step (waits for it)."""
        return self._previous((self.i - 1) & 1) if self.i else None

    def close(self):
        """Release the captured graph and its static buffers.  A CUDA graph that holds captured NCCL kernels must be
        destroyed BEFORE its process group is: `destroy_process_group()` waits for it otherwise."""
        if self.i:
            torch.cuda.current_stream(self.device).synchronize()
        self.graph, self.static = None, None
        self.slots = [None, None]


class PackedDataset:
    """The training set kept PACKED in device memory ("keep data 2-bit-packed on device"): the
    reference re-encodes and re-copies its float one-hot batch every step (train.py:266-295: 16*L bytes per sequence
    over PCIe per step); here the whole set is packed once -- 12*ceil(L/32) bytes per sequence, 228 B at 600 bp, so
    10^6 sequences are 228 MB of the 180 GB -- and a step takes a slice or a gathered batch of it.

        ds = PackedDataset.from_ascii(seqs_uint8, targets, device)      # or .from_onehot(x_float, targets, device)
        for planes, y in ds.batches(4096, shuffle=True, generator=g):
            pipe.step_device(planes, y, length=ds.length)
    """

    def __init__(self, planes, targets, length):
        self.planes, self.targets, self.length = planes, targets, int(length)

    @classmethod
    def from_ascii(cls, seqs, targets, device, chunk=1 << 16):
        """seqs: uint8 (N, L) ASCII bases (host or device); packed on the device chunk by chunk"""
        N, L = seqs.shape
        planes = torch.empty((N, 3, ops.planes_words(L)), dtype=torch.int32, device=device)
        for i in range(0, N, chunk):
            ops.pack_ascii(seqs[i:i + chunk].to(device, non_blocking=True).contiguous(), out=planes[i:i + chunk])
        return cls(planes, torch.as_tensor(targets, dtype=torch.float32).to(device), L)

    @classmethod
    def from_onehot(cls, x, targets, device, chunk=1 << 14):
        """x: float32 one-hot (N, 4, L) (host or device); every column is checked to be one-hot or all-zero"""
        N, _, L = x.shape
        planes = torch.empty((N, 3, ops.planes_words(L)), dtype=torch.int32, device=device)
        for i in range(0, N, chunk):
            ops.pack_onehot(x[i:i + chunk].to(device, non_blocking=True), check=True, out=planes[i:i + chunk])
        return cls(planes, torch.as_tensor(targets, dtype=torch.float32).to(device), L)

    def __len__(self):
        return self.planes.shape[0]

    def batches(self, batch_size, shuffle=False, generator=None, drop_last=False):
        """(planes (B,3,W) int32, targets (B,F)) device tensors; contiguous slices, or gathered rows when shuffled"""
        N = len(self)
        order = torch.randperm(N, device=self.planes.device, generator=generator) if shuffle else None
        for i in range(0, N, batch_size):
            if drop_last and i + batch_size > N:
                break
            if order is None:
                yield self.planes[i:i + batch_size], self.targets[i:i + batch_size]
            else:
                idx = order[i:i + batch_size]
                yield self.planes.index_select(0, idx), self.targets.index_select(0, idx)
