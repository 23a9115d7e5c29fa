"""The training step of the reference's Trainer (explainn/selene/__init__.py:283-295) fed from HOST batches.

    inputs, targets -> device; predictions = model(inputs); loss = criterion(...); zero_grad;
    loss.backward(); optimizer.step(); loss.item()

The reference does these strictly in sequence: a blocking fp32 H2D of the one-hot batch (16*L bytes per
sequence), the step, then `loss.item()` which drains the device every step.  `StepPipeline` keeps the same
per-step work -- every batch is copied from pinned host memory, every loss is read back to the host -- but
(i) batch i+1 is copied on a copy stream into the other of two device slots while step i computes, and
(ii) loss i is copied to pinned host memory asynchronously and handed back one call later, so the device
never idles on the host.  Gradient all-reduce over ranks goes through parallel.GradBucket.
"""
import torch


class StepPipeline:
    def __init__(self, model, criterion, optimizer, bucket=None, device=None):
        self.model, self.criterion, self.optimizer, self.bucket = model, criterion, optimizer, bucket
        self.device = device if device is not None else next(model.parameters()).device
        if self.device.type != "cuda":
            raise RuntimeError("StepPipeline runs on a CUDA device only (explainn_b200 has no CPU fallback)")
        self.copy_stream = torch.cuda.Stream(self.device)
        self.slots = [None, None]            # (x_dev, y_dev)
        self.copied = [torch.cuda.Event(), torch.cuda.Event()]
        self.consumed = [None, None]         # event: the step that read slot s has finished with it
        self.loss_host = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.loss_ready = [None, None]
        self.i = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _slot_buffers(self, s, x_host, y_host):
        cur = self.slots[s]
        if cur is None or cur[0].shape != x_host.shape or cur[1].shape != y_host.shape:
            cur = (torch.empty(x_host.shape, dtype=x_host.dtype, device=self.device),
                   torch.empty(y_host.shape, dtype=y_host.dtype, device=self.device))
            self.slots[s] = cur
        return cur

    def step(self, x_host, y_host):
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
        pred = self.model(x_dev)
        loss = self.criterion(pred, y_dev)
        self.optimizer.zero_grad()
        loss.backward()
        if self.bucket is not None:
            self.bucket.allreduce_mean()
        self.optimizer.step()
        self.loss_host[s].copy_(loss.detach().reshape(1), non_blocking=True)
        self.d2h_bytes += 4
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
        return float(self.loss_host[s][0])

    def flush(self):
        """Loss of the most recent step (waits for it)."""
        return self._previous((self.i - 1) & 1) if self.i else None
