"""Drop-in for the reference's `architectures` package on the hot path.

Mirrors explainn/architectures/__init__.py: `ExplaiNN` (:41-114), `PWM` (:116-170),
`get_loss` / `get_metrics` / `get_optimizer` (:446-464), with the same constructor
arguments, `_options`, `state_dict()` keys and shapes (`linears.0.weight` (U,4,k),
`linears.6.weight` (100U,n,1), `linears.10.weight` (U,100,1), `final.weight` (F,U) ...),
and a `linears` container that still answers `linears(x_rep)`, `linears[:3](x_rep)`
(test.py:151,159) and `linears[0].weight` (interpret.py:154, train.py:318-324).

Every layer object is a parameter holder only: the arithmetic runs in the fused
sm_100a kernels of explainn_b200/csrc through the C ABI (ops.py).  Calling a
single layer on its own raises -- there is no torch-library or CPU fallback.
"""
import math
from collections import OrderedDict

import torch
import torch.nn as nn

from . import ops
from . import torch_ops  # noqa: F401  (registers torch.ops.explainn_b200.*)

_FULL_KEYS = [str(i) for i in range(14)]


def _fused_only(self, *a, **k):
    raise RuntimeError(
        f"{type(self).__name__} is a parameter holder inside the fused ExplaiNN unit bank; call "
        "`model.linears(x)`, `model.linears[:3](x)`, `model.linears[:4](x)` or `model(x)` instead")


class FusedConv1d(nn.Conv1d):
    """nn.Conv1d initialisation, parameters and state_dict; no stand-alone forward."""
    forward = _fused_only


class FusedBatchNorm1d(nn.BatchNorm1d):
    forward = _fused_only


class ExpAct(nn.Module):
    """Exponential activation (architectures/__init__.py:12-19), fused into the front-end kernel."""
    forward = _fused_only


class UnSqueeze(nn.Module):
    forward = _fused_only


class _FusedMaxPool1d(nn.MaxPool1d):
    forward = _fused_only


class _FusedFlatten(nn.Flatten):
    forward = _fused_only


class _FusedReLU(nn.ReLU):
    forward = _fused_only


class _FusedDropout(nn.Dropout):
    forward = _fused_only


def _check_param(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} lives on {t.device}: move the model to a CUDA device "
                           "(explainn_b200 has no CPU fallback)")
    if t.dtype != torch.float32:
        raise TypeError(f"{name}: the B200 path computes in float32, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def _bn_momentum(bn):
    if bn.momentum is None:
        raise NotImplementedError("cumulative-average BatchNorm (momentum=None) is not supported")
    return bn.momentum


class FusedLinear(nn.Linear):
    """`final` (architectures/__init__.py:104): nn.Linear parameters, our kernels through the
    `explainn_b200::linear` custom op (autograd registered on the op: torch_ops.py)."""

    _arena = None      # (flat gradient arena, [offset of dW, offset of db]) for the duration of ExplaiNN.forward

    def forward(self, x):
        w = _check_param(self.weight, "final.weight")
        b = _check_param(self.bias, "final.bias")
        if x.dtype != torch.float32 or not x.is_cuda:
            raise RuntimeError("final(x): x must be a float32 CUDA tensor")
        x2 = x.contiguous().view(-1, w.shape[1])
        arena, offs = self._arena if (self._arena is not None and torch.is_grad_enabled()) else (None, [])
        y = torch.ops.explainn_b200.linear(x2, w, b, arena, offs)
        return y.view(*x.shape[:-1], w.shape[0])


class UnitBank(nn.Sequential):
    """`linears` of the reference (architectures/__init__.py:72-102) as one fused operator.

    Index / slice access keeps working (`linears[0].weight`, `linears[:3]`).  Supported
    call forms: the full bank -> unit outputs (B,U); `[:3]` -> activations (B,U,Lc);
    `[:4]` -> max-pooled activations (B,U,n).  Input is the float one-hot (B,4,L) or the
    reference's U-times repeated (B,4U,L) tensor (only its first 4 channels are read), or
    already packed planes wrapped in `PackedDNA`.
    """

    def _kind(self):
        keys = list(self._modules.keys())
        if keys == _FULL_KEYS:
            return "outputs"
        if keys == ["0", "1", "2"]:
            return "acts"
        if keys == ["0", "1", "2", "3"]:
            return "pool"
        raise NotImplementedError(
            f"slice {keys} of the fused unit bank is not callable; supported: full, [:3], [:4]")

    # What happens to a float input column that is neither one-hot nor all-zero (the packed layout cannot hold
    # it; the reference's Conv1d would compute on it):
    #   "strict": raise at once (one device->host read per call) -- the default;
    #   "lazy"  : count such columns in a device-side flag without a sync; `check_inputs()` raises later
    #             (StepPipeline reads the flag together with every loss it hands back).
    input_check = "strict"

    def _bad_flag(self, device):
        flags = self.__dict__.setdefault("_bad_flags", {})
        if device not in flags:
            flags[device] = torch.zeros(1, dtype=torch.int32, device=device)
        return flags[device]

    def check_inputs(self):
        """Raise if any float batch packed since the last check had non-one-hot columns ("lazy" mode)."""
        for dev, flag in self.__dict__.get("_bad_flags", {}).items():
            if int(flag.item()) != 0:
                flag.zero_()
                raise ValueError("input is not one-hot: a batch had columns that are neither one-hot nor all-zero "
                                 "(explainn_b200 computes on the packed 2-bit layout)")

    def _pack(self, x):
        if isinstance(x, PackedDNA):
            return x.planes, x.length
        if not isinstance(x, torch.Tensor) or x.dim() != 3:
            raise ValueError("expected a (B, 4, L) or (B, 4*cnn_units, L) float32 tensor")
        if not x.is_cuda:
            raise RuntimeError("explainn_b200 runs on CUDA tensors only; there is no CPU fallback")
        U = self[0].out_channels
        if x.shape[1] not in (4, 4 * U):
            raise ValueError(f"expected 4 or {4 * U} input channels, got {x.shape[1]}")
        if self.input_check == "strict" and not torch.cuda.is_current_stream_capturing():
            return ops.pack_onehot(x, check=True), x.shape[2]
        return torch.ops.explainn_b200.pack_onehot(x, self._bad_flag(x.device)), x.shape[2]

    def _check_kernel_size(self, training_path):
        k = self[0].kernel_size[0]
        lim = ops.kernel_size_limit(1 if training_path else 0)
        if k > lim:
            what = ("training (BatchNorm batch statistics, backward) and linears[:3] activations"
                    if training_path else "the fused front end")
            raise NotImplementedError(f"kernel_size {k}: {what} support kernel_size <= {lim} "
                                      f"(eval forward / interpret / PWM scans: <= {ops.kernel_size_limit(0)})")

    def _front_affine(self, planes, L):
        """(scale, shift) of BN1 folded with the conv bias; batch statistics in train mode."""
        conv, bn1 = self[0], self[1]
        cw = _check_param(conv.weight, "linears.0.weight")
        cb = _check_param(conv.bias, "linears.0.bias")
        g1 = _check_param(bn1.weight, "linears.1.weight")
        be1 = _check_param(bn1.bias, "linears.1.bias")
        k = cw.shape[2]
        if bn1.training:
            cnt1, cnt2 = ops.cooc_counts(planes, L, k)
            n_pos = planes.shape[0] * (L - k + 1)
            scale, shift, _, _ = ops.bn1_train_prepare(cnt1, cnt2, n_pos, cw, cb, g1, be1, bn1.eps,
                                                       _bn_momentum(bn1), bn1.running_mean, bn1.running_var)
            ops.counters_advance_list([bn1.num_batches_tracked])
        else:
            scale, shift = ops.bn_fold_eval(g1, be1, bn1.running_mean, bn1.running_var, cb, bn1.eps)
        return cw, scale, shift

    _arena = None      # (flat gradient arena, offsets of the 12 parameter gradients) during ExplaiNN.forward
    _draw = None       # device-resident dropout draw counter (int64 (1,)), used while a CUDA graph is captured

    def _dropout_seed(self, p, dropout_mask, device):
        """(host seed, device draw counter).  Eager: a fresh seed per call from torch's CPU generator (so
        torch.manual_seed reproduces a run).  Under CUDA-graph capture a host seed would be baked into the graph:
        the kernels then also mix in a device counter that the graph itself advances on every replay."""
        if not (p > 0.0 and dropout_mask is None):
            return 0, None
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        if not torch.cuda.is_current_stream_capturing():
            return seed, None
        if self._draw is None or self._draw.device != device:
            raise RuntimeError("capturing a training step with dropout: call linears.enable_device_draws() first "
                               "(the draw counter must exist before the capture starts)")
        ops.counters_advance(self._draw)
        return seed, self._draw

    def enable_device_draws(self, device):
        if self._draw is None or self._draw.device != torch.device(device):
            self.__dict__["_draw"] = torch.zeros(1, dtype=torch.int64, device=device)

    def forward(self, x, dropout_mask=None):
        kind = self._kind()
        planes, L = self._pack(x)
        conv = self[0]
        if L < conv.kernel_size[0]:
            raise ValueError("sequence shorter than the kernel")
        self._check_kernel_size(training_path=(kind == "acts" or self[1].training))
        if kind in ("acts", "pool"):
            with torch.no_grad():
                cw, scale, shift = self._front_affine(planes, L)
                op = torch.ops.explainn_b200.frontend_acts if kind == "acts" else torch.ops.explainn_b200.frontend_pool
                return op(planes, L, cw.detach(), scale, shift)
        bn1, fc1, bn2, drop, fc2, bn3 = self[1], self[6], self[7], self[9], self[10], self[11]
        names = ["linears.0.weight", "linears.0.bias", "linears.1.weight", "linears.1.bias", "linears.6.weight",
                 "linears.6.bias", "linears.7.weight", "linears.7.bias", "linears.10.weight", "linears.10.bias",
                 "linears.11.weight", "linears.11.bias"]
        params = [conv.weight, conv.bias, bn1.weight, bn1.bias, fc1.weight, fc1.bias, bn2.weight, bn2.bias,
                  fc2.weight, fc2.bias, bn3.weight, bn3.bias]
        params = [_check_param(p, nme) for p, nme in zip(params, names)]
        # the kernels address W1 as (100U, n) and W2 as (U, 100): same memory as the reference's (.., 1) shapes
        modes = (bn1.training, bn2.training, bn3.training)
        need_grad = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        eps = [bn1.eps, bn2.eps, bn3.eps]
        p = float(drop.p) if drop.training else 0.0
        if dropout_mask is not None and dropout_mask.dtype != torch.uint8:
            raise TypeError("dropout_mask: expected a uint8 (B, cnn_units, 100) tensor")
        if all(modes):
            if planes.shape[0] < 2:
                raise ValueError("Expected more than 1 value per channel when training")
            seed, seed_dev = self._dropout_seed(p, dropout_mask, planes.device)
            arena, offs = self._arena if (need_grad and self._arena is not None) else (None, [])
            outs = torch.ops.explainn_b200.unit_bank_train(planes, L, params, eps, p, seed, seed_dev, dropout_mask,
                                                           arena, offs)
            stats1, stats2, stats3 = outs[8], outs[9], outs[10]
            B = planes.shape[0]
            world = _world_for_stats()
            counts = [float(B * (L - conv.kernel_size[0] + 1) * world), float(B * world), float(B * world)]
            with torch.no_grad():
                bns = [b_ for b_ in (bn1, bn2, bn3)]
                if all(b_.running_mean is not None and b_.running_var is not None for b_ in bns):
                    nbt = [b_.num_batches_tracked for b_ in bns]
                    torch.ops.explainn_b200.bn_running_update(
                        [stats1, stats2, stats3], [b_.running_mean for b_ in bns], [b_.running_var for b_ in bns],
                        nbt if all(t is not None for t in nbt) else [], [_bn_momentum(b_) for b_ in bns], counts)
            return outs[0]
        if need_grad:
            raise NotImplementedError("backward through eval-mode BatchNorm is not implemented: "
                                      "call model.train() for training steps or wrap inference in torch.no_grad()")
        if any(modes):
            raise NotImplementedError("mixed train/eval BatchNorm modes inside the unit bank are not supported")
        with torch.no_grad():
            seed, _ = self._dropout_seed(p, dropout_mask, planes.device)
            running = [bn1.running_mean, bn1.running_var, bn2.running_mean, bn2.running_var, bn3.running_mean,
                       bn3.running_var]
            return torch.ops.explainn_b200.unit_bank_eval(planes, L, [q.detach() for q in params], running, eps, p,
                                                          seed, dropout_mask)


def _world_for_stats():
    from . import torch_ops
    return torch_ops._sync["world"] if torch_ops._sync["sum"] else 1


class PackedDNA:
    """A batch already in the packed `planes` layout (include/explainn_b200.h) plus its length."""

    def __init__(self, planes, length):
        self.planes = planes
        self.length = int(length)

    @classmethod
    def from_onehot(cls, x):
        return cls(ops.pack_onehot(x), x.shape[2])

    def revcomp(self):
        return PackedDNA(ops.revcomp_planes(self.planes, self.length), self.length)

    def __len__(self):
        return self.planes.shape[0]


class _Model(nn.Module):

    def load_weights(self, weight_file):
        """Legacy loader (architectures/__init__.py:27-39): values are matched to keys by
        position and a trailing size-1 dimension is squeezed."""
        sd = torch.load(weight_file)
        keys = list(self.state_dict().keys())
        remapped = OrderedDict()
        for key, v in zip(keys, sd.values()):
            if v.dim() > 1 and v.shape[-1] == 1:
                v = v.squeeze(-1)
            remapped[key] = v
        self.load_state_dict(remapped)


class ExplaiNN(_Model):
    """ExplaiNN: explainable neural networks -- B200-native drop-in (architectures/__init__.py:41-114)."""

    def __init__(self, cnn_units, kernel_size, sequence_length, n_features=1, weights_file=None):
        super().__init__()
        self._options = {
            "cnn_units": cnn_units,
            "kernel_size": kernel_size,
            "sequence_length": sequence_length,
            "n_features": n_features,
            "weights_file": weights_file,
        }
        if kernel_size > ops.KERNEL_SIZE_MAX:
            raise NotImplementedError(f"kernel_size {kernel_size}: the fused front end supports kernel_size <= "
                                      f"{ops.KERNEL_SIZE_MAX} ({ops.KERNEL_SIZE_MAX_TRAIN} for training)")
        n = math.floor((sequence_length - kernel_size + 1) / 7.)
        self._channels_after_maxpool = n
        # same construction order as the reference, so the same seed gives the same initial weights
        self.linears = UnitBank(
            FusedConv1d(in_channels=4 * cnn_units, out_channels=cnn_units, kernel_size=kernel_size,
                        groups=cnn_units),
            FusedBatchNorm1d(cnn_units),
            ExpAct(),
            _FusedMaxPool1d(7, 7),
            _FusedFlatten(),
            UnSqueeze(),
            FusedConv1d(in_channels=n * cnn_units, out_channels=100 * cnn_units, kernel_size=1, groups=cnn_units),
            FusedBatchNorm1d(100 * cnn_units, 1e-05, 0.1, True),
            _FusedReLU(),
            _FusedDropout(0.3),
            FusedConv1d(in_channels=100 * cnn_units, out_channels=cnn_units, kernel_size=1, groups=cnn_units),
            FusedBatchNorm1d(cnn_units, 1e-05, 0.1, True),
            _FusedReLU(),
            _FusedFlatten(),
        )
        self.final = FusedLinear(cnn_units, n_features)
        if weights_file is not None:
            self.load_weights(weights_file)

    def forward(self, x, dropout_mask=None):
        """x: float32 one-hot (B,4,L) on CUDA (or PackedDNA) -> logits (B,F).  The reference's
        `x.repeat(1, cnn_units, 1)` (:111) is never materialised."""
        arena = None
        if torch.is_grad_enabled() and self.training:
            arena = self._grad_arena()
        if arena is None:
            return self.final(self.linears(x, dropout_mask=dropout_mask))
        self.linears._arena = (arena, self._arena_offs[2:])
        self.final._arena = (arena, self._arena_offs[:2])
        try:
            return self.final(self.linears(x, dropout_mask=dropout_mask))
        finally:
            self.linears._arena = None
            self.final._arena = None

    # ---- flat gradient arena -------------------------------------------------------------------------------
    # Layout (element offsets, each a multiple of 64 = 256 bytes):
    #   [ final.weight, final.bias | conv_w, conv_b, g1, be1, w1, b1, g2, be2, w2, b2, g3, be3 ]
    # with the conv / BN1 gradients -- the last ones the backward produces -- stored LAST in memory order
    # (`arena_split`): everything before the split is complete when the MLP backward finishes and is
    # all-reduced while the conv weight gradient is still being computed (parallel.GradBucket).
    ARENA_ALIGN = 64

    def _arena_plan(self):
        names = ["final.weight", "final.bias", "linears.0.weight", "linears.0.bias", "linears.1.weight",
                 "linears.1.bias", "linears.6.weight", "linears.6.bias", "linears.7.weight", "linears.7.bias",
                 "linears.10.weight", "linears.10.bias", "linears.11.weight", "linears.11.bias"]
        pd = dict(self.named_parameters())
        order = [0, 1, 12, 13, 10, 11, 8, 9, 6, 7, 2, 3, 4, 5]     # memory order: MLP segment first, conv/BN1 last
        offs, off, split = [0] * len(names), 0, 0
        for i in order:
            if i == 2:
                split = off
            offs[i] = off
            off += -(-pd[names[i]].numel() // self.ARENA_ALIGN) * self.ARENA_ALIGN
        return names, offs, off, split

    def _grad_arena(self):
        """the arena of THIS forward/backward: the persistent one (use_grad_arena) or a fresh allocation"""
        names, offs, total, split = self._arena_plan()
        self.__dict__["_arena_offs"] = offs
        self.__dict__["arena_split"] = split
        dev = self.final.weight.device
        buf = self.__dict__.get("_arena_buf")
        if buf is not None:
            if buf.numel() != total or buf.device != dev:
                raise RuntimeError("the persistent gradient arena no longer matches the parameters: call "
                                   "use_grad_arena() again after changing / moving the model")
            return buf
        return torch.zeros(total, dtype=torch.float32, device=dev) if self.__dict__.get("_arena_zero") else \
            torch.empty(total, dtype=torch.float32, device=dev)

    def use_grad_arena(self, persistent=True):
        """Keep ONE flat gradient buffer across steps (data-parallel buckets and CUDA graphs need stable addresses).
        p.grad of every parameter then aliases it: a gradient kept from an earlier step is overwritten by the next
        backward -- clone it if it must survive.  Returns the buffer."""
        if not persistent:
            self.__dict__.pop("_arena_buf", None)
            return None
        names, offs, total, split = self._arena_plan()
        self.__dict__["_arena_buf"] = torch.zeros(total, dtype=torch.float32, device=self.final.weight.device)
        self.__dict__["_arena_offs"] = offs
        self.__dict__["arena_split"] = split
        return self.__dict__["_arena_buf"]

    def check_inputs(self):
        """see UnitBank.input_check"""
        self.linears.check_inputs()


class PWM(nn.Module):
    """PWM scorer (architectures/__init__.py:116-170): log-odds correlation on both strands,
    per-sequence max or sum, as one fused scan kernel."""

    def __init__(self, pwms, sequence_length, scoring="sum"):
        super().__init__()
        groups, _, kernel_size = pwms.shape
        self._options = {
            "groups": groups,
            "kernel_size": kernel_size,
            "sequence_length": sequence_length,
            "scoring": scoring,
        }
        self.conv1d = FusedConv1d(in_channels=4 * groups, out_channels=groups, kernel_size=kernel_size,
                                  groups=groups)
        self.conv1d.bias.data = torch.zeros(groups)
        self.conv1d.weight.data = torch.as_tensor(pwms, dtype=torch.float32).clone()
        for p in self.conv1d.parameters():
            p.requires_grad = False

    def forward(self, x, return_argmax=False):
        if isinstance(x, PackedDNA):
            planes, L = x.planes, x.length
        else:
            if not x.is_cuda:
                raise RuntimeError("explainn_b200 runs on CUDA tensors only; there is no CPU fallback")
            planes, L = ops.pack_onehot(x), x.shape[2]
        w = _check_param(self.conv1d.weight, "conv1d.weight")
        mode = "max" if self._options["scoring"] == "max" else "sum"
        res = torch.ops.explainn_b200.pwm_scan(planes, L, w.detach(), mode, bool(return_argmax and mode == "max"))
        return (res[0], res[1] if len(res) > 1 else None) if return_argmax else res[0]


class _FusedLossMixin:
    """nn.MSELoss / nn.BCEWithLogitsLoss with the default arguments, computed by the `explainn_b200::loss` custom op
    (one reduction kernel forward, one elementwise kernel backward; no library kernels on the step path)."""
    _kind = None

    def forward(self, input, target):
        if self.reduction != "mean" or getattr(self, "weight", None) is not None or \
                getattr(self, "pos_weight", None) is not None:
            raise NotImplementedError("the fused loss implements the reference's configuration only: 'mean' "
                                      "reduction, no weights (architectures/__init__.py:446-456)")
        if not (input.is_cuda and target.is_cuda):
            raise RuntimeError("explainn_b200 losses run on CUDA tensors only; there is no CPU fallback")
        if input.dtype != torch.float32 or target.dtype != torch.float32:
            raise TypeError("the fused loss computes in float32")
        if input.shape != target.shape:
            raise ValueError(f"input {tuple(input.shape)} and target {tuple(target.shape)} must have the same shape")
        return torch.ops.explainn_b200.loss(input.contiguous(), target.contiguous(), self._kind)


class FusedMSELoss(_FusedLossMixin, nn.MSELoss):
    _kind = "mse"


class FusedBCEWithLogitsLoss(_FusedLossMixin, nn.BCEWithLogitsLoss):
    _kind = "bce_logits"


def pwm_best_sites(pwm_module, x):
    """Best site per (sequence, PWM) with the strand rule of the reference's cudamo.py: both strands are scored by the
    same forward correlation (`Conv1d` + `MaxPool1d(return_indices=True)`, cudamo.py:217-228 -- the FIRST maximum of a
    strand wins), and the forward strand is chosen only if its score is STRICTLY greater (`fwd[0] > rev[0]`,
    cudamo.py:254-265); on a tie the reverse strand wins.  Returns (score (B,G) float32, strand (B,G) int32 with
    0 = forward / 1 = reverse complement, start (B,G) int32 = position on the chosen strand's own string).

    One scan of the REVERSE-COMPLEMENTED batch gives exactly this: its concatenated axis is [rc strand | forward
    strand], the scan keeps the first maximum of that axis, so an index below L-w+1 means "rc strand >= forward
    strand, first rc maximum" and anything else "forward strictly greater, first forward maximum"."""
    if pwm_module._options["scoring"] != "max":
        raise ValueError("best sites need a PWM module with scoring='max'")
    packed = x if isinstance(x, PackedDNA) else PackedDNA.from_onehot(x)
    score, arg = pwm_module(packed.revcomp(), return_argmax=True)
    Lc = packed.length - pwm_module._options["kernel_size"] + 1
    strand = (arg < Lc).to(torch.int32)
    return score, strand, torch.where(arg < Lc, arg, arg - Lc)


def get_loss(input_data="binary"):
    """architectures/__init__.py:446-456: nn.BCEWithLogitsLoss() for "binary", nn.MSELoss() otherwise (subclasses
    of those, so isinstance checks and attributes hold)"""
    return FusedBCEWithLogitsLoss() if input_data == "binary" else FusedMSELoss()


def get_metrics(input_data="binary"):
    """architectures/__init__.py:458-461: the same keys and `(y_true, y_score)` call signature; the functions are
    the device versions of explainn_b200/metrics.py (csrc/metrics.cu), not sklearn / scipy on the host."""
    from . import metrics
    return metrics.get_metrics(input_data)


def get_optimizer(params, lr=1e-03):
    """architectures/__init__.py:463-464: Adam with torch defaults; the update runs in the fused
    xnn_adam_step kernel and the state layout stays torch.optim.Adam's."""
    from .optim import FusedAdam
    return FusedAdam(params, lr=lr)
