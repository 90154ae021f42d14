"""ResNet-18/34/50 with torchvision-identical parameter / buffer names, running on the
hand-written sm_100a kernels.

Parity target: ``torchvision.models.resnet18(num_classes=10)`` as built by the reference
(train_ddp.py:153-156; torchvision/models/resnet.py:59-103 BasicBlock, :108-163 Bottleneck,
:166-285 ResNet) — same module tree, same ``state_dict`` keys and shapes, same initialisation
(kaiming_normal fan_out for convs, BN gamma=1 beta=0, default Linear init), stock ImageNet stem
(7x7/2 conv + 3x3/2 max-pool) even on 32x32 inputs.

Two execution paths share the parameters:
  * native (CUDA + bf16): ``ResNetEngine`` — a static plan of kernel launches over preallocated
    NHWC bf16 activations.  conv = tcgen05 implicit GEMM whose epilogue also produces the
    BatchNorm batch statistics; BN+ReLU(+residual) one elementwise kernel; backward is
    hand-scheduled (fused residual-sum + ReLU-mask + BN backward, dgrad, wgrad) and writes
    gradients straight into the flat gradient arena, notifying the gradient reducer bucket by
    bucket.  The whole thing is CUDA-graph capturable (no allocation, no host sync).
  * torch ops (CPU, fp32/TF32, fp16 autocast): the same math through ATen, used for the gloo
    plumbing configuration and as the in-repo baseline.
"""
from __future__ import annotations

import math
import os
from typing import Callable, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


# ===================================================================================== modules
class Conv2d(nn.Module):
    """Bias-free convolution parameter holder.  ``weight`` is logically ``[Cout, Cin, R, S]``
    (state_dict compatible) but stored channels_last, i.e. physically KRSC, which is the layout
    the implicit-GEMM kernels read."""

    def __init__(self, cin: int, cout: int, kernel_size: int, stride: int = 1, padding: int = 0):
        super().__init__()
        self.in_channels, self.out_channels = cin, cout
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding
        w = torch.empty(cout, cin, kernel_size, kernel_size)
        nn.init.kaiming_normal_(w, mode="fan_out", nonlinearity="relu")
        self.weight = nn.Parameter(w.contiguous(memory_format=torch.channels_last))

    def forward(self, x):
        return F.conv2d(x, self.weight, None, self.stride, self.padding)

    def extra_repr(self) -> str:
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, "
                f"stride={self.stride}, padding={self.padding}, bias=False")


class BatchNorm2d(nn.Module):
    def __init__(self, c: int, eps: float = 1e-5, momentum: float = 0.1):
        super().__init__()
        self.num_features, self.eps, self.momentum = c, eps, momentum
        self.weight = nn.Parameter(torch.ones(c))
        self.bias = nn.Parameter(torch.zeros(c))
        self.register_buffer("running_mean", torch.zeros(c))
        self.register_buffer("running_var", torch.ones(c))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))

    def forward(self, x):
        if self.training:
            self.num_batches_tracked.add_(1)
        return F.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias,
                            self.training, self.momentum, self.eps)

    def extra_repr(self) -> str:
        return f"{self.num_features}, eps={self.eps}, momentum={self.momentum}"


class Linear(nn.Module):
    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.in_features, self.out_features = cin, cout
        self.weight = nn.Parameter(torch.empty(cout, cin))
        self.bias = nn.Parameter(torch.empty(cout))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        bound = 1 / math.sqrt(cin)
        nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x):
        return F.linear(x, self.weight, self.bias)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample=None):
        super().__init__()
        self.conv1 = Conv2d(inplanes, planes, 3, stride, 1)
        self.bn1 = BatchNorm2d(planes)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = Conv2d(planes, planes, 3, 1, 1)
        self.bn2 = BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample=None):
        super().__init__()
        self.conv1 = Conv2d(inplanes, planes, 1)
        self.bn1 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, 3, stride, 1)
        self.bn2 = BatchNorm2d(planes)
        self.conv3 = Conv2d(planes, planes * 4, 1)
        self.bn3 = BatchNorm2d(planes * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes: int = 1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = Conv2d(3, 64, 7, 2, 3)
        self.bn1 = BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = Linear(512 * block.expansion, num_classes)
        self.num_classes = num_classes
        self.native = "auto"      # "auto" | "always" | "never"
        self.native_fp32 = True   # without autocast: native fp32/TF32 kernels (False -> ATen ops)
        self._engines: dict = {}
        self._grad_sink: Optional[Callable] = None  # set by the DDP wrapper (bucket-ready hook)
        self._last_engine = None

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(Conv2d(self.inplanes, planes * block.expansion, 1, stride),
                                       BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    # ------------------------------------------------------------------ dispatch
    def _use_native(self, x) -> bool:
        if self.native == "never" or not x.is_cuda:
            return False
        if self.native == "always":
            return True
        if torch.is_autocast_enabled("cuda"):
            return torch.get_autocast_dtype("cuda") in (torch.bfloat16, torch.float16)
        return self.native_fp32          # no autocast: fp32 activations, TF32 tensor-core math

    def _native_dtype(self):
        if torch.is_autocast_enabled("cuda"):
            dt = torch.get_autocast_dtype("cuda")
            if dt in (torch.float16, torch.bfloat16):
                return dt
        if self.native == "always":
            return getattr(self, "native_dtype", torch.bfloat16)
        return torch.float32

    def engine_for(self, shape, training: bool, dtype=None):
        dtype = dtype or self._native_dtype()
        key = (tuple(shape), bool(training), dtype)
        eng = self._engines.get(key)
        if eng is None:
            eng = ResNetEngine(self, shape, training, dtype)
            self._engines[key] = eng
        return eng

    def forward(self, x):
        if self._use_native(x):
            eng = self.engine_for(x.shape, self.training)
            self._last_engine = eng
            if self.training and torch.is_grad_enabled():
                # autograd needs one differentiable input to call backward; a private dummy is
                # used (NOT a parameter: post-accumulate-grad hooks fire even for None
                # gradients, which would report a parameter to the reducer twice)
                anchor = self.__dict__.get("_anchor")
                if anchor is None or anchor.device != x.device:
                    anchor = torch.zeros((), device=x.device, requires_grad=True)
                    self.__dict__["_anchor"] = anchor
                return _EngineFn.apply(x, anchor, eng)
            return eng.forward(x)
        return self._forward_torch(x)

    def _forward_torch(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x)


def resnet18(num_classes: int = 1000) -> ResNet:
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)


def resnet34(num_classes: int = 1000) -> ResNet:
    return ResNet(BasicBlock, [3, 4, 6, 3], num_classes)


def resnet50(num_classes: int = 1000) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes)


def resnet101(num_classes: int = 1000) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 23, 3], num_classes)


def resnet152(num_classes: int = 1000) -> ResNet:
    return ResNet(Bottleneck, [3, 8, 36, 3], num_classes)


ARCHS = {"resnet18": resnet18, "resnet34": resnet34, "resnet50": resnet50, "resnet101": resnet101,
         "resnet152": resnet152}


def build_model(device, arch: str = "resnet18", num_classes: int = 10) -> ResNet:
    """Reference train_ddp.py:153-156 (``resnet18(num_classes=10).to(device)``).  The other
    torchvision depths share the block types and therefore the kernels; 18 and 50 are the
    benchmarked ones (BASELINE.json configs 2 and 4)."""
    return ARCHS[arch](num_classes).to(device)


# ===================================================================================== engine
class _EngineFn(torch.autograd.Function):
    """Bridges the hand-scheduled engine into autograd: forward returns the logits, backward
    runs the engine's backward pass, which deposits every parameter gradient directly in the
    gradient arena (hence no gradients are returned to autograd)."""

    @staticmethod
    def forward(ctx, x, anchor, eng):
        ctx.eng = eng
        return eng.forward(x).detach()

    @staticmethod
    def backward(ctx, dlogits):
        ctx.eng.backward(dlogits)
        return None, None, None


class _Unit:
    """conv -> BN (-> +residual) (-> ReLU) with its buffers."""
    __slots__ = ("conv", "bn", "x_shape", "y_shape", "raw", "y", "sum", "sq", "mean", "invstd",
                 "d_raw", "g", "relu", "name", "fsplit", "raw32", "dsplit", "bscratch", "bcounter")


def _live_taps(H, W, P, Q, k, stride, pad):
    """Filter taps that touch at least one real (non-padding) pixel."""
    n = 0
    for r in range(k):
        for s_ in range(k):
            hmin, hmax = r - pad, (P - 1) * stride + r - pad
            wmin, wmax = s_ - pad, (Q - 1) * stride + s_ - pad
            if hmax < 0 or hmin >= H or wmax < 0 or wmin >= W:
                continue
            n += 1
    return n


def _pick_splits(tiles: int, ksteps: int, sm_target: int = 0) -> int:
    """Split the k loop over CTAs when the tile grid alone would leave most SMs idle: a single
    CTA can only pull ~50 GB/s of operands through TMA, so small-M layers are spread out."""
    import os
    if sm_target <= 0:
        sm_target = int(os.environ.get("DPT_SPLITK_TARGET", "64"))
    if tiles > 32 or ksteps < 4 or os.environ.get("DPT_NO_SPLITK", "0") == "1":
        return 1
    s = min(ksteps, max(1, sm_target // tiles))
    per = -(-ksteps // s)
    return max(1, -(-ksteps // per))


class ResNetEngine:
    STEM_LDK = 152  # 7*7*3 = 147 padded so that rows are 16-byte aligned for TMA

    def __init__(self, model: ResNet, in_shape, training: bool, dtype=torch.bfloat16) -> None:
        self.dtype = dtype
        from ..ops import _lib
        from ..ops import functional as Fn
        self.Fn = Fn
        self._lib = _lib
        self.m = model
        self.training = training
        self.dev = model.fc.weight.device
        N, C, H, W = in_shape
        assert C == 3, "stem expects RGB input"
        self.N = N
        self.in_shape = tuple(in_shape)
        bf = dict(device=self.dev, dtype=self.dtype)
        f32 = dict(device=self.dev, dtype=torch.float32)
        self.units: list[_Unit] = []
        self._zero32_plan: list = []
        self._zero32_n = 0
        bns = [mod for mod in model.modules() if isinstance(mod, BatchNorm2d)]
        total_c = sum(b.num_features for b in bns)
        # one arena for everything that must be zero at the start of a step (forward batch
        # statistics, backward per-channel sums, grid-barrier counters): a single memset
        self.stats = torch.zeros(4 * total_c + len(bns) + 8, **f32)
        self._bwd_off = 2 * total_c
        self._cnt_off = 4 * total_c
        self.saved = torch.zeros(2 * total_c, **f32)
        self._stat_off = 0

        # ---- stem ----------------------------------------------------------------------
        P, Q = Fn.conv_out_hw(H, W, 7, 7, 2, 3)
        self.stem_hw = (P, Q)
        self.cols = torch.empty((N * P * Q, self.STEM_LDK), **bf)
        self.stem_w = torch.zeros((64, self.STEM_LDK), **bf)
        self.stem = self._unit("conv1", model.conv1, model.bn1, (N, H, W, 3), (N, P, Q, 64), True)
        PP, QQ = Fn.conv_out_hw(P, Q, 3, 3, 2, 1)
        self.pool_y = torch.empty((N, PP, QQ, 64), **bf)
        self.pool_idx = torch.empty((N, PP, QQ, 64), device=self.dev, dtype=torch.uint8)
        self.d_pool_in = torch.empty((N, P, Q, 64), **bf) if training else None

        # ---- residual stages -------------------------------------------------------------
        self.blocks = []
        shape = (N, PP, QQ, 64)
        for lname in ("layer1", "layer2", "layer3", "layer4"):
            for bi, blk in enumerate(getattr(model, lname)):
                shape = self._plan_block(f"{lname}.{bi}", blk, shape)
        self.feat_shape = shape
        K = model.fc.out_features
        self.pooled = torch.empty((N, shape[3]), **f32)
        self.logits = torch.empty((N, K), **f32)
        self.d_feat = torch.empty(shape, **bf) if training else None
        self._last_input = None
        # materialise the fp32 split-K accumulators
        self.zero32 = torch.zeros(max(self._zero32_n, 1), **f32) if training else None

        def z32(handle):
            off, n, shp = self._zero32_plan[handle]
            return self.zero32[off:off + n].view(shp)
        if training:
            for u in self.units:
                if u.raw32 is not None:
                    u.raw32 = z32(u.raw32)
            for b in self.blocks:
                if isinstance(b.get("dx"), int):
                    b["dx"] = z32(b["dx"])
                b["d_mid"] = [z32(t) if isinstance(t, int) else t for t in b.get("d_mid", [])]
        # weight-gradient kernels are off the critical path (nothing in backward consumes
        # them): they run on a second stream, concurrently with the dgrad / BN-backward chain
        self.side = torch.cuda.Stream(device=self.dev) if training else None
        # the 1x1 downsample branch (forward conv+BN, backward BN) is independent of the main
        # branch of its block: it runs on a third stream and joins where its result is consumed
        self.side2 = torch.cuda.Stream(device=self.dev) if training else None
        self._ev2 = [torch.cuda.Event() for _ in range(32)] if training else []
        self._ev2_i = 0
        self._ev_pool = [torch.cuda.Event() for _ in range(64)] if training else []
        self._ev_i = 0
        import os
        self._fuse_bn_bwd = os.environ.get("DPT_FUSE_BN_BWD", "0") == "1"
        self._branch_streams = os.environ.get("DPT_BRANCH_STREAMS", "1") == "1"
        self._z32_ev = torch.cuda.Event() if training else None
        self._z32_pending = False

    # ------------------------------------------------------------------ planning
    def _unit(self, name, conv, bn, x_shape, y_shape, relu) -> _Unit:
        u = _Unit()
        u.name, u.conv, u.bn, u.x_shape, u.y_shape, u.relu = name, conv, bn, x_shape, y_shape, relu
        bf = dict(device=self.dev, dtype=self.dtype)
        C = y_shape[3]
        u.raw = torch.empty(y_shape, **bf)
        u.y = torch.empty(y_shape, **bf)
        o = self._stat_off
        u.sum, u.sq = self.stats[o:o + C], self.stats[o + C:o + 2 * C]
        u.mean, u.invstd = self.saved[o:o + C], self.saved[o + C:o + 2 * C]
        self._stat_off += 2 * C
        u.d_raw = torch.empty(y_shape, **bf) if self.training else None
        u.g = torch.empty(y_shape, **bf) if self.training else None
        u.fsplit, u.raw32, u.dsplit = 1, None, 1
        bo = self._bwd_off + o
        u.bscratch = self.stats[bo:bo + 2 * C]
        ci = self._cnt_off + len(self.units)
        u.bcounter = self.stats[ci:ci + 1].view(torch.int32)
        if self.training and conv.in_channels % 8 == 0:
            N, H, W, Cin = x_shape
            _, P, Q, Cout = y_shape
            k, st, pd = conv.kernel_size, conv.stride, conv.padding
            bn_ = 128 if Cout >= 128 else 64
            tiles = -(-(N * P * Q) // 128) * -(-Cout // bn_)
            ksteps = _live_taps(H, W, P, Q, k, st, pd) * -(-Cin // 64)
            u.fsplit = _pick_splits(tiles, ksteps)
            if u.fsplit > 1:
                u.raw32 = self._zero32_alloc(y_shape)
            # dgrad: rows are input pixels of one stride-parity class, k runs over (taps, Cout)
            bn_d = 128 if Cin >= 128 else 64
            rows = N * -(-H // st) * -(-W // st)
            dtiles = -(-rows // 128) * -(-Cin // bn_d) * (st * st)
            dk = max(1, _live_taps(H, W, P, Q, k, st, pd) // (st * st)) * -(-Cout // 64)
            u.dsplit = _pick_splits(dtiles, dk)
        self.units.append(u)
        return u

    def _zero32_alloc(self, shape):
        """fp32 scratch that must be zero at the start of every step (split-K accumulators):
        carved from one flat buffer so a single memset clears all of it."""
        n = 1
        for d in shape:
            n *= d
        self._zero32_plan.append((self._zero32_n, n, tuple(shape)))
        self._zero32_n += (n + 63) // 64 * 64
        return len(self._zero32_plan) - 1

    def _conv_unit(self, name, conv, bn, x_shape, relu) -> _Unit:
        N, H, W, _ = x_shape
        P, Q = self.Fn.conv_out_hw(H, W, conv.kernel_size, conv.kernel_size, conv.stride,
                                   conv.padding)
        return self._unit(name, conv, bn, x_shape, (N, P, Q, conv.out_channels), relu)

    def _plan_block(self, name, blk, x_shape):
        bf = dict(device=self.dev, dtype=self.dtype)
        b = {"name": name, "x_shape": x_shape, "kind": "basic" if isinstance(blk, BasicBlock) else "bottle"}
        if b["kind"] == "basic":
            u1 = self._conv_unit(f"{name}.conv1", blk.conv1, blk.bn1, x_shape, True)
            u2 = self._conv_unit(f"{name}.conv2", blk.conv2, blk.bn2, u1.y_shape, True)
            b["units"] = [u1, u2]
        else:
            u1 = self._conv_unit(f"{name}.conv1", blk.conv1, blk.bn1, x_shape, True)
            u2 = self._conv_unit(f"{name}.conv2", blk.conv2, blk.bn2, u1.y_shape, True)
            u3 = self._conv_unit(f"{name}.conv3", blk.conv3, blk.bn3, u2.y_shape, True)
            b["units"] = [u1, u2, u3]
        last = b["units"][-1]
        b["ds"] = None
        if blk.downsample is not None:
            b["ds"] = self._conv_unit(f"{name}.downsample", blk.downsample[0], blk.downsample[1],
                                      x_shape, False)
        if self.training:
            u1 = b["units"][0]
            first_block = len(self.blocks) == 0     # its dx feeds max-pool backward (bf16 only)
            if u1.dsplit > 1 and not first_block:
                b["dx"] = self._zero32_alloc(x_shape)           # fp32 split-K accumulator
            else:
                u1.dsplit = 1
                b["dx"] = torch.empty(x_shape, **bf)           # grad wrt block input (main path)
            fused = self._can_fuse_dgrad(u1, b["ds"])
            b["dx_ds"] = torch.empty(x_shape, **bf) if (b["ds"] is not None and not fused) else None
            if b["ds"] is not None:
                b["ds"].dsplit = 1
            b["d_mid"] = []
            for i, u in enumerate(b["units"][:-1]):
                nxt = b["units"][i + 1]                         # its dgrad produces grad wrt u.y
                if nxt.dsplit > 1:
                    b["d_mid"].append(self._zero32_alloc(u.y_shape))
                else:
                    b["d_mid"].append(torch.empty(u.y_shape, **bf))
        self.blocks.append(b)
        return last.y_shape

    # ------------------------------------------------------------------ parameter access
    def _w(self, conv: Conv2d):
        """bf16 KRSC weight for the tensor-core kernels (shadow arena slice when available)."""
        p = conv.weight
        arena = getattr(p, "_dpt_arena", None)
        Cout, Cin, R, S = p.shape
        if self.dtype == torch.float32:
            # FP32 mode: the tensor cores read the fp32 masters directly (as TF32), no shadow copy
            if arena is not None:
                return arena.flat_param_view(p).view(Cout, R, S, Cin)
            w = p.detach()
            if w.dim() == 4 and not w.is_contiguous():
                return w.permute(0, 2, 3, 1)          # channels_last storage is already KRSC
            return w.permute(0, 2, 3, 1).contiguous() if (R > 1 or S > 1) else w.view(Cout, R, S, Cin)
        if arena is not None and arena.shadow is not None:
            if arena.shadow.dtype != self.dtype:
                arena.set_shadow_dtype(self.dtype)
            return arena.shadow_view(p).view(Cout, R, S, Cin)
        cache = getattr(self, "_wcache", None)
        if cache is None:
            cache = self._wcache = {}
        t = cache.get(id(p))
        if t is None:
            t = cache[id(p)] = torch.empty((Cout, R, S, Cin), device=self.dev, dtype=self.dtype)
        t.copy_(p.detach().permute(0, 2, 3, 1))
        return t

    def _gw(self, p):
        """fp32 gradient destination of ``p`` in physical layout (arena slice if present)."""
        arena = getattr(p, "_dpt_arena", None)
        if arena is not None:
            return arena.flat_grad_view(p)
        if p.grad is None:
            p.grad = torch.zeros_like(p, memory_format=torch.preserve_format)
        g = p.grad
        if g.dim() == 4 and not g.is_contiguous():
            assert g.permute(0, 2, 3, 1).is_contiguous(), "conv gradients must be channels_last"
            return g.permute(0, 2, 3, 1).reshape(-1)
        return g.view(-1)

    # ------------------------------------------------------------------ forward
    def _ev2_next(self):
        ev = self._ev2[self._ev2_i % len(self._ev2)]
        self._ev2_i += 1
        return ev

    def _fwd_unit(self, u: _Unit, x, residual=None, residual_event=None):
        Fn = self.Fn
        conv, bn = u.conv, u.bn
        if residual_event is not None:
            # the residual comes from the downsample branch on side2: run our conv first, then
            # join right before the BN kernel that consumes it
            return self._fwd_unit_join(u, x, residual, residual_event)
        if self.training and u.fsplit > 1:
            # tiny tile grid: split-K over CTAs into an fp32 accumulator, then one column-strip
            # kernel does statistics + normalisation (+residual, +ReLU)
            Fn.conv_fprop(x, self._w(conv), conv.stride, conv.padding, out=u.raw32, splits=u.fsplit)
            Fn.bn_fwd_train_small_f32(u.raw32, bn.weight, bn.bias, residual=residual, relu=u.relu,
                                      out=u.y, x_bf16=u.raw, save_mean=u.mean, save_invstd=u.invstd,
                                      running_mean=bn.running_mean, running_var=bn.running_var,
                                      num_batches_tracked=bn.num_batches_tracked,
                                      momentum=bn.momentum, eps=bn.eps)
        elif self.training:
            Fn.conv_fprop(x, self._w(conv), conv.stride, conv.padding, out=u.raw,
                          stats=(u.sum, u.sq))
            Fn.bn_fwd_train(u.raw, u.sum, u.sq, bn.weight, bn.bias, residual=residual, relu=u.relu,
                            out=u.y, save_mean=u.mean, save_invstd=u.invstd,
                            running_mean=bn.running_mean, running_var=bn.running_var,
                            num_batches_tracked=bn.num_batches_tracked, momentum=bn.momentum,
                            eps=bn.eps)
        else:
            Fn.conv_fprop(x, self._w(conv), conv.stride, conv.padding, out=u.raw)
            Fn.bn_fwd_eval(u.raw, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                           residual=residual, relu=u.relu, out=u.y, eps=bn.eps)
        return u.y

    def _fwd_unit_join(self, u: _Unit, x, residual, ev):
        Fn = self.Fn
        conv, bn = u.conv, u.bn
        split = u.fsplit > 1
        if split:
            Fn.conv_fprop(x, self._w(conv), conv.stride, conv.padding, out=u.raw32, splits=u.fsplit)
        else:
            Fn.conv_fprop(x, self._w(conv), conv.stride, conv.padding, out=u.raw, stats=(u.sum, u.sq))
        torch.cuda.current_stream().wait_event(ev)
        self._lib.pdl_skip_next()
        if split:
            Fn.bn_fwd_train_small_f32(u.raw32, bn.weight, bn.bias, residual=residual, relu=u.relu,
                                      out=u.y, x_bf16=u.raw, save_mean=u.mean, save_invstd=u.invstd,
                                      running_mean=bn.running_mean, running_var=bn.running_var,
                                      num_batches_tracked=bn.num_batches_tracked,
                                      momentum=bn.momentum, eps=bn.eps)
        else:
            Fn.bn_fwd_train(u.raw, u.sum, u.sq, bn.weight, bn.bias, residual=residual, relu=u.relu,
                            out=u.y, save_mean=u.mean, save_invstd=u.invstd,
                            running_mean=bn.running_mean, running_var=bn.running_var,
                            num_batches_tracked=bn.num_batches_tracked, momentum=bn.momentum,
                            eps=bn.eps)
        return u.y

    def forward(self, x_nchw: torch.Tensor) -> torch.Tensor:
        Fn = self.Fn
        m = self.m
        x_nchw = x_nchw.contiguous()
        if x_nchw.dtype != torch.float32:
            x_nchw = x_nchw.float()
        self._last_input = x_nchw
        Fn.set_lowp_dtype(self.dtype)
        if self.training:
            self.stats.zero_()
            if self._zero32_n:
                with self._fork_side():      # off the critical path: first used in layer2
                    self.zero32.zero_()
                self._z32_ev.record(self.side)
                self._z32_pending = True
        # ---- stem: im2col + GEMM (+BN statistics in the epilogue) ----------------------
        st = self.stem
        N, P, Q, _ = st.y_shape
        Fn.im2col_nchw(x_nchw, 7, 7, 2, 3, self.STEM_LDK, out=self.cols)
        Fn.pad_rows_to_bf16(m.conv1.weight, 64, 147, self.STEM_LDK, out=self.stem_w)
        M = N * P * Q
        if self.training:
            Fn.gemm(0, self.cols, self.stem_w, st.raw, M, 64, 147, self.STEM_LDK, self.STEM_LDK, 64,
                    out_mode=0, stats=(st.sum, st.sq))
            bn = m.bn1
            Fn.bn_fwd_train(st.raw, st.sum, st.sq, bn.weight, bn.bias, relu=True, out=st.y,
                            save_mean=st.mean, save_invstd=st.invstd, running_mean=bn.running_mean,
                            running_var=bn.running_var, num_batches_tracked=bn.num_batches_tracked,
                            momentum=bn.momentum, eps=bn.eps)
        else:
            Fn.gemm(0, self.cols, self.stem_w, st.raw, M, 64, 147, self.STEM_LDK, self.STEM_LDK, 64,
                    out_mode=0)
            bn = m.bn1
            Fn.bn_fwd_eval(st.raw, bn.weight, bn.bias, bn.running_mean, bn.running_var, relu=True,
                           out=st.y, eps=bn.eps)
        Fn.maxpool_fwd(st.y, out=self.pool_y, idx=self.pool_idx)
        x = self.pool_y
        # ---- residual stages ------------------------------------------------------------
        for b in self.blocks:
            b["x"] = x
            if self.training and self._z32_pending and any(u.fsplit > 1 for u in b["units"] + ([b["ds"]] if b["ds"] else [])):
                torch.cuda.current_stream().wait_event(self._z32_ev)
                self._lib.pdl_skip_next()
                self._z32_pending = False
            ds_ev = None
            if b["ds"] is None:
                idn = x
            elif self.training and self._branch_streams:
                # downsample branch on side2, concurrently with the main branch
                ev = self._ev2_next()
                ev.record(torch.cuda.current_stream())
                self.side2.wait_event(ev)
                self._lib.pdl_skip_next()
                with torch.cuda.stream(self.side2):
                    idn = self._fwd_unit(b["ds"], x)
                    ds_ev = self._ev2_next()
                    ds_ev.record(self.side2)
            else:
                idn = self._fwd_unit(b["ds"], x)
            h = x
            for u in b["units"][:-1]:
                h = self._fwd_unit(u, h)
            x = self._fwd_unit(b["units"][-1], h, residual=idn, residual_event=ds_ev)
        # ---- head -----------------------------------------------------------------------
        Fn.head_fwd(x, m.fc.weight, m.fc.bias, pooled=self.pooled, logits=self.logits)
        return self.logits

    # ------------------------------------------------------------------ backward
    def _ready(self, *params) -> None:
        sink = self.m._grad_sink
        if sink is not None:
            sink(params)

    def _bn_bwd(self, u: _Unit, dy, dy2=None):
        """Fused (gradient-pair sum + ReLU mask +) BatchNorm backward of unit ``u``.
        Leaves grad wrt the conv output in ``u.d_raw`` and the masked gradient in ``u.g``."""
        bn = u.bn
        nvec = u.raw.numel() // 8
        if self._fuse_bn_bwd and nvec <= 148 * 256:
            # small tensor: one launch (grid barrier between the reduction and the dx pass)
            self.Fn.bn_bwd_fused(dy, u.raw, u.mean, u.invstd, bn.weight, self._gw(bn.weight),
                                 self._gw(bn.bias), u.bscratch, u.bcounter, dy2=dy2,
                                 y=u.y if u.relu else None, g_out=u.g, dx=u.d_raw)
            return u.d_raw
        masked = u.relu or dy2 is not None
        self.Fn.bn_bwd(dy, u.raw, u.mean, u.invstd, bn.weight, self._gw(bn.weight),
                       self._gw(bn.bias), dy2=dy2, y=u.y if u.relu else None,
                       g_out=u.g if masked else None, dx=u.d_raw)
        return u.d_raw

    def _fork_side(self):
        """Side stream picks up everything enqueued on the current stream so far."""
        ev = self._ev_pool[self._ev_i % len(self._ev_pool)]
        self._ev_i += 1
        ev.record(torch.cuda.current_stream())
        self.side.wait_event(ev)
        self._lib.pdl_skip_next()     # first kernel on the side stream follows an event wait
        return torch.cuda.stream(self.side)

    def _wgrad(self, u: _Unit, x_in):
        conv = u.conv
        Cout, Cin, R, S = conv.weight.shape
        with self._fork_side():
            self.Fn.conv_wgrad(u.d_raw, x_in, (Cout, R, S, Cin), conv.stride, conv.padding,
                               out=self._gw(conv.weight))
            self._ready(conv.weight, u.bn.weight, u.bn.bias)

    def _dgrad(self, u: _Unit, x_shape, out, u2: Optional[_Unit] = None):
        conv = u.conv
        splits = u.dsplit if out.dtype == torch.float32 else 1
        if u2 is None:
            self.Fn.conv_dgrad(u.d_raw, self._w(conv), x_shape, conv.stride, conv.padding, out=out,
                               splits=splits)
        else:
            self.Fn.conv_dgrad(u.d_raw, self._w(conv), x_shape, conv.stride, conv.padding, out=out,
                               dy2=u2.d_raw, w2=self._w(u2.conv), pad2=u2.conv.padding,
                               splits=splits)
        return out

    @staticmethod
    def _can_fuse_dgrad(u1, ds) -> bool:
        """The fused dgrad kernel accumulates two convs that read the same input in one TMEM
        tile; it needs the same stride and the same number of output channels (BasicBlock)."""
        return (ds is not None and u1.conv.stride == ds.conv.stride
                and u1.conv.out_channels == ds.conv.out_channels)

    def _bwd_block(self, b, dy, dy2):
        units, ds = b["units"], b["ds"]
        k = len(units)
        # block output: sum the incoming pair, apply the output-ReLU mask, BN backward
        self._bn_bwd(units[-1], dy, dy2)
        g_skip = units[-1].g                       # gradient flowing into the identity branch
        if ds is not None and self._branch_streams:
            ev = self._ev2_next()
            ev.record(torch.cuda.current_stream())
            self.side2.wait_event(ev)
            self._lib.pdl_skip_next()
            with torch.cuda.stream(self.side2):
                self._bn_bwd(ds, g_skip)
                dev_ = self._ev2_next()
                dev_.record(self.side2)
            b["_ds_bwd_ev"] = dev_
        for i in range(k - 1, 0, -1):
            u, prev = units[i], units[i - 1]
            d_in = self._dgrad(u, prev.y_shape, b["d_mid"][i - 1])
            self._wgrad(u, prev.y)
            self._bn_bwd(prev, d_in)
        u1 = units[0]
        if ds is None:
            self._dgrad(u1, b["x_shape"], b["dx"])
            self._wgrad(u1, b["x"])
            return b["dx"], g_skip
        ds_ev = None
        if self._branch_streams:
            # g_skip exists since the block-output BN backward; the downsample BN backward does
            # not depend on the main branch, so it ran on side2 (issued below, before the chain)
            ds_ev = b.get("_ds_bwd_ev")
            torch.cuda.current_stream().wait_event(ds_ev)
            self._lib.pdl_skip_next()
        else:
            self._bn_bwd(ds, g_skip)
        if self._can_fuse_dgrad(u1, ds):
            # conv1 and the 1x1 downsample read the same block input: one fused dgrad kernel
            self._dgrad(u1, b["x_shape"], b["dx"], ds)
            pair = (b["dx"], None)
        else:
            self._dgrad(u1, b["x_shape"], b["dx"])
            self._dgrad(ds, b["x_shape"], b["dx_ds"])
            pair = (b["dx"], b["dx_ds"])
        self._wgrad(u1, b["x"])
        self._wgrad(ds, b["x"])
        return pair

    def backward(self, dlogits: torch.Tensor) -> None:
        Fn = self.Fn
        m = self.m
        dlogits = dlogits.contiguous()
        Fn.set_lowp_dtype(self.dtype)
        if self._z32_pending:
            torch.cuda.current_stream().wait_event(self._z32_ev)
            self._lib.pdl_skip_next()
            self._z32_pending = False
        # input gradient on the critical path first (it READS fc.weight: with the optimizer fused
        # into the all-reduce, reporting fc ready lets a bucket kernel rewrite the weights, so the
        # side stream is forked only after this kernel is enqueued — same rule as every conv:
        # dgrad before wgrad); weight / bias gradient on the side stream
        Fn.head_bwd(dlogits, m.fc.weight, self.pooled, self.feat_shape, None, None, dx=self.d_feat,
                    part="dx")
        with self._fork_side():
            Fn.head_bwd(dlogits, m.fc.weight, self.pooled, self.feat_shape, self._gw(m.fc.weight),
                        self._gw(m.fc.bias), part="dw")
            self._ready(m.fc.weight, m.fc.bias)
        dy, dy2 = self.d_feat, None
        for b in reversed(self.blocks):
            dy, dy2 = self._bwd_block(b, dy, dy2)
        # ---- stem: max-pool backward sums the pair, then BN+ReLU backward, then wgrad -------
        st = self.stem
        Fn.maxpool_bwd(dy, self.pool_idx, st.y_shape, out=self.d_pool_in, dy2=dy2)
        self._bn_bwd(st, self.d_pool_in)
        N, P, Q, _ = st.y_shape
        M = N * P * Q
        with self._fork_side():
            Fn.gemm(2, st.d_raw, self.cols, self._gw(m.conv1.weight), 64, 147, M, 64, self.STEM_LDK,
                    147, out_mode=2, splits=max(1, min(64, M // 512)))
            self._ready(m.conv1.weight, m.bn1.weight, m.bn1.bias)
        torch.cuda.current_stream().wait_stream(self.side)   # join
        self._lib.pdl_skip_next()
