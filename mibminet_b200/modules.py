"""Drop-in `nn.Module`s with the reference's operator surface, running the int8 CUDA path.

* `edgeEEGNet`      <- QuantLab/quantlab/{PhysionetMMMI,BCI-CompIV-2a}/edgeEEGNet/edgeEEGnet.py:11-178
                       same constructor keywords, same state_dict keys, `forward(x[B,1,C,T]) -> [B,N]`.
* `edgeEEGNetQuant` <- edge-eegnet_torch/edgeEEGnet_quant.py:10-107, `forward(x[B,C,T]) -> [B,N]`.
* `EEGNet`          <- QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:11-168 (temporal conv first, Nf = 128).

Only inference is implemented (the training behaviour of INQ / STE / dropout is out of scope): the
forward quantises the float input like `quant1` (STEActivation with floorToZero,
QuantLab/quantlab/indiv/ste_ops.py:88-102), runs the integer network on the GPU and de-quantises the
logits with `quant6.absMaxValue` -- i.e. it returns what the reference's *integer* deployment
(golden_model.py) returns, which the reference itself accepts within 10 LSB of its float-simulated
forward on trained nets (test/python/golden_model/testcase.py:42).
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .convert import NF, params_from_state_dict
from .engine import EEGNetInt8, Int8Net


class _Quant(nn.Module):
    """state_dict-compatible stand-in for STEActivation: one buffer `absMaxValue[1]`."""

    def __init__(self):
        super().__init__()
        self.register_buffer("absMaxValue", torch.ones(1))


class _InqParam(nn.Module):
    """state_dict-compatible stand-in for INQConv2d / INQLinear: weight, weightFrozen, sParam (+bias)."""

    def __init__(self, shape, bias: Optional[int] = None):
        super().__init__()
        self.weight = nn.Parameter(torch.zeros(shape), requires_grad=False)
        self.register_buffer("weightFrozen", torch.full(shape, float("nan")))
        self.register_buffer("sParam", torch.ones(1))
        if bias is not None:
            self.bias = nn.Parameter(torch.zeros(bias), requires_grad=False)


class edgeEEGNet(nn.Module):
    """Quantized edgeEEGNet (inference), integer pipeline on the GPU."""

    def __init__(self, F1=16, D=1, F2=None, C=64, T=480, N=4, Nf=128, p_dropout=0.5,
                 dropout_type='TimeDropout2d', quantWeight=True, quantAct=True,
                 weightInqSchedule=None, weightInqNumLevels=255, weightInqStrategy="matnitude",
                 weightInqInitMethod="uniform", actSTENumLevels=255, actSTEStartEpoch=2,
                 floorToZero=False, actFirstLayerNumLevels=None, weightFirstLayerNumLevels=None,
                 first_layer_only=False, reorder_bn=True):
        super().__init__()
        if F2 is None:
            F2 = F1 * D
        # constraints of the integer tool-chain (golden_model.py:61-83): the reference's own defaults (floorToZero=False,
        # Nf=128) describe nets its integer deployment cannot run either, so they are refused here, at construction
        if weightInqNumLevels != 255 or actSTENumLevels != 255:
            raise ValueError("the integer pipeline exists for 255 levels only (golden_model.py:61-62)")
        if not floorToZero:
            raise ValueError("the integer pipeline needs floorToZero=True (golden_model.py:65): truncating activations")
        if F2 != F1 * D or D != 1:
            raise ValueError("F2 must equal F1 * D with D == 1 (golden_model.py:83,169)")
        if Nf != NF:
            raise ValueError("the integer tool-chain only supports Nf == 64 (golden_model.py:216-219); pass Nf=64")
        self.F1, self.D, self.F2, self.C, self.T, self.N, self.Nf = F1, D, F2, C, T, N, Nf
        self.reorder_bn = reorder_bn
        n_features = (T // 8) // 8
        self.quant1 = _Quant()
        self.conv1 = _InqParam((F1, 1, C, 1))
        self.batch_norm1 = nn.BatchNorm2d(F1, momentum=0.01, eps=0.001)
        self.quant2 = _Quant()
        self.conv2 = _InqParam((F1 * D, 1, 1, Nf))
        self.batch_norm2 = nn.BatchNorm2d(F1 * D, momentum=0.01, eps=0.001)
        self.quant3 = _Quant()
        self.sep_conv1 = _InqParam((D * F1, 1, 1, 16))
        self.quant4 = _Quant()
        self.sep_conv2 = _InqParam((F2, D * F1, 1, 1))
        self.batch_norm3 = nn.BatchNorm2d(F2, momentum=0.01, eps=0.001)
        self.quant5 = _Quant()
        self.fc = _InqParam((N, F2 * n_features), bias=N)
        self.quant6 = _Quant()
        self._engine: Optional[Int8Net] = None
        self._engine_key = None
        self.eval()

    # any parameter change invalidates the packed integer network: load_state_dict / .to() / .cuda() directly, in-place
    # edits of parameters or buffers (quant1.absMaxValue.fill_(...)) through the tensors' version counters
    def load_state_dict(self, state_dict, strict: bool = True, **kw):
        sd = {k: (torch.as_tensor(v) if not torch.is_tensor(v) else v) for k, v in state_dict.items()}
        r = super().load_state_dict(sd, strict=strict, **kw)
        self._engine = None
        return r

    def _apply(self, fn, *a, **kw):
        self._engine = None
        return super()._apply(fn, *a, **kw)

    def _state_key(self):
        return tuple((t.data_ptr(), t._version) for t in self.state_dict(keep_vars=True).values())

    def integer_network(self, device=None) -> Int8Net:
        """The packed integer network (built lazily from the current state_dict)."""
        dev = torch.device(device) if device is not None else self.quant1.absMaxValue.device
        if dev.type != "cuda":
            raise RuntimeError("edgeEEGNet (mibminet_b200) runs on CUDA only: call .cuda() first")
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        key = self._state_key()
        if self._engine is None or self._engine.device_index != idx or self._engine_key != key:
            sd = {k: v.detach().cpu().numpy() for k, v in self.state_dict().items()}
            params = params_from_state_dict(sd, C=self.C, T=self.T, F1=self.F1, D=self.D, F2=self.F2, N=self.N)
            self._engine = Int8Net(params, device=torch.device("cuda", idx), reorder_bn=self.reorder_bn)
            self._engine_key = key
        return self._engine

    @torch.no_grad()
    def forward_int8(self, x_int8: torch.Tensor) -> torch.Tensor:
        """int8 [B, (1,) C, T] -> int8 logits [B, N] (the benchmarked path)."""
        return self.integer_network(x_int8.device).forward(x_int8)

    @torch.no_grad()
    def forward(self, x: torch.Tensor, with_stats: bool = False):
        # input dimensions: (s, 1, C, T)
        if x.dim() == 4:
            x = x[:, 0]
        net = self.integer_network(x.device)
        if x.dtype == torch.int8:
            y = net.forward(x)
        else:
            y = net.forward(net.quantize_input(x))
        out = net.dequantize_logits(y)
        if with_stats:
            stats = [('conv1_w', self.conv1.weight.data), ('conv2_w', self.conv2.weight.data),
                     ('sep_conv1_w', self.sep_conv1.weight.data), ('sep_conv2_w', self.sep_conv2.weight.data),
                     ('fc_w', self.fc.weight.data), ('fc_b', self.fc.bias.data)]
            return stats, out
        return out

    def forward_with_tensor_stats(self, x):
        return self.forward(x, with_stats=True)


class EEGNet(nn.Module):
    """Quantized EEGNet / Q-EEGNet (inference): QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:11-168 -- temporal conv
    (1, 128) first, then the depthwise spatial conv.  Same constructor keywords and state_dict keys; the forward runs the
    integer pipeline of mibmi_eegnet.cu (no integer model of this order exists in the reference; it tracks the reference's
    float-simulated forward within a few LSB, tests/test_eegnet_order.py)."""

    def __init__(self, F1=8, D=2, F2=None, C=22, T=1125, N=4, p_dropout=0.5,
                 dropout_type='TimeDropout2d', quantWeight=True, quantAct=True,
                 weightInqSchedule=None, weightInqNumLevels=255, weightInqStrategy="matnitude",
                 weightInqInitMethod="uniform", actSTENumLevels=255, actSTEStartEpoch=2,
                 floorToZero=False, actFirstLayerNumLevels=None, weightFirstLayerNumLevels=None,
                 first_layer_only=False, Nf=128):
        super().__init__()
        if F2 is None:
            F2 = F1 * D
        if weightInqNumLevels != 255 or actSTENumLevels != 255:
            raise ValueError("the integer pipeline exists for 255 levels only")
        self.F1, self.D, self.F2, self.C, self.T, self.N, self.Nf = F1, D, F2, C, T, N, Nf
        n_features = (T // 8) // 8
        self.quant1 = _Quant()
        self.conv1 = _InqParam((F1, 1, 1, Nf))
        self.batch_norm1 = nn.BatchNorm2d(F1, momentum=0.01, eps=0.001)
        self.quant2 = _Quant()
        self.conv2 = _InqParam((D * F1, 1, C, 1))
        self.batch_norm2 = nn.BatchNorm2d(D * F1, momentum=0.01, eps=0.001)
        self.quant3 = _Quant()
        self.sep_conv1 = _InqParam((D * F1, 1, 1, 16))
        self.quant4 = _Quant()
        self.sep_conv2 = _InqParam((F2, D * F1, 1, 1))
        self.batch_norm3 = nn.BatchNorm2d(F2, momentum=0.01, eps=0.001)
        self.quant5 = _Quant()
        self.fc = _InqParam((N, F2 * n_features), bias=N)
        self.quant6 = _Quant()
        self._engine: Optional[EEGNetInt8] = None
        self._engine_key = None
        if not floorToZero:
            raise ValueError("the integer pipeline needs floorToZero=True: truncating activations")
        self.eval()

    def load_state_dict(self, state_dict, strict: bool = True, **kw):
        sd = {k: (torch.as_tensor(v) if not torch.is_tensor(v) else v) for k, v in state_dict.items()}
        r = super().load_state_dict(sd, strict=strict, **kw)
        self._engine = None
        return r

    def _apply(self, fn, *a, **kw):
        self._engine = None
        return super()._apply(fn, *a, **kw)

    def integer_network(self, device=None) -> EEGNetInt8:
        dev = torch.device(device) if device is not None else self.quant1.absMaxValue.device
        if dev.type != "cuda":
            raise RuntimeError("EEGNet (mibminet_b200) runs on CUDA only: call .cuda() first")
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        key = tuple((t.data_ptr(), t._version) for t in self.state_dict(keep_vars=True).values())
        if self._engine is None or self._engine.device_index != idx or self._engine_key != key:
            sd = {k: v.detach().cpu().numpy() for k, v in self.state_dict().items()}
            self._engine = EEGNetInt8.from_state_dict(sd, C=self.C, T=self.T, F1=self.F1, D=self.D, F2=self.F2, N=self.N,
                                                      device=torch.device("cuda", idx))
            self._engine_key = key
        return self._engine

    @torch.no_grad()
    def forward_int8(self, x_int8: torch.Tensor) -> torch.Tensor:
        return self.integer_network(x_int8.device).forward(x_int8)

    @torch.no_grad()
    def forward(self, x: torch.Tensor, with_stats: bool = False):
        # input dimensions: (s, 1, C, T)
        if x.dim() == 4:
            x = x[:, 0]
        net = self.integer_network(x.device)
        y = net.forward(x if x.dtype == torch.int8 else net.quantize_input(x))
        out = net.dequantize_logits(y)
        if with_stats:
            stats = [('conv1_w', self.conv1.weight.data), ('conv2_w', self.conv2.weight.data),
                     ('sep_conv1_w', self.sep_conv1.weight.data), ('sep_conv2_w', self.sep_conv2.weight.data),
                     ('fc_w', self.fc.weight.data), ('fc_b', self.fc.bias.data)]
            return stats, out
        return out

    def forward_with_tensor_stats(self, x):
        return self.forward(x, with_stats=True)


class edgeEEGNetQuant(nn.Module):
    """Signature of edge-eegnet_torch/edgeEEGnet_quant.py:10-107 over the Wolf integer scheme.

    The reference module's arithmetic after `tq.convert` lives in PyTorch's qnnpack backend (affine
    quint8, float requant multipliers) and is not pinned by any reference test; this class keeps the
    constructor / forward signature and computes with the integer parameters given to
    `load_integer_params` (a WolfParams or a QuantLab-format state_dict)."""

    def __init__(self, qconfig=None, F1=16, D=1, F2=None, C=22, T=1125, N=4, p_dropout=0.5,
                 dropout_type='TimeDropout2D'):
        super().__init__()
        self.qconfig = qconfig
        self.F1, self.D, self.F2, self.C, self.T, self.N = F1, D, (F2 if F2 is not None else F1 * D), C, T, N
        self._params = None
        self._engine: Optional[Int8Net] = None

    def load_integer_params(self, params_or_state_dict):
        from .convert import WolfParams
        if isinstance(params_or_state_dict, WolfParams):
            self._params = params_or_state_dict
        else:
            sd = {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else v) for k, v in params_or_state_dict.items()}
            self._params = params_from_state_dict(sd, C=self.C, T=self.T, F1=self.F1, D=self.D, F2=self.F2, N=self.N)
        self._engine = None
        return self

    @torch.no_grad()
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        # input dimensions: (s, C, T)
        if self._params is None:
            raise RuntimeError("edgeEEGNetQuant: call load_integer_params() first")
        if not x.is_cuda:
            raise RuntimeError("edgeEEGNetQuant (mibminet_b200) runs on CUDA only")
        idx = x.device.index if x.device.index is not None else torch.cuda.current_device()
        if self._engine is None or self._engine.device_index != idx:
            self._engine = Int8Net(self._params, device=torch.device("cuda", idx))
        net = self._engine
        y = net.forward(x if x.dtype == torch.int8 else net.quantize_input(x))
        return net.dequantize_logits(y)
