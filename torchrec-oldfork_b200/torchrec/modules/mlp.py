"""Perceptron / MLP (reference torchrec/modules/mlp.py).  Dense layers stay on cuBLAS through
PyTorch (BASELINE.json north_star).  The only B200-specific piece is the memory layout of a layer
whose fan-in is not a multiple of 4 floats (the 479-wide first over-arch layer of Criteo DLRM):
RowPitchedLinear keeps the weight rows at a 16-byte pitch with zeros in the pad columns, so that an
input produced at the same pitch (InteractionArch's output) runs through cuBLAS's aligned TF32
kernels as a fan-in-480 GEMM; same parameter shapes, same state_dict, same arithmetic."""
from typing import Callable, List, Optional, Union

import torch
from torch import nn
from torch.nn import functional as F

ZERO_PADDED_ATTR = "_b200_zero_padded_rows"  # set on tensors whose row padding is known to hold zeros


def mark_zero_padded(t: torch.Tensor) -> torch.Tensor:
    setattr(t, ZERO_PADDED_ATTR, True)
    return t


def _pitch(width: int) -> int:
    return (width + 3) // 4 * 4


def _fusable(x: torch.Tensor, out_features: int, bias) -> bool:
    return (x.is_cuda and x.dim() == 2 and x.dtype == torch.float32 and bias is not None and
            out_features % 4 == 0 and out_features <= 1024 and x.shape[0] > 0)


class _LinearReLUFn(torch.autograd.Function):
    """relu(x W^T + b).  Forward: one cuBLASLt GEMM with the bias + ReLU epilogue
    (torch._addmm_activation).  Backward: ReLU mask and bias gradient in one pass over grad
    (b200_relu_bias_grad) instead of threshold_backward + sum(0); the two GEMMs stay on cuBLAS.
    `pitch` > 0: x and W are views of zero-padded buffers with that row pitch (RowPitchedLinear)."""

    @staticmethod
    def forward(ctx, x, weight, bias, pitch: int):
        K = x.shape[1]
        if pitch:
            x = x.as_strided((x.shape[0], pitch), (pitch, 1), x.storage_offset())
            weight = weight.as_strided((weight.shape[0], pitch), (pitch, 1), weight.storage_offset())
        elif not x.is_contiguous():
            x = x.contiguous()
        y = torch._addmm_activation(bias, x, weight.t(), use_gelu=False)
        ctx.save_for_backward(x, weight, y)
        ctx.K = K
        return y

    @staticmethod
    def backward(ctx, gy):
        from b200_sparse import ops as _b200
        x, weight, y = ctx.saved_tensors
        g, gb = _b200.relu_bias_grad(gy, y)
        gx = gw = None
        if ctx.needs_input_grad[0]:
            gx = (g @ weight)[:, :ctx.K]  # pitched: the pad column is g . 0 = 0
        if ctx.needs_input_grad[1]:
            gw = (g.t() @ x)[:, :ctx.K]
        return gx, gw, (gb if ctx.needs_input_grad[2] else None), None


class _PitchedLinearFn(torch.autograd.Function):
    """y = x W^T + b on the pitch-wide buffers behind x [B, K] and W [N, K] (pad columns are zero)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        K, Kp = x.shape[1], x.stride(0)
        xf = x.as_strided((x.shape[0], Kp), (Kp, 1), x.storage_offset())
        wf = weight.as_strided((weight.shape[0], Kp), (Kp, 1), weight.storage_offset())
        ctx.save_for_backward(xf, wf)
        ctx.K, ctx.has_bias = K, bias is not None
        return torch.addmm(bias, xf, wf.t()) if bias is not None else xf @ wf.t()

    @staticmethod
    def backward(ctx, gy):
        xf, wf = ctx.saved_tensors
        gy = gy.contiguous()
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            gx = (gy @ wf)[:, :ctx.K]  # pad column = gy . 0 = 0: stays a zero-padded pitched view
        if ctx.needs_input_grad[1]:
            gw = (gy.t() @ xf)[:, :ctx.K]
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = gy.sum(0)
        return gx, gw, gb


class RowPitchedLinear(nn.Linear):
    """nn.Linear (same parameters / state_dict) whose weight storage is [out, round_up(in, 4)]."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, device=None, dtype=None) -> None:
        super().__init__(in_features, out_features, bias=bias, device=device, dtype=dtype)
        self._repitch()

    def _pitched(self) -> bool:
        w, Kp = self.weight, _pitch(self.in_features)
        return (w.dim() == 2 and w.stride(1) == 1 and w.stride(0) == Kp and
                w.untyped_storage().nbytes() >= (w.storage_offset() + w.shape[0] * Kp) * w.element_size())

    def _repitch(self) -> None:
        w, Kp = self.weight, _pitch(self.in_features)
        if Kp == self.in_features or w.device.type != "cuda" or w.dtype != torch.float32 or self._pitched():
            return
        full = torch.zeros((self.out_features, Kp), dtype=w.dtype, device=w.device)
        full[:, :self.in_features].copy_(w.detach())
        self.weight.data = full[:, :self.in_features]

    def _apply(self, fn, *args, **kwargs):
        r = super()._apply(fn, *args, **kwargs)
        self._repitch()
        return r

    def _pitched_input(self, input: torch.Tensor) -> bool:
        Kp = _pitch(self.in_features)
        return (Kp != self.in_features and input.is_cuda and input.dim() == 2 and input.dtype == torch.float32 and
                getattr(input, ZERO_PADDED_ATTR, False) and input.stride(1) == 1 and input.stride(0) == Kp and
                input.storage_offset() % 4 == 0 and self._pitched())

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        if self._pitched_input(input):
            return _PitchedLinearFn.apply(input, self.weight, self.bias)
        return F.linear(input, self.weight, self.bias)

    def forward_relu(self, input: torch.Tensor) -> torch.Tensor:
        """relu(linear(input)) with the bias + ReLU fused into the GEMM epilogue and a one-pass backward."""
        if _fusable(input, self.out_features, self.bias):
            if self._pitched_input(input):
                return _LinearReLUFn.apply(input, self.weight, self.bias, _pitch(self.in_features))
            if self.weight.is_contiguous():
                return _LinearReLUFn.apply(input, self.weight, self.bias, 0)
        return torch.relu(self.forward(input))


class Perceptron(nn.Module):
    def __init__(self, in_size: int, out_size: int, bias: bool = True,
                 activation: Union[nn.Module, Callable[[torch.Tensor], torch.Tensor]] = torch.relu,
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._out_size, self._in_size = out_size, in_size
        self._linear = RowPitchedLinear(in_size, out_size, bias=bias, device=device)
        self._activation_fn = activation

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        if self._activation_fn is torch.relu or isinstance(self._activation_fn, nn.ReLU):
            return self._linear.forward_relu(input)
        return self._activation_fn(self._linear(input))


class MLP(nn.Module):
    def __init__(self, in_size: int, layer_sizes: List[int], bias: bool = True,
                 activation: Union[str, Callable[[], nn.Module], nn.Module,
                                   Callable[[torch.Tensor], torch.Tensor]] = torch.relu,
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        if activation == "relu":
            activation = torch.relu
        elif activation == "sigmoid":
            activation = torch.sigmoid
        if not callable(activation) and not isinstance(activation, nn.Module):
            raise ValueError("activation must be 'relu', 'sigmoid', a callable or an nn.Module")
        sizes = [in_size] + list(layer_sizes)
        self._mlp = nn.Sequential(*[
            Perceptron(sizes[i], sizes[i + 1], bias=bias, activation=activation, device=device)
            for i in range(len(layer_sizes))])

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return self._mlp(input)
