"""Tensor-level entry points: torch CUDA tensors in, torch CUDA tensors out, every op a call
into libflowcuda.so on torch's current stream.  torch is used for device memory and streams
only; no torch kernel computes flow math here."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import CouplingSpec, check


def _dev_check(*ts):
    dev = None
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda:
            raise _lib.FlowCudaError("flowcuda ops need CUDA tensors (there is no CPU path)")
        if t.dtype != torch.float32:
            raise _lib.FlowCudaError(f"flowcuda ops are float32; got {t.dtype}")
        if not t.is_contiguous():
            raise _lib.FlowCudaError("flowcuda ops need contiguous tensors")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise _lib.FlowCudaError(f"flowcuda ops need all tensors on one device; got {dev} and {t.device}")


def _on_tensor_device(fn):
    """Run an op in the device context (and on the current stream) of its first CUDA tensor argument: the kernels launch
    in the CURRENT context, which in a process driving several GPUs need not be the tensors' device."""
    import functools

    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        dev = None
        for a in list(args) + list(kwargs.values()):
            if isinstance(a, torch.Tensor) and a.is_cuda:
                dev = a.device
                break
        if dev is None or dev.index is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)

    return wrapped


def _f32(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(dtype=torch.float32).contiguous()


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _stream():
    # current stream of the current device; _dev_check has made the tensors' device current
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def flip_mask_of(flips: Sequence[bool]) -> int:
    m = 0
    for i, f in enumerate(flips):
        if f:
            m |= 1 << i
    return m


# ---------------------------------------------------------------------------- row #1
@_on_tensor_device
def realnvp_log_prob(params: torch.Tensor, n_layers: int, hidden: int, flip_mask: int, y: torch.Tensor,
                     want_z: bool = False):
    y = _f32(y)
    _dev_check(params, y)
    n = y.shape[0]
    lp = torch.empty(n, device=y.device, dtype=torch.float32)
    z = torch.empty_like(y) if want_z else None
    check(_lib.load().flowcuda_realnvp_log_prob(_ptr(params), n_layers, hidden, flip_mask, _ptr(y), _ptr(lp), _ptr(z),
                                                n, _stream()), "realnvp_log_prob")
    return (lp, z) if want_z else lp


@_on_tensor_device
def realnvp_sample(params: torch.Tensor, n_layers: int, hidden: int, flip_mask: int, z: torch.Tensor) -> torch.Tensor:
    z = _f32(z)
    _dev_check(params, z)
    x = torch.empty_like(z)
    check(_lib.load().flowcuda_realnvp_sample(_ptr(params), n_layers, hidden, flip_mask, _ptr(z), _ptr(x), z.shape[0],
                                              _stream()), "realnvp_sample")
    return x


@_on_tensor_device
def realnvp_forward(layer_params: torch.Tensor, hidden: int, flip: bool, x: torch.Tensor) -> torch.Tensor:
    x = _f32(x)
    _dev_check(layer_params, x)
    y = torch.empty_like(x)
    check(_lib.load().flowcuda_realnvp_forward(_ptr(layer_params), hidden, int(flip), _ptr(x), _ptr(y), x.shape[0],
                                               _stream()), "realnvp_forward")
    return y


@_on_tensor_device
def realnvp_inverse(layer_params: torch.Tensor, hidden: int, flip: bool, y: torch.Tensor):
    y = _f32(y)
    _dev_check(layer_params, y)
    x = torch.empty_like(y)
    ls = torch.empty(y.shape[0], 1, device=y.device, dtype=torch.float32)
    check(_lib.load().flowcuda_realnvp_inverse(_ptr(layer_params), hidden, int(flip), _ptr(y), _ptr(x), _ptr(ls),
                                               y.shape[0], _stream()), "realnvp_inverse")
    return x, ls


@_on_tensor_device
def realnvp_nll_grad(params: torch.Tensor, n_layers: int, hidden: int, flip_mask: int, y: torch.Tensor,
                     lr: float = 0.0, params_out: Optional[torch.Tensor] = None,
                     workspace: Optional[torch.Tensor] = None):
    """(loss[1], grad[L*(4H+2)]); with lr != 0 also writes p - lr*g into params_out."""
    y = _f32(y)
    _dev_check(params, y, params_out)
    lib = _lib.load()
    n = y.shape[0]
    need = lib.flowcuda_realnvp_nll_grad_workspace_bytes(n_layers, hidden, n)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty((need + 3) // 4, device=y.device, dtype=torch.float32)
    loss = torch.empty(1, device=y.device, dtype=torch.float32)
    grad = torch.empty_like(params)
    check(lib.flowcuda_realnvp_nll_grad(_ptr(params), n_layers, hidden, flip_mask, _ptr(y), n, _ptr(loss), _ptr(grad),
                                        float(lr), _ptr(params_out), _ptr(workspace),
                                        workspace.numel() * workspace.element_size(), _stream()), "realnvp_nll_grad")
    return loss, grad


# ---------------------------------------------------------------------------- row #5
@_on_tensor_device
def bernstein_constrain_thetas(theta_raw: torch.Tensor) -> torch.Tensor:
    t = _f32(theta_raw)
    _dev_check(t)
    order = t.shape[-1]
    out = torch.empty_like(t)
    check(_lib.load().flowcuda_bernstein_constrain_thetas(_ptr(t), _ptr(out), t.numel() // order, order, _stream()),
          "bernstein_constrain_thetas")
    return out


@_on_tensor_device
def bernstein_forward_and_log_det(theta: torch.Tensor, x: torch.Tensor, want_y: bool = True, want_ld: bool = True):
    """theta [M] (shared) or [*x.shape, M] (one row per element); x any shape."""
    theta, x = _f32(theta), _f32(x)
    _dev_check(theta, x)
    order = theta.shape[-1]
    if theta.dim() == 1:
        stride = 0
    else:
        if tuple(theta.shape[:-1]) != tuple(x.shape):
            theta = theta.expand(*x.shape, order).contiguous()
        stride = order
    y = torch.empty_like(x) if want_y else None
    ld = torch.empty_like(x) if want_ld else None
    check(_lib.load().flowcuda_bernstein_forward_and_log_det(_ptr(theta), order, stride, _ptr(x), _ptr(y), _ptr(ld),
                                                             x.numel(), _stream()), "bernstein_forward_and_log_det")
    return y, ld


@_on_tensor_device
def bernstein_forward_and_log_det_vjp(theta: torch.Tensor, x: torch.Tensor, ybar=None, ldbar=None,
                                      want_xbar: bool = True, want_thetabar: bool = True):
    """Reverse mode of bernstein_forward_and_log_det (csrc/bernstein_grad.cu): cotangents ybar / ldbar (shape of
    x, either may be None) -> (xbar, thetabar); thetabar is [M] for a shared theta, theta's shape for rows."""
    theta, x = _f32(theta), _f32(x)
    _dev_check(theta, x)
    order = theta.shape[-1]
    shared = theta.dim() == 1
    if not shared and tuple(theta.shape[:-1]) != tuple(x.shape):
        raise ValueError("bernstein vjp: per-element theta must be [*x.shape, M]")
    if ybar is not None:
        ybar = _f32(ybar.expand_as(x))
        _dev_check(ybar)
    if ldbar is not None:
        ldbar = _f32(ldbar.expand_as(x))
        _dev_check(ldbar)
    lib = _lib.load()
    stride = 0 if shared else order
    xbar = torch.empty_like(x) if want_xbar else None
    tbar = torch.empty_like(theta) if want_thetabar else None
    nbytes = lib.flowcuda_bernstein_vjp_workspace_bytes(order, stride, x.numel()) if want_thetabar else 0
    ws = torch.empty(max(nbytes, 16) // 4, device=x.device, dtype=torch.float32)
    check(lib.flowcuda_bernstein_forward_and_log_det_vjp(_ptr(theta), order, stride, _ptr(x), _ptr(ybar), _ptr(ldbar),
                                                         _ptr(xbar), _ptr(tbar), x.numel(), _ptr(ws), ws.numel() * 4,
                                                         _stream()), "bernstein_forward_and_log_det_vjp")
    return xbar, tbar


@_on_tensor_device
def bernstein_constrain_thetas_vjp(theta_raw: torch.Tensor, thetabar: torch.Tensor) -> torch.Tensor:
    t, tb = _f32(theta_raw), _f32(thetabar)
    _dev_check(t, tb)
    if t.shape != tb.shape:
        raise ValueError("bernstein_constrain_thetas_vjp: thetabar must have theta_raw's shape")
    order = t.shape[-1]
    out = torch.empty_like(t)
    check(_lib.load().flowcuda_bernstein_constrain_thetas_vjp(_ptr(t), _ptr(tb), _ptr(out), t.numel() // order, order,
                                                              _stream()), "bernstein_constrain_thetas_vjp")
    return out


@_on_tensor_device
def bernstein_inverse(theta: torch.Tensor, y: torch.Tensor, n_points: int = 200) -> torch.Tensor:
    """theta [E,M] (or [M] with E=1), y [N,E] -> x [E,N] (the reference's vmap output layout)."""
    theta, y = _f32(theta), _f32(y)
    _dev_check(theta, y)
    if theta.dim() == 1:
        theta = theta[None, :].contiguous()
    E, order = theta.shape
    if y.dim() != 2 or y.shape[1] != E:
        raise ValueError("vmap got inconsistent sizes for array axes to be mapped: y must be [N, %d]" % E)
    n = y.shape[0]
    x = torch.empty(E, n, device=y.device, dtype=torch.float32)
    ws = torch.empty(E * n_points, device=y.device, dtype=torch.float32)
    check(_lib.load().flowcuda_bernstein_inverse(_ptr(theta), order, E, n_points, _ptr(y), _ptr(x), n, _ptr(ws),
                                                 _stream()), "bernstein_inverse")
    return x


@_on_tensor_device
def bernstein_flow_log_prob(theta: torch.Tensor, scal: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    theta, scal, x = _f32(theta), _f32(scal), _f32(x)
    _dev_check(theta, scal, x)
    lp = torch.empty_like(x)
    check(_lib.load().flowcuda_bernstein_flow_log_prob(_ptr(theta), theta.shape[-1], _ptr(scal), _ptr(x), _ptr(lp),
                                                       x.numel(), _stream()), "bernstein_flow_log_prob")
    return lp


# ---------------------------------------------------------------------------- rows #2-#4
def make_spec(event_dim: int, num_layers: int, hidden: Sequence[int], transform: int, num_bins: int = 0,
              range_min: float = 0.0, range_max: float = 1.0, base: int = _lib.BASE_NORMAL,
              precision: int = _lib.PRECISION_FP32) -> CouplingSpec:
    if len(hidden) > _lib.MAX_HIDDEN_LAYERS:
        raise ValueError("at most %d hidden layers" % _lib.MAX_HIDDEN_LAYERS)
    sp = CouplingSpec()
    sp.event_dim, sp.num_layers, sp.num_hidden = event_dim, num_layers, len(hidden)
    for i, h in enumerate(hidden):
        sp.hidden[i] = h
    sp.transform, sp.num_bins = transform, num_bins
    sp.range_min, sp.range_max = range_min, range_max
    sp.base, sp.precision = base, precision
    return sp


def coupling_param_count(spec: CouplingSpec) -> int:
    n = _lib.load().flowcuda_coupling_param_count(C.byref(spec))
    if n < 0:
        check(-1, "coupling_param_count")
    return n


def _workspace(spec: CouplingSpec, n: int, device) -> Tuple[Optional[torch.Tensor], int]:
    need = _lib.load().flowcuda_coupling_workspace_bytes(C.byref(spec), n)
    if need == 0:
        return None, 0
    ws = torch.empty((need + 3) // 4, device=device, dtype=torch.float32)
    return ws, ws.numel() * 4


class PackedWeights:
    """A workspace whose head holds a flow's weights as tensor-core tile images (flowcuda_*_pack_weights), sized
    for batches up to `max_batch`.  Passing it to log_prob / sample skips the per-call packing pass; valid only
    while the parameters it was packed from are unchanged (the caller's promise — the library cannot see
    writes through raw pointers, e.g. its own optimiser kernels)."""

    def __init__(self, workspace: torch.Tensor, max_batch: int, kind: str):
        self.workspace, self.max_batch, self.kind = workspace, int(max_batch), kind

    @property
    def nbytes(self) -> int:
        return self.workspace.numel() * self.workspace.element_size()


@_on_tensor_device
def coupling_pack_weights(spec: CouplingSpec, params: torch.Tensor, max_batch: int) -> PackedWeights:
    _dev_check(params)
    lib = _lib.load()
    need = lib.flowcuda_coupling_workspace_bytes(C.byref(spec), int(max_batch))
    ws = torch.empty((need + 3) // 4, device=params.device, dtype=torch.float32)
    check(lib.flowcuda_coupling_pack_weights(C.byref(spec), _ptr(params), _ptr(ws), ws.numel() * 4, _stream()),
          "coupling_pack_weights")
    return PackedWeights(ws, max_batch, "coupling")


@_on_tensor_device
def made_pack_weights(spec: "_lib.MadeSpec", params: torch.Tensor, max_batch: int) -> PackedWeights:
    _dev_check(params)
    lib = _lib.load()
    need = lib.flowcuda_made_workspace_bytes(C.byref(spec), int(max_batch))
    ws = torch.empty((need + 3) // 4, device=params.device, dtype=torch.float32)
    check(lib.flowcuda_made_pack_weights(C.byref(spec), _ptr(params), _ptr(ws), ws.numel() * 4, _stream()),
          "made_pack_weights")
    return PackedWeights(ws, max_batch, "made")


def _use_packed(packed: Optional[PackedWeights], kind: str, n: int) -> bool:
    if packed is None:
        return False
    if packed.kind != kind:
        raise _lib.FlowCudaError(f"packed weights of a {packed.kind} flow given to a {kind} call")
    if n > packed.max_batch:
        raise _lib.FlowCudaError(f"batch {n} exceeds the packed workspace's max_batch {packed.max_batch}")
    return True


@_on_tensor_device
def coupling_log_prob(spec: CouplingSpec, params: torch.Tensor, x: torch.Tensor, want_z: bool = False,
                      want_log_det: bool = False, want_log_prob: bool = True, layers: Optional[Tuple[int, int]] = None,
                      with_base: bool = True, packed: Optional[PackedWeights] = None):
    """Data -> base direction (Transformed.log_prob).  Returns (log_prob, z, log_det) with None
    for outputs that were not requested."""
    x = _f32(x).reshape(x.shape[0], -1)
    _dev_check(params, x)
    n = x.shape[0]
    lb, le = layers if layers is not None else (0, spec.num_layers)
    lp = torch.empty(n, device=x.device, dtype=torch.float32) if want_log_prob else None
    z = torch.empty_like(x) if want_z else None
    ld = torch.empty(n, device=x.device, dtype=torch.float32) if want_log_det else None
    if _use_packed(packed, "coupling", n):
        check(_lib.load().flowcuda_coupling_log_prob_packed(C.byref(spec), _ptr(params), _ptr(x), _ptr(lp), _ptr(z), _ptr(ld),
                                                            n, lb, le, int(with_base), _ptr(packed.workspace), packed.nbytes,
                                                            _stream()), "coupling_log_prob_packed")
        return lp, z, ld
    ws, wsb = _workspace(spec, n, x.device)
    check(_lib.load().flowcuda_coupling_log_prob(C.byref(spec), _ptr(params), _ptr(x), _ptr(lp), _ptr(z), _ptr(ld), n,
                                                 lb, le, int(with_base), _ptr(ws), wsb, _stream()), "coupling_log_prob")
    return lp, z, ld


@_on_tensor_device
def coupling_sample(spec: CouplingSpec, params: torch.Tensor, z: torch.Tensor, want_log_det: bool = False,
                    layers: Optional[Tuple[int, int]] = None, packed: Optional[PackedWeights] = None):
    """Base -> data direction (Transformed.sample with the noise given).  Returns (x, log_det)."""
    z = _f32(z).reshape(z.shape[0], -1)
    _dev_check(params, z)
    n = z.shape[0]
    lb, le = layers if layers is not None else (0, spec.num_layers)
    x = torch.empty_like(z)
    ld = torch.empty(n, device=z.device, dtype=torch.float32) if want_log_det else None
    if _use_packed(packed, "coupling", n):
        check(_lib.load().flowcuda_coupling_sample_packed(C.byref(spec), _ptr(params), _ptr(z), _ptr(x), _ptr(ld), n, lb, le,
                                                          _ptr(packed.workspace), packed.nbytes, _stream()),
              "coupling_sample_packed")
        return x, ld
    ws, wsb = _workspace(spec, n, z.device)
    check(_lib.load().flowcuda_coupling_sample(C.byref(spec), _ptr(params), _ptr(z), _ptr(x), _ptr(ld), n, lb, le,
                                               _ptr(ws), wsb, _stream()), "coupling_sample")
    return x, ld


@_on_tensor_device
def coupling_nll_grad(spec: CouplingSpec, params: torch.Tensor, x: torch.Tensor, on_layer=None):
    """(loss[1], grad[param_count]) of -mean(log_prob(x)) — jax.value_and_grad(loss_fn) of the toy scripts.

    on_layer(layer, grad_slice, stream): optional host callback, called from inside the C call one or more times per
    flow layer (disjoint views of `grad` that together cover the layer) as soon as everything that writes the slice has
    been enqueued; `stream` is a
    torch.cuda.ExternalStream on which the slice is complete in stream order (flowcuda_coupling_nll_grad_layers:
    the hook of data-parallel training — reduce a layer's gradient while the others are still being computed)."""
    x = _f32(x).reshape(x.shape[0], -1)
    _dev_check(params, x)
    lib = _lib.load()
    n = x.shape[0]
    need = lib.flowcuda_coupling_nll_grad_workspace_bytes(C.byref(spec), n)
    ws = torch.empty(max(1, (need + 3) // 4), device=x.device, dtype=torch.float32)
    loss = torch.empty(1, device=x.device, dtype=torch.float32)
    grad = torch.empty_like(params)
    if on_layer is None:
        check(lib.flowcuda_coupling_nll_grad(C.byref(spec), _ptr(params), _ptr(x), n, _ptr(loss), _ptr(grad), _ptr(ws),
                                             ws.numel() * 4, _stream()), "coupling_nll_grad")
        return loss, grad
    failure = []

    def _cb(user, layer, offset, count, stream):  # no exception may cross the C frames
        try:
            # a NULL handle is the legacy default stream = torch's default stream; torch.cuda.ExternalStream(0) does NOT
            # denote it (its work was seen to run unordered), so it is mapped by hand
            st = torch.cuda.ExternalStream(int(stream), device=x.device) if stream else torch.cuda.default_stream(x.device)
            on_layer(int(layer), grad[offset:offset + count], st)
        except BaseException as e:  # noqa: BLE001
            failure.append(e)

    cb = _lib.LAYER_GRAD_FN(_cb)
    rc = lib.flowcuda_coupling_nll_grad_layers(C.byref(spec), _ptr(params), _ptr(x), n, _ptr(loss), _ptr(grad), _ptr(ws),
                                               ws.numel() * 4, _stream(), C.cast(cb, C.c_void_p), None)
    if failure:
        raise failure[0]
    check(rc, "coupling_nll_grad_layers")
    return loss, grad


# ---------------------------------------------------------------------------- row #8
def make_made_spec(event_dim: int, num_layers: int, hidden: Sequence[int], base: int = _lib.BASE_NORMAL,
                   precision: int = _lib.PRECISION_FP32) -> _lib.MadeSpec:
    """`precision` selects the engine of made_nll_grad's GEMMs: FP32 CUDA cores or tcgen05 bf16x3."""
    if len(hidden) > _lib.MAX_HIDDEN_LAYERS:
        raise ValueError("at most %d hidden layers" % _lib.MAX_HIDDEN_LAYERS)
    sp = _lib.MadeSpec()
    sp.event_dim, sp.num_layers, sp.num_hidden = event_dim, num_layers, len(hidden)
    for i, h in enumerate(hidden):
        sp.hidden[i] = h
    sp.base = base
    sp.precision = precision
    return sp


def made_param_count(spec: _lib.MadeSpec) -> int:
    n = _lib.load().flowcuda_made_param_count(C.byref(spec))
    if n < 0:
        check(-1, "made_param_count")
    return n


@_on_tensor_device
def made_log_prob(spec: _lib.MadeSpec, params: torch.Tensor, x: torch.Tensor, want_u: bool = False,
                  want_log_det: bool = False, want_log_prob: bool = True, with_base: bool = True,
                  workspace: Optional[torch.Tensor] = None, packed: Optional[PackedWeights] = None):
    """MAF density pass (data -> base).  Returns (log_prob, u, log_det), None where not requested."""
    x = _f32(x).reshape(x.shape[0], spec.event_dim)
    _dev_check(params, x)
    lib = _lib.load()
    n = x.shape[0]
    lp = torch.empty(n, device=x.device, dtype=torch.float32) if want_log_prob else None
    u = torch.empty_like(x) if want_u else None
    ld = torch.empty(n, device=x.device, dtype=torch.float32) if want_log_det else None
    if _use_packed(packed, "made", n):
        check(lib.flowcuda_made_log_prob_packed(C.byref(spec), _ptr(params), _ptr(x), _ptr(lp), _ptr(u), _ptr(ld), n,
                                                int(with_base), _ptr(packed.workspace), packed.nbytes, _stream()),
              "made_log_prob_packed")
        return lp, u, ld
    need = lib.flowcuda_made_workspace_bytes(C.byref(spec), n)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty((need + 3) // 4, device=x.device, dtype=torch.float32)
    check(lib.flowcuda_made_log_prob(C.byref(spec), _ptr(params), _ptr(x), _ptr(lp), _ptr(u), _ptr(ld), n,
                                     int(with_base), _ptr(workspace), workspace.numel() * workspace.element_size(),
                                     _stream()), "made_log_prob")
    return lp, u, ld


@_on_tensor_device
def made_log_prob_fp32(spec: _lib.MadeSpec, params: torch.Tensor, x: torch.Tensor, want_u: bool = False):
    """The MAF density pass with every product in FP32 on the CUDA cores (flowcuda_made_log_prob_fp32): the
    full-precision partner of `made_log_prob`.  Returns (log_prob, u or None)."""
    x = _f32(x).reshape(x.shape[0], spec.event_dim)
    _dev_check(params, x)
    lib = _lib.load()
    n = x.shape[0]
    lp = torch.empty(n, device=x.device, dtype=torch.float32)
    u = torch.empty_like(x) if want_u else None
    need = lib.flowcuda_made_log_prob_fp32_workspace_bytes(C.byref(spec), n)
    ws = torch.empty(max(1, (need + 3) // 4), device=x.device, dtype=torch.float32)
    check(lib.flowcuda_made_log_prob_fp32(C.byref(spec), _ptr(params), _ptr(x), _ptr(lp), _ptr(u), n, _ptr(ws), ws.numel() * 4,
                                          _stream()), "made_log_prob_fp32")
    return lp, u


@_on_tensor_device
def made_sample(spec: _lib.MadeSpec, params: torch.Tensor, z: torch.Tensor) -> torch.Tensor:
    """MAF sampling with the base noise given (base -> data), sequential over dimensions."""
    z = _f32(z).reshape(z.shape[0], spec.event_dim)
    _dev_check(params, z)
    x = torch.empty_like(z)
    check(_lib.load().flowcuda_made_sample(C.byref(spec), _ptr(params), _ptr(z), _ptr(x), z.shape[0], _stream()),
          "made_sample")
    return x


@_on_tensor_device
def made_nll_grad(spec: _lib.MadeSpec, params: torch.Tensor, x: torch.Tensor, workspace: Optional[torch.Tensor] = None):
    """(loss[1], grad[param_count]) of -mean(log_prob(x)) for the MADE flow — jax.value_and_grad(loss_fn)."""
    x = _f32(x).reshape(x.shape[0], spec.event_dim)
    _dev_check(params, x)
    lib = _lib.load()
    n = x.shape[0]
    need = lib.flowcuda_made_nll_grad_workspace_bytes(C.byref(spec), n)
    if need == 0:
        raise _lib.FlowCudaError("made_nll_grad: unsupported spec (needs two hidden layers)")
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty((need + 3) // 4, device=x.device, dtype=torch.float32)
    loss = torch.empty(1, device=x.device, dtype=torch.float32)
    grad = torch.empty_like(params)
    check(lib.flowcuda_made_nll_grad(C.byref(spec), _ptr(params), _ptr(x), n, _ptr(loss), _ptr(grad), _ptr(workspace),
                                     workspace.numel() * workspace.element_size(), _stream()), "made_nll_grad")
    return loss, grad


# ---------------------------------------------------------------------------- rows #7, #9
@_on_tensor_device
def sum_scaled(v: torch.Tensor, scale: float) -> torch.Tensor:
    v = _f32(v).reshape(-1)
    _dev_check(v)
    out = torch.empty(1, device=v.device, dtype=torch.float32)
    ws = torch.empty(1024, device=v.device, dtype=torch.float64)
    check(_lib.load().flowcuda_sum_scaled(_ptr(v), v.numel(), float(scale), _ptr(out), _ptr(ws), _stream()), "sum_scaled")
    return out


@_on_tensor_device
def nll(log_prob: torch.Tensor) -> torch.Tensor:
    """-jnp.mean(log_prob) (real_nvp.py:293)."""
    return sum_scaled(log_prob, -1.0 / max(1, log_prob.numel()))


@_on_tensor_device
def sgd_step(p: torch.Tensor, g: torch.Tensor, lr: float, grad_scale: float = 1.0) -> None:
    _dev_check(p, g)
    check(_lib.load().flowcuda_sgd_step(_ptr(p), _ptr(g), p.numel(), float(lr), float(grad_scale), _stream()), "sgd_step")


@_on_tensor_device
def adam_step(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int, lr: float,
              b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8, grad_scale: float = 1.0) -> None:
    _dev_check(p, g, m, v)
    check(_lib.load().flowcuda_adam_step(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), int(step), float(lr), float(b1),
                                         float(b2), float(eps), float(grad_scale), _stream()), "adam_step")


# ---------------------------------------------------------------------------- data step (§8f-2, §8f-3)
DENSITY_KINDS = {"banana": 0, "energy_1": 1, "energy_2": 2, "energy_3": 3, "energy_4": 4, "gaussian_blobs": 5}


def density_sample(kind, seed: int, n: int, offset: int = 0, num_blobs: int = 8, r: float = 2.0,
                   device=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x[n,2] = samples offset..offset+n of the density's Philox stream (csrc/densities.cu)."""
    k = DENSITY_KINDS[kind] if isinstance(kind, str) else int(kind)
    if out is None:
        out = torch.empty(int(n), 2, device=device or "cuda", dtype=torch.float32)
    _dev_check(out)
    if out.numel() != 2 * int(n):
        raise _lib.FlowCudaError("density_sample: out must hold n*2 floats")
    with torch.cuda.device(out.device):
        check(_lib.load().flowcuda_density_sample(k, int(seed) & (2 ** 64 - 1), int(offset), int(num_blobs), float(r),
                                                  _ptr(out), int(n), _stream()), "density_sample")
    return out


@_on_tensor_device
def dequantize_u8(bytes_: torch.Tensor, seed: int, offset: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(uint8 + U[0,1)) / 256 on the device (prepare_data, src/maf_mnist.py:101-106)."""
    if not bytes_.is_cuda or bytes_.dtype != torch.uint8 or not bytes_.is_contiguous():
        raise _lib.FlowCudaError("dequantize_u8 needs a contiguous CUDA uint8 tensor (there is no CPU path)")
    if out is None:
        out = torch.empty(bytes_.shape, device=bytes_.device, dtype=torch.float32)
    _dev_check(out)
    if out.numel() != bytes_.numel():
        raise _lib.FlowCudaError("dequantize_u8: out must have as many elements as the input")
    with torch.cuda.device(out.device):
        check(_lib.load().flowcuda_dequantize_u8(_ptr(bytes_), int(seed) & (2 ** 64 - 1), int(offset), _ptr(out),
                                                 bytes_.numel(), _stream()), "dequantize_u8")
    return out


def base_noise(shape, seed: int, uniform: bool = False, offset: int = 0, device=None) -> torch.Tensor:
    """N(0,1) (or U[0,1)) base noise for Transformed.sample(seed=...), generated on the device."""
    out = torch.empty(tuple(shape), device=device or "cuda", dtype=torch.float32)
    _dev_check(out)
    with torch.cuda.device(out.device):
        check(_lib.load().flowcuda_base_noise(1 if uniform else 0, int(seed) & (2 ** 64 - 1), int(offset), _ptr(out),
                                              out.numel(), _stream()), "base_noise")
    return out


@_on_tensor_device
def rotate2d(x: torch.Tensor, cos_theta: float, sin_theta: float, inverse: bool = False) -> torch.Tensor:
    """y = x @ R (R = [[c,-s],[s,c]]^T) or x @ R^T, x[...,2] (src/rotation_flow.py:23-27)."""
    if x.shape[-1] != 2:
        raise _lib.FlowCudaError(f"rotate2d needs a trailing dimension of 2, got {tuple(x.shape)}")
    xf = _f32(x)
    _dev_check(xf)
    y = torch.empty_like(xf)
    check(_lib.load().flowcuda_rotate2d(_ptr(xf), float(cos_theta), float(sin_theta), 1 if inverse else 0, _ptr(y),
                                        xf.numel() // 2, _stream()), "rotate2d")
    return y


# ---------------------------------------------------------------------------- elementwise bijectors / base densities
def _rows_inner(x_shape, p_shape):
    """Broadcast pattern of parameters of shape `p_shape` against data of shape `x_shape` as (rows, inner) of the
    kernels' index rule row = (i / inner) % rows; None when the pattern is neither leading- nor trailing-broadcast."""
    xs, ps = tuple(x_shape), tuple(p_shape)
    pn = 1
    for d in ps:
        pn *= d
    if pn == 1:
        return 1, 1
    ps = (1,) * (len(xs) - len(ps)) + ps if len(ps) <= len(xs) else None
    if ps is None:
        return None
    # strip leading 1s (leading broadcast: rows repeat) and trailing 1s (trailing broadcast: `inner` elements share a row)
    lead = 0
    while lead < len(ps) and ps[lead] == 1:
        lead += 1
    trail = len(ps)
    while trail > lead and ps[trail - 1] == 1:
        trail -= 1
    if ps[lead:trail] != xs[lead:trail]:
        return None
    inner = 1
    for d in xs[trail:]:
        inner *= d
    return pn, inner


def _bcast_params(x, *ps):
    """(rows, inner, params...) with the parameters contiguous float32 on x's device; falls back to materialising the
    broadcast when the pattern is irregular."""
    ps = [None if p is None else _f32(torch.as_tensor(p)).to(x.device) for p in ps]
    live = [p for p in ps if p is not None]
    if not live:
        return 1, 1, ps
    shape = torch.broadcast_shapes(*[p.shape for p in live])
    ri = _rows_inner(x.shape, shape)
    if ri is None:
        shape = torch.broadcast_shapes(shape, x.shape)
        ri = (int(x.numel()), 1) if tuple(shape) == tuple(x.shape) else None
        if ri is None:
            raise ValueError(f"parameters of shape {tuple(shape)} do not broadcast against data of shape {tuple(x.shape)}")
    ps = [None if p is None else p.expand(shape).contiguous() for p in ps]
    return ri[0], ri[1], ps


@_on_tensor_device
def affine_elementwise(x, shift=None, scale=None, log_scale=None, inverse=False, want_y=True, want_ld=True):
    """distrax.ScalarAffine / tfb.Scale / tfb.Shift: (y, log_det) elementwise, parameters broadcast."""
    x = _f32(x)
    _dev_check(x)
    sc, is_log = (log_scale, 1) if log_scale is not None else (scale, 0)
    rows, inner, (sh, sc) = _bcast_params(x, shift, sc)
    y = torch.empty_like(x) if want_y else None
    ld = torch.empty_like(x) if want_ld else None
    check(_lib.load().flowcuda_affine_elementwise(_ptr(x), _ptr(sh), _ptr(sc), is_log, int(inverse), _ptr(y), _ptr(ld), x.numel(),
                                                  rows, inner, _stream()), "affine_elementwise")
    return y, ld


def add_elementwise(a, b):
    """a + b on the device (log-det bookkeeping of Chain / Transformed): the affine kernel with shift = b."""
    a = _f32(a)
    b = _f32(torch.as_tensor(b)).to(a.device)
    if b.shape != a.shape:
        b = b.expand(a.shape).contiguous()
    return affine_elementwise(a, shift=b, want_ld=False)[0]


@_on_tensor_device
def softclip_elementwise(x, low, high, hinge_softness, inverse=False):
    x = _f32(x)
    _dev_check(x)
    y, ld = torch.empty_like(x), torch.empty_like(x)
    check(_lib.load().flowcuda_softclip_elementwise(_ptr(x), float(low), float(high), float(hinge_softness), int(inverse), _ptr(y),
                                                    _ptr(ld), x.numel(), _stream()), "softclip_elementwise")
    return y, ld


@_on_tensor_device
def rqs_elementwise(x, params, range_min, range_max, inverse=False):
    """distrax.RationalQuadraticSpline(params[..., 3K+1], range_min, range_max) applied to x[...]."""
    x = _f32(x)
    _dev_check(x)
    params = _f32(torch.as_tensor(params)).to(x.device)
    P = params.shape[-1]
    ri = _rows_inner(x.shape, params.shape[:-1])
    if ri is None:
        shape = torch.broadcast_shapes(params.shape[:-1], x.shape)
        if tuple(shape) != tuple(x.shape):
            raise ValueError(f"spline parameters {tuple(params.shape)} do not broadcast against data {tuple(x.shape)}")
        params = params.expand(*shape, P).contiguous()
        ri = (int(x.numel()), 1)
    y, ld = torch.empty_like(x), torch.empty_like(x)
    check(_lib.load().flowcuda_rqs_elementwise(_ptr(x), _ptr(params), (P - 1) // 3, float(range_min), float(range_max), int(inverse),
                                               _ptr(y), _ptr(ld), x.numel(), ri[0], ri[1], _stream()), "rqs_elementwise")
    return y, ld


@_on_tensor_device
def base_log_prob(x, a=None, b=None, uniform=False):
    """Normal(loc=a, scale=b).log_prob(x) / Uniform(low=a, high=b).log_prob(x), elementwise."""
    x = _f32(x)
    _dev_check(x)
    rows, inner, (a, b) = _bcast_params(x, a, b)
    out = torch.empty_like(x)
    check(_lib.load().flowcuda_base_log_prob(_ptr(x), _ptr(a), _ptr(b), int(uniform), _ptr(out), x.numel(), rows, inner, _stream()),
          "base_log_prob")
    return out


@_on_tensor_device
def sum_last(x, ndims):
    """Sum over the last `ndims` axes in a fixed order (distrax.Independent / Block)."""
    if ndims == 0:
        return x
    x = _f32(x)
    _dev_check(x)
    batch = x.shape[: x.dim() - ndims]
    cols = 1
    for d in x.shape[x.dim() - ndims:]:
        cols *= d
    rows = x.numel() // max(cols, 1)
    out = torch.empty(rows, device=x.device, dtype=torch.float32)
    check(_lib.load().flowcuda_sum_last(_ptr(x), _ptr(out), rows, cols, _stream()), "sum_last")
    return out.reshape(batch)


# ---------------------------------------------------------------------------- diagnostics
@_on_tensor_device
def selftest_umma(a: torch.Tensor, w: torch.Tensor, mode: int, want_unaligned: bool = False):
    """d[128,n] = a[128,k] @ w[k,n] through the tcgen05 primitives (mode 0 tf32x3, 1 bf16x3, 2 tf32, 3 bf16)."""
    a, w = _f32(a), _f32(w)
    _dev_check(a, w)
    k, n = w.shape
    lib = _lib.load_selftest()
    need = lib.flowcuda_selftest_umma_workspace_bytes(n, k)
    ws = torch.empty((need + 3) // 4, device=a.device, dtype=torch.float32)
    d = torch.zeros(128, n, device=a.device, dtype=torch.float32)
    d2 = torch.zeros(128, n, device=a.device, dtype=torch.float32) if want_unaligned else None
    rc = lib.flowcuda_selftest_umma(_ptr(a), _ptr(w), _ptr(d), _ptr(d2), n, k, mode, _ptr(ws), ws.numel() * 4, _stream())
    if rc != 0:
        raise _lib.FlowCudaError(f"selftest_umma failed with code {rc}: {lib.flowcuda_selftest_last_error().decode()}")
    return (d, d2) if want_unaligned else d
