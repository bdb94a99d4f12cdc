"""torch.autograd.Function wrappers over the C ABI for steps 3-5 (rotation head, SE(3) update,
point-matching losses).  Tensors that arrive on the CPU are staged through the current CUDA
device (the kernels are the only implementation; there is no CPU fallback) and results are
returned on the input's device, as the reference's functions would."""
import torch

from . import _lib
from ._lib import RAC_POSE_SAVED_FLOATS, RAC_ROT


def _cu(t, dtype=torch.float32):
    if t.is_cuda and t.dtype == dtype and t.is_contiguous():     # the common case: nothing to stage or convert
        return t.detach()
    dev = _lib.require_cuda()
    return t.detach().to(device=dev, dtype=dtype).contiguous()


_counters = {}


def _counter(dev):
    c = _counters.get(dev.index)
    if c is None:
        c = torch.zeros(1, dtype=torch.int32, device=dev)
        _counters[dev.index] = c
    return c


class PoseUpdateFn(torch.autograd.Function):
    """ref: rotation_representation.py:131-184 (calculate_T_CO_pred)."""

    @staticmethod
    def forward(ctx, model_output, T_CO_init, rot_id, fx, fy):
        L = _lib.lib()
        y = _cu(model_output)
        t64 = T_CO_init.dtype == torch.float64
        T_in = _cu(T_CO_init, torch.float64 if t64 else torch.float32)
        B = y.shape[0]
        T_out = torch.empty((B, 4, 4), dtype=torch.float32, device=y.device)
        saved = torch.empty((B, RAC_POSE_SAVED_FLOATS), dtype=torch.float32, device=y.device)
        _lib.check(L.rac_pose_update_fwd(_lib.ptr(y), rot_id, _lib.ptr(T_in), int(t64), float(fx), float(fy), B,
                                         _lib.ptr(T_out), _lib.ptr(saved), _lib.stream_ptr()))
        ctx.save_for_backward(saved)
        ctx.rot_id = rot_id
        ctx.out_dev = model_output.device
        ctx.D = y.shape[1]
        return T_out.to(model_output.device)

    @staticmethod
    def backward(ctx, gT):
        (saved,) = ctx.saved_tensors
        L = _lib.lib()
        g = _cu(gT)
        B = g.shape[0]
        gy = torch.empty((B, ctx.D), dtype=torch.float32, device=g.device)
        _lib.check(L.rac_pose_update_bwd(_lib.ptr(g), _lib.ptr(saved), ctx.rot_id, B, _lib.ptr(gy), _lib.stream_ptr()))
        return gy.to(ctx.out_dev), None, None, None, None


class RotationFn(torch.autograd.Function):
    """ref: rotation_representation.py:4-19 (6D), :22-39 (SVD), :41-72 (exp map)."""

    @staticmethod
    def forward(ctx, x, rot_id, eps):
        L = _lib.lib()
        xin = _cu(x)
        B = xin.shape[0]
        R = torch.empty((B, 3, 3), dtype=torch.float32, device=xin.device)
        saved = torch.empty((B, RAC_POSE_SAVED_FLOATS), dtype=torch.float32, device=xin.device)
        _lib.check(L.rac_rotation_fwd(_lib.ptr(xin), rot_id, float(eps), B, _lib.ptr(R), _lib.ptr(saved), _lib.stream_ptr()))
        ctx.save_for_backward(saved)
        ctx.rot_id = rot_id
        ctx.in_shape = tuple(xin.shape)
        ctx.out_dev = x.device
        ctx.in_dtype = x.dtype
        return R.to(device=x.device, dtype=x.dtype)

    @staticmethod
    def backward(ctx, gR):
        (saved,) = ctx.saved_tensors
        L = _lib.lib()
        g = _cu(gR)
        B = g.shape[0]
        gin = torch.empty(ctx.in_shape, dtype=torch.float32, device=g.device)
        _lib.check(L.rac_rotation_bwd(_lib.ptr(g), _lib.ptr(saved), ctx.rot_id, B, _lib.ptr(gin), _lib.stream_ptr()))
        return gin.to(device=ctx.out_dev, dtype=ctx.in_dtype), None, None


class AddLossFn(torch.autograd.Function):
    """ref: loss.py:5-33,81-122.  T_a/T_b in the reference's argument order (see include/rac.h)."""

    @staticmethod
    def forward(ctx, T_a, T_b, T_prev, points, mode, batch_mean):
        L = _lib.lib()
        a, b, pts = _cu(T_a), _cu(T_b), _cu(points)
        prev = _cu(T_prev) if T_prev is not None else None
        B, N = pts.shape[0], pts.shape[1]
        dev = a.device
        per = torch.empty(B, dtype=torch.float32, device=dev)
        mean = torch.empty(1, dtype=torch.float32, device=dev)
        need_a, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        ga = torch.empty((B, 4, 4), dtype=torch.float32, device=dev) if need_a else None
        gb = torch.empty((B, 4, 4), dtype=torch.float32, device=dev) if need_b else None
        _lib.check(L.rac_add_loss_fwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(prev), _lib.ptr(pts), B, N, mode,
                                      _lib.ptr(per), _lib.ptr(mean), _lib.ptr(ga), _lib.ptr(gb),
                                      _lib.ptr(_counter(dev)), _lib.stream_ptr()))
        ctx.save_for_backward(*(t for t in (ga, gb) if t is not None))
        ctx.has = (need_a, need_b)
        ctx.batch_mean = batch_mean
        ctx.B = B
        ctx.devs = (T_a.device, T_b.device)
        out = mean[0] if batch_mean else per
        return out.to(T_b.device)

    @staticmethod
    def backward(ctx, gout):
        L = _lib.lib()
        saved = list(ctx.saved_tensors)
        B = ctx.B
        g = _cu(gout)
        w = (g.reshape(1).expand(B) / B).contiguous() if ctx.batch_mean else g.reshape(B).contiguous()
        res = []
        for i, has in enumerate(ctx.has):
            if not has:
                res.append(None)
                continue
            G = saved.pop(0)
            gT = torch.empty_like(G)
            _lib.check(L.rac_add_loss_bwd(_lib.ptr(G), _lib.ptr(w), B, _lib.ptr(gT), _lib.stream_ptr()))
            res.append(gT.to(ctx.devs[i]))
        return res[0], res[1], None, None, None, None


class TransformPtsFn(torch.autograd.Function):
    """ref: loss.py:60-78 (transform_pts); T (B,M,4,4) or (B,4,4)."""

    @staticmethod
    def forward(ctx, T, pts):
        L = _lib.lib()
        Tc, p = _cu(T), _cu(pts)
        B, N = p.shape[0], p.shape[1]
        M = Tc.shape[1] if Tc.dim() == 4 else 1
        out = torch.empty((B, M, N, 3), dtype=torch.float32, device=p.device)
        _lib.check(L.rac_transform_pts_fwd(_lib.ptr(Tc), _lib.ptr(p), B, M, N, _lib.ptr(out), _lib.stream_ptr()))
        ctx.save_for_backward(Tc, p)
        ctx.dims = (B, M, N, T.dim())
        ctx.devs = (T.device, pts.device)
        if T.dim() == 3:
            out = out.reshape(B, N, 3)
        return out.to(pts.device)

    @staticmethod
    def backward(ctx, gout):
        L = _lib.lib()
        Tc, p = ctx.saved_tensors
        B, M, N, tdim = ctx.dims
        g = _cu(gout)
        gT = torch.empty((B, M, 4, 4), dtype=torch.float32, device=g.device) if ctx.needs_input_grad[0] else None
        gp = torch.empty((B, N, 3), dtype=torch.float32, device=g.device) if ctx.needs_input_grad[1] else None
        _lib.check(L.rac_transform_pts_bwd(_lib.ptr(g), _lib.ptr(Tc), _lib.ptr(p), B, M, N, _lib.ptr(gT), _lib.ptr(gp),
                                           _lib.stream_ptr()))
        if gT is not None:
            gT = (gT.reshape(B, 4, 4) if tdim == 3 else gT).to(ctx.devs[0])
        if gp is not None:
            gp = gp.to(ctx.devs[1])
        return gT, gp


def rot_id(rot_repr):
    if rot_repr not in RAC_ROT:
        assert False   # ref: rotation_representation.py:157
    return RAC_ROT[rot_repr]


_LOSS_SLOT = {   # loss name -> (mode, pred_is_a): which argument slot of the loss the NEW pose takes in the drivers' calls
    "add_l1": (_lib.LOSS_ADD_L1, 0),                     # compute_ADD_L1_loss(T_gt, T_new)            train_model.py:246
    "add_l2": (_lib.LOSS_ADD_L2, 0),                     # compute_ADD_L2_loss(T_gt, T_new)
    "add_l1_disentangled": (_lib.LOSS_DISENTANGLED, 0),  # compute_disentangled_ADD_L1_loss(T_gt, T_new)  :280 (swapped, Q7)
    "add_l1_disentl_scaled": (_lib.LOSS_DISENT_SCALED, 1),   # compute_scaled_disentl_ADD_L1_loss(T_prev, T_new, T_gt)  :283
}


def pose_loss_step(y, T_in, rot_repr, fx, fy, T_gt, pts, loss_name, T_prev=None, out=None, feedback=None):
    """calculate_T_CO_pred + loss + batch mean + backward to y in ONE launch (rac_pose_loss_step), for callers that own
    their loop.  y (B,D), T_in (B,4,4) f32/f64, T_gt (B,4,4), pts (B,N,3), all CUDA.  -> (T_new, loss (1,), grad_y,
    per_sample); `out` may hand in preallocated (T_new, loss, grad_y, per_sample) tensors (CUDA-graph capture);
    `feedback` (B,4,4) f32 also receives T_new and may be T_in / T_prev themselves (the loop's T_CO_pred = T_CO_pred_new)."""
    L = _lib.lib()
    mode, pred_is_a = _LOSS_SLOT[loss_name]
    yy = _cu(y)
    t64 = T_in.dtype == torch.float64
    Ti = _cu(T_in, torch.float64 if t64 else torch.float32)
    Tg, P = _cu(T_gt), _cu(pts)
    Tp = _cu(T_prev) if T_prev is not None else None
    B, N = P.shape[0], P.shape[1]
    assert yy.shape[0] == B and tuple(Ti.shape) == (B, 4, 4) and tuple(Tg.shape) == (B, 4, 4) and P.shape[2] == 3
    dev = yy.device
    if out is None:
        out = (torch.empty((B, 4, 4), dtype=torch.float32, device=dev), torch.empty(1, dtype=torch.float32, device=dev),
               torch.empty_like(yy), torch.empty(B, dtype=torch.float32, device=dev))
    T_new, loss, grad_y, per = out
    _lib.check(L.rac_pose_loss_step(_lib.ptr(yy), RAC_ROT[rot_repr], _lib.ptr(Ti), int(t64), float(fx), float(fy), _lib.ptr(Tg),
                                    _lib.ptr(Tp), _lib.ptr(P), B, N, mode, pred_is_a, _lib.ptr(T_new), _lib.ptr(per), _lib.ptr(loss),
                                    _lib.ptr(grad_y), _lib.ptr(feedback), _lib.ptr(_counter(dev)), _lib.stream_ptr()))
    return T_new, loss, grad_y, per
