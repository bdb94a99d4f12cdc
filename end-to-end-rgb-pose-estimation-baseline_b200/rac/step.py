"""RenderCompareStep — one refinement iteration of the hot path captured in a CUDA graph.

The reference's inner loop (train_model.py:240-250) launches ~60 tiny kernels plus a Python loop per
iteration; here one iteration is memset + 4 render kernels + pose-update fwd + loss fwd + loss bwd + pose-update
bwd, and at batch 64 the GPU finishes them faster than Python can enqueue them.  The step therefore owns static
device buffers and replays a captured graph:

    step = RenderCompareStep(mesh, B, (W, H), rot_repr="6D", loss="add_l1_disentangled")
    step.T_pred.copy_(...); step.K...; step.gt.copy_(...); step.y.copy_(cnn_output); step.T_gt...; step.pts...
    step.run()            # -> step.x (B,C,H,W), step.T_new (B,4,4), step.loss (1,), step.grad_y (B,D)

`step.x` is what the CNN consumes; `step.grad_y` is dLoss/d(cnn output), to be fed to `cnn_out.backward(grad_y)`.
A caller that wants the CNN between render and pose update uses run_render() / run_pose_loss() (two graphs).

Three replayable forms of the iteration:
    run_chain()  the loop's true dependency chain (train_model.py:240-250): render -> pose update + loss + backward ->
                 T_pred <- T_new, strictly in that order on one stream — in the real loop y = CNN(x), so the pose / loss
                 kernels can never start before the render of the same iteration has finished, and the next render
                 needs the updated pose.  This is the figure bench.py reports as `value`;
    run()        render and pose/loss as two parallel graph branches (legal only while y does not depend on x: an upper
                 bound, reported as `value_overlapped`); T_pred is left alone;
    run_render() / run_pose_loss()   the two halves, for a caller that runs its CNN in between.
Kernels only — the captured work is exactly the C-ABI calls of include/rac.h (no CPU fallback).
"""
import torch

from . import _lib
from .render import render_concat_device

_LOSSES = ("add_l1", "add_l2", "add_l1_disentangled", "add_l1_disentl_scaled")


class RenderCompareStep:
    def __init__(self, meshes, B, img_size, rot_repr="6D", loss="add_l1_disentangled", n_points=1000,
                 use_norm_depth=False, fx=None, fy=None, device=None, fused_pose_loss=True):
        import loss as L                       # the shadow modules (reference signatures)
        import rotation_representation as RR
        assert loss in _LOSSES, loss
        self.dev = device or _lib.require_cuda()
        W, H = (img_size, img_size) if isinstance(img_size, int) else (int(img_size[0]), int(img_size[1]))
        self.B, self.W, self.H, self.rot_repr, self.loss_name = B, W, H, rot_repr, loss
        self.meshes, self.use_norm_depth = meshes, bool(use_norm_depth)
        D = {"6D": 9, "SVD": 12, "3D": 6}[rot_repr]
        C = 7 if use_norm_depth else 6
        f32 = dict(dtype=torch.float32, device=self.dev)
        # static inputs
        self.T_pred = torch.eye(4, **f32).repeat(B, 1, 1)
        self.T_prev = torch.eye(4, **f32).repeat(B, 1, 1)      # only read by the scaled loss
        self.K = torch.zeros((B, 3, 3), **f32)
        self.gt = torch.zeros((B, H, W, 3), **f32)
        self.y = torch.zeros((B, D), **f32)
        self.T_gt = torch.eye(4, **f32).repeat(B, 1, 1)
        self.pts = torch.zeros((B, n_points, 3), **f32)
        # static outputs
        self.x = torch.empty((B, C, H, W), **f32)
        self.T_new = torch.empty((B, 4, 4), **f32)
        self.loss = torch.zeros(1, **f32)
        self.grad_y = torch.zeros((B, D), **f32)
        self._fx, self._fy = fx, fy
        self._L, self._RR = L, RR
        # the post-CNN half as ONE launch (rac_pose_loss_step) instead of the four kernels behind the reference-named
        # functions; False keeps the reference-named path (what the drivers call) for parity tests
        self.fused_pose_loss = bool(fused_pose_loss)
        self._per_sample = torch.empty(B, **f32)
        self._g_all = self._g_render = self._g_pose = self._g_chain = None

    # ---- the eager bodies (also what gets captured) ----
    def _render(self):
        render_concat_device(self.T_pred, self.meshes, self.K, (self.W, self.H), self.gt,
                             use_norm_depth=self.use_norm_depth, out=self.x)

    def _pose_loss(self, feed_back=False):
        """feed_back (chain mode): T_pred <- T_new inside the same launch; the scaled loss then reads the iteration's
        input pose from T_pred itself (train_model.py:283 passes T_CO_pred), so T_prev needs no copy either."""
        if self.fused_pose_loss:
            from .ops import pose_loss_step
            K0 = self._Ks_host[0]
            scaled = self.loss_name == "add_l1_disentl_scaled"
            pose_loss_step(self.y, self.T_pred, self.rot_repr, float(K0[0, 0]), float(K0[1, 1]), self.T_gt, self.pts, self.loss_name,
                           T_prev=(self.T_pred if feed_back else self.T_prev) if scaled else None,
                           out=(self.T_new, self.loss, self.grad_y, self._per_sample), feedback=self.T_pred if feed_back else None)
            if feed_back and scaled:
                self.T_prev.copy_(self.T_new)      # keep the public T_prev buffer coherent for a caller that reads it
            return
        L, RR = self._L, self._RR
        y = self.y.detach().requires_grad_(True)
        Ks = self._Ks_host
        T_new = RR.calculate_T_CO_pred(y, self.T_pred, self.rot_repr, Ks)
        if self.loss_name == "add_l1":
            l = L.compute_ADD_L1_loss(self.T_gt, T_new, self.pts)
        elif self.loss_name == "add_l2":
            l = L.compute_ADD_L2_loss(self.T_gt, T_new, self.pts)
        elif self.loss_name == "add_l1_disentangled":
            l = L.compute_disentangled_ADD_L1_loss(self.T_gt, T_new, self.pts)      # train_model.py:280 order (Q7)
        else:
            l = L.compute_scaled_disentl_ADD_L1_loss(self.T_prev, T_new, self.T_gt, self.pts)
        (gy,) = torch.autograd.grad(l, y)
        self.T_new.copy_(T_new.detach())
        self.loss.copy_(l.detach().reshape(1))
        self.grad_y.copy_(gy)

    def _intrinsics_host(self):
        """calculate_T_CO_pred uses the FIRST sample's fx, fy, read on the host (Q2, Q5): fixed at capture."""
        if self._fx is not None:
            K0 = torch.tensor([[self._fx, 0, 0], [0, self._fy, 0], [0, 0, 1]], dtype=torch.float32)
        else:
            K0 = self.K[0].detach().cpu()
        return K0.unsqueeze(0).repeat(self.B, 1, 1)

    def capture(self):
        """Warm up on a side stream, then capture three graphs: render, pose+loss, and both."""
        self._Ks_host = self._intrinsics_host()
        s = torch.cuda.Stream(device=self.dev)
        s.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(s):
            for _ in range(2):
                self._render()
                self._pose_loss()
        torch.cuda.current_stream(self.dev).wait_stream(s)
        torch.cuda.synchronize(self.dev)
        self._g_render, self._g_pose, self._g_all = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._g_render):
            self._render()
        with torch.cuda.graph(self._g_pose):
            self._pose_loss()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self._g_all):
            # render and pose-update/loss are independent within one iteration (both read T_pred): two graph
            # branches, so the latency-bound pose / loss kernels run under the rasteriser
            main = torch.cuda.current_stream(self.dev)
            s.wait_stream(main)
            with torch.cuda.stream(s):
                self._pose_loss()
            self._render()
            main.wait_stream(s)
        self.launches_per_run = _lib.launch_count() - n0     # librac kernels replayed by every run()
        self._g_chain = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self._g_chain):
            self._render()
            if self.fused_pose_loss:
                self._pose_loss(feed_back=True)              # T_pred <- T_new inside the launch (train_model.py:250)
            else:
                self._pose_loss()
                self.T_pred.copy_(self.T_new)                # train_model.py:250: the next iteration renders under T_new
                self.T_prev.copy_(self.T_new)                # (the scaled loss reads the iteration's input pose, train_model.py:283)
        self.launches_per_chain = _lib.launch_count() - n0
        return self

    def run(self):
        if self._g_all is None:
            self.capture()
        self._g_all.replay()

    def run_chain(self):
        if self._g_chain is None:
            self.capture()
        self._g_chain.replay()

    def run_render(self):
        if self._g_render is None:
            self.capture()
        self._g_render.replay()

    def run_pose_loss(self):
        if self._g_pose is None:
            self.capture()
        self._g_pose.replay()

    def run_eager(self):
        """Same work without graphs (debugging / parity tests)."""
        if getattr(self, "_Ks_host", None) is None:
            self._Ks_host = self._intrinsics_host()
        self._render()
        self._pose_loss()
