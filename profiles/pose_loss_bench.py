"""Steps 3-5 (rotation head + SE(3) update + point-matching loss, forward and backward) at the bench shape (B = 64,
N = 1000): this library's fused kernels against the reference's formulation run through stock torch CUDA ops on the
same GPU (the restatement in oracle/pose_loss_ref.py, which is line-for-line the reference's torch code and is used
here only as the thing being compared against — diagnostic evidence, not a product path).  Reports host-inclusive
time per fwd+bwd (eager Python), device time under a CUDA graph, and the number of kernels launched.
Usage (GPU box): python profiles/pose_loss_bench.py"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200"), ROOT]


def main():
    import torch
    import bench
    import rac
    import loss as L
    import rotation_representation as RR
    from oracle import pose_loss_ref as O
    dev = torch.device("cuda", 0)
    mesh, sets, K = bench.make_inputs(seed=1000)
    s = sets[0]
    B = bench.B
    Ks_cpu = torch.tensor(K)
    Ks_dev = Ks_cpu.to(dev)
    T_in = torch.tensor(s["T_init"], device=dev)
    T_gt = torch.tensor(s["T_gt"], device=dev)
    pts = torch.tensor(s["pts"], device=dev)
    rng = torch.Generator(device=dev); rng.manual_seed(1)
    cases = [("6D", 9, "add_l1_disentangled"), ("SVD", 12, "add_l1_disentangled"), ("3D", 6, "add_l1")]
    print("%-5s %-22s %-28s %10s %12s %9s" % ("rot", "loss", "implementation", "eager us", "graph us", "kernels"))
    for rot, D, loss in cases:
        y0 = 0.1 * torch.randn((B, D), device=dev, generator=rng)
        if rot != "3D":
            y0[:, 0] += 1; y0[:, 4] += 1
        y0[:, -1] = 1.0

        def ours():
            y = y0.clone().requires_grad_(True)
            Tn = RR.calculate_T_CO_pred(y, T_in, rot, Ks_cpu)
            l = L.compute_disentangled_ADD_L1_loss(T_gt, Tn, pts) if loss == "add_l1_disentangled" else L.compute_ADD_L1_loss(T_gt, Tn, pts)
            l.backward()
            return y.grad

        def stock():
            y = y0.clone().requires_grad_(True)
            Tn = O.pose_update(y, T_in, rot, Ks_cpu)
            l = O.disentangled(T_gt, Tn, pts) if loss == "add_l1_disentangled" else O.add_l1(T_gt, Tn, pts)
            l.backward()
            return y.grad

        for name, fn in (("librac_b200 (fused kernels)", ours), ("stock torch ops (reference)", stock)):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            n0 = rac.launch_count()
            t0 = time.perf_counter()
            for _ in range(50):
                fn()
            torch.cuda.synchronize()
            eager = (time.perf_counter() - t0) / 50 * 1e6
            ours_launches = (rac.launch_count() - n0) / 50
            graph_us = float("nan")
            nk = float("nan")
            try:
                g = torch.cuda.CUDAGraph()
                st = torch.cuda.Stream()
                with torch.cuda.stream(st):
                    fn()
                    torch.cuda.synchronize()
                    with torch.cuda.graph(g, stream=st):
                        fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                g.replay(); torch.cuda.synchronize()
                e0.record()
                for _ in range(50):
                    g.replay()
                e1.record(); torch.cuda.synchronize()
                graph_us = e0.elapsed_time(e1) / 50 * 1e3
            except Exception as ex:      # e.g. torch.linalg.svd is not capturable
                graph_us = float("nan")
                sys.stderr.write("graph capture failed for %s/%s: %s\n" % (rot, name, str(ex)[:120]))
            from torch.profiler import profile, ProfilerActivity
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                fn(); torch.cuda.synchronize()
            nk = sum(1 for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()
                     and "memset" not in e.name.lower())
            print("%-5s %-22s %-28s %10.1f %12.1f %9d   (library launches per call: %.0f)" % (rot, loss, name, eager, graph_us, nk, ours_launches))


if __name__ == "__main__":
    main()
