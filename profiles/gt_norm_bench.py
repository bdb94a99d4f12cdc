"""rac_assemble_input (prepare_model_input on existing renders: gt_norm + assemble_render kernels) at the bench shape, with
the gt crop moved by bulk asynchronous copies (gt_norm_tma_kernel, default) and by the register-staged kernel
(RAC_GT_TMA=0): device time per call (CUDA events, rotating 4 input/output sets > L2), algorithmic GB/s
(12 B/px gt in + 12 B/px render in + 24 B/px out) and a checksum showing the two are bit-identical.
Usage (GPU box): python profiles/gt_norm_bench.py"""
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200"), ROOT]


def child():
    import torch
    import rac
    from rac import render as R
    R.set_gt_cache(False)
    B, H, W = 64, 240, 320
    g = torch.Generator(device="cuda").manual_seed(3)
    sets = [(torch.rand((B, H, W, 3), device="cuda", generator=g), torch.rand((B, H, W, 3), device="cuda", generator=g)) for _ in range(4)]
    outs = None
    for i in range(8):
        x = rac.assemble_input_device(sets[i % 4][0], sets[i % 4][1], None, False)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(5):
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(40):
            x = rac.assemble_input_device(sets[i % 4][0], sets[i % 4][1], None, False)
        t1.record()
        torch.cuda.synchronize()
        best = min(best, t0.elapsed_time(t1) / 40 * 1e3)
    h = hashlib.sha1()
    for a, b in sets:
        h.update(rac.assemble_input_device(a, b, None, False).cpu().numpy().tobytes())
    algo = 48.0 * W * H * B
    print("%7.1f us per call  %6.0f GB/s algorithmic (%.0f MB)   x sha1 %s" % (best, algo / (best * 1e-6) / 1e9, algo / 1e6, h.hexdigest()[:12]))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        child()
    else:
        for tma in ("1", "0"):
            env = dict(os.environ, RAC_GT_TMA=tma)
            r = subprocess.run([sys.executable, __file__, "child"], capture_output=True, text=True, env=env)
            print("RAC_GT_TMA=%s  %s" % (tma, (r.stdout.strip().splitlines() or [r.stderr[-400:]])[-1]), flush=True)
