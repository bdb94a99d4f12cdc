import cProfile, pstats, sys, os, io
sys.path[:0]=["/root/repo/end-to-end-rgb-pose-estimation-baseline_b200","/root/repo"]
import numpy as np, torch
import bench, data_loaders as DL, loss as L, rotation_representation as RR
from rac import mesh as M
dev=torch.device("cuda",0)
mesh,sets,K=bench.make_inputs(seed=1000); B,W,H=bench.B,bench.W,bench.H
ply="/tmp/m.ply"; M.save_ply(ply,mesh); paths=[ply]*B
Ks=torch.tensor(K.astype(np.float64)); rng=np.random.default_rng(0)
hs=[dict(gt=rng.integers(0,256,(B,H,W,3),dtype=np.uint8)/255.0,T=torch.tensor(s["T_init"].astype(np.float64)),Tg=torch.tensor(s["T_gt"]),y=torch.tensor(s["y"]).pin_memory(),pts=torch.tensor(s["pts"])) for s in sets]
def batch(i):
    s=hs[i%4]; T_pred=s["T"]; T_gt=s["Tg"].to(dev); pts=s["pts"].to(dev)
    for j in range(3):
        imgs,depths=DL.render_batch(T_pred if j else T_pred.numpy(),paths,Ks,(W,H),False)
        T_dev=torch.tensor(T_pred).to(dev) if j else T_pred.to(dev)
        x=DL.prepare_model_input(imgs,s["gt"],depths,False).to(dev)
        y=s["y"].to(dev,non_blocking=True).requires_grad_(True)
        T_new=RR.calculate_T_CO_pred(y,T_dev,bench.ROT,Ks)
        l=L.compute_disentangled_ADD_L1_loss(T_gt,T_new,pts)
        l.backward()
        T_pred=T_new.detach().cpu().numpy()
import contextlib
with contextlib.redirect_stdout(io.StringIO()):
    for i in range(4): batch(i)
    pr=cProfile.Profile(); pr.enable()
    for i in range(20): batch(i)
    pr.disable()
st=pstats.Stats(pr); st.sort_stats("tottime").print_stats(28)
