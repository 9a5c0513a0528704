# A/B of the k = 64 stage layout (4 x 16 on the 4-deep ring vs 2 x 32 on the 2-deep ring), 50k x 20k naive
for kc in 16 32 16 32; do F2P_GATED_KC=$kc python - <<PY
import sys, torch
sys.path.insert(0, "6d-pose-estimation-for-unseen-categories_b200")
from fmap2pointmap_solvers import naive_fmap2pointmap
from fmap2pointmap_solvers.synth import make_pair
for k in (64, 128):
    p = make_pair(50000, 20000, k, seed=3)
    C, ex, ey = p.C12.cuda(), p.evecs_x.cuda(), p.evecs_y.cuda()
    for _ in range(5): naive_fmap2pointmap(C, ex, ey)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(20): naive_fmap2pointmap(C, ex, ey)
    b.record(); torch.cuda.synchronize()
    print("kc=$kc k=%d ms=%.4f" % (k, a.elapsed_time(b) / 20))
PY
done
