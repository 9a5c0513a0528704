import sys, ctypes, numpy as np, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/6d-pose-estimation-for-unseen-categories_b200')
from fmap2pointmap_solvers import _lib
from fmap2pointmap_solvers.batched import thresholds
from oracle import f2p_oracle as O
L=_lib.load()
for M in [32, 64, 256, 1000]:
    g = torch.Generator().manual_seed(1000 + M)
    ncad, npc = 3 * M + 7, M + 3
    CAD = torch.rand(ncad, 3, generator=g) * 10
    Rm = torch.linalg.qr(torch.randn(3, 3, generator=g)).Q
    p0 = torch.randint(0, ncad, (M,), generator=g)
    p1 = torch.sort(torch.randint(0, npc, (M,), generator=g)).values
    PC = torch.rand(npc, 3, generator=g) * 10
    good = torch.rand(M, generator=g) < 0.6
    PC[p1[good]] = CAD[p0[good]] @ Rm.t() + 0.05 * torch.randn(int(good.sum()), 3, generator=g)
    p = torch.stack([p0, p1], 0).contiguous()
    thr = (ctypes.c_float * 4)(*thresholds(17.3))
    out = torch.zeros(2, M, dtype=torch.int64, device="cuda"); meta = torch.zeros(2, dtype=torch.int32, device="cuda")
    scores = torch.full((3, M), float("nan"), device="cuda")
    ws = torch.empty(L.f2p_filter_workspace_bytes(M)+256, dtype=torch.uint8, device="cuda")
    cd, pd, pp = CAD.cuda(), PC.cuda(), p.cuda()
    rc = L.f2p_spatial_filter(cd.data_ptr(), ncad, pd.data_ptr(), npc, pp.data_ptr(), M, thr, out.data_ptr(), meta.data_ptr(), meta.data_ptr()+4, None, scores.data_ptr(), ws.data_ptr(), ws.numel(), None)
    torch.cuda.synchronize()
    s0 = O.vote_scores(CAD[p0], PC[p1]); got = scores[0].cpu().numpy()
    bad = np.nonzero(got.view(np.int32) != s0.view(np.int32))[0]
    print("M",M,"mismatches",len(bad), bad[:10], got[bad[:5]], s0[bad[:5]])
