"""Ground-truth correspondences of a sample on the device: the step right BEFORE the path in the
reference's dataset code.

`find_positives(pc1, pc2, r)` keeps the reference's signature and result (`dataset/object.py:281-288`: int64 `[N,2]`
array of (pc1 index, pc2 index) pairs within distance r, in `np.argwhere` order); `get_overlap` mirrors `:311-317`;
`find_positives_with_overlap` returns all three from one pass.  NumPy arrays in -> NumPy arrays out (as the dataset
code expects), tensors in -> tensors out.  The arithmetic runs in the inputs' own dtype (float64 or float32) in
`positives_kernel` (csrc/positives.cu); there is no CPU path.
"""
import numpy as np
import torch

from . import _lib


def _prep(pc, device):
    t = torch.from_numpy(np.ascontiguousarray(pc)) if isinstance(pc, np.ndarray) else pc.detach()
    if t.dtype not in (torch.float32, torch.float64):
        t = t.to(torch.float64)          # NumPy promotes integer inputs to float64 in the subtraction
    if t.dim() != 2 or t.shape[1] != 3:
        raise ValueError(f"point clouds must be [N,3], got {tuple(t.shape)}")
    return t.to(device).contiguous()


def find_positives_with_overlap(pc1, pc2, r=0.2):
    """(P int64 [N,2], overlap_12 int8 [n1], overlap_21 int8 [n2])."""
    lib = _lib.load()
    as_numpy = isinstance(pc1, np.ndarray)
    device = _lib.cuda_device_for(pc1, pc2)
    with torch.cuda.device(device):
        a, b = _prep(pc1, device), _prep(pc2, device)
        if a.dtype != b.dtype:           # NumPy's subtraction promotes float32 with float64 to float64
            a, b = a.to(torch.float64), b.to(torch.float64)
        n1, n2 = int(a.shape[0]), int(b.shape[0])
        f64 = 1 if a.dtype == torch.float64 else 0
        total = torch.zeros(1, dtype=torch.int64, device=device)
        ov12 = torch.zeros(n1, dtype=torch.int8, device=device)
        ov21 = torch.zeros(n2, dtype=torch.int8, device=device)
        ws = torch.empty(lib.f2p_positives_workspace_bytes(n1) + 256, dtype=torch.uint8, device=device)
        st = _lib.stream_ptr(device)
        rc = lib.f2p_find_positives_count(a.data_ptr(), n1, b.data_ptr(), n2, f64, float(r), total.data_ptr(),
                                          ov12.data_ptr(), ov21.data_ptr(), ws.data_ptr(), ws.numel(), st)
        _lib.check(rc, "f2p_find_positives_count")
        n = int(total.item())            # the one host sync: N sizes the result
        pairs = torch.empty((n, 2), dtype=torch.int64, device=device)
        rc = lib.f2p_find_positives_write(a.data_ptr(), n1, b.data_ptr(), n2, f64, float(r), pairs.data_ptr(), n,
                                          ws.data_ptr(), ws.numel(), st)
        _lib.check(rc, "f2p_find_positives_write")
        if as_numpy:
            return pairs.cpu().numpy(), ov12.cpu().numpy(), ov21.cpu().numpy()
        src = pc1.device
        return pairs.to(src), ov12.to(src), ov21.to(src)


def find_positives(pc1, pc2, r=0.2):
    """Drop-in for `base_object_dataset.find_positives` (dataset/object.py:281-288)."""
    return find_positives_with_overlap(pc1, pc2, r)[0]


def get_overlap(l_1, l_2, p):
    """`base_object_dataset.get_overlap` (dataset/object.py:311-317) on an existing pair list (host-side index scatter)."""
    if isinstance(p, np.ndarray):
        o12, o21 = np.zeros(l_1, dtype=np.byte), np.zeros(l_2, dtype=np.byte)
        o12[p[:, 0]] = 1; o21[p[:, 1]] = 1
        return o12, o21
    o12 = torch.zeros(l_1, dtype=torch.int8, device=p.device); o21 = torch.zeros(l_2, dtype=torch.int8, device=p.device)
    o12[p[:, 0]] = 1; o21[p[:, 1]] = 1
    return o12, o21
