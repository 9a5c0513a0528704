"""The on-disk hand-off between the path and pose estimation.

`scripts/eval.py:104-119` attaches `p_pred`, `C_pred` and `ir` to the last element (the `Obj` dict) of every
un-collated sample `[CAD, PC, Obj]` and `torch.save`s it as `<idx>_obj_<obj_id>.pt`; `scripts/test_RANSAC.py:331-376`
and `scripts/test_teaser.py` `torch.load` every `.pt` of that directory in sorted order and read
`result[-1]['p_pred'].T.cpu().detach().numpy()`, `result[0]['xyz']`, `result[-1]['ir']` ...
These helpers write and read exactly that layout for a whole batch of solver outputs, so the batched / multi-GPU path
feeds the pose scripts unchanged.  Plain torch serialization; no arithmetic happens here.
"""
import os
from typing import List, Sequence

import torch


def attach_results(samples: Sequence[list], p_preds: Sequence[torch.Tensor], C_pred, irs: Sequence) -> None:
    """`b[-1].update({"p_pred": ..., "C_pred": ..., "ir": ...})` for every sample (scripts/eval.py:104-108)."""
    if not (len(samples) == len(p_preds) == len(irs)) or len(C_pred) != len(samples):
        raise ValueError("samples, p_preds, C_pred and irs must have one entry per sample")
    for b, ppp, ccc, ir in zip(samples, p_preds, C_pred, irs):
        if ppp.dtype != torch.int64 or ppp.dim() != 2 or ppp.shape[0] != 2:
            raise ValueError(f"p_pred must be int64 [2,M], got {ppp.dtype} {tuple(ppp.shape)}")
        # the batched solvers return every p_pred as a VIEW into one batch-wide result buffer; torch.save serialises a
        # view's whole storage, so a dense copy is attached (a few KB per file, as in the reference layout)
        if ppp.untyped_storage().nbytes() != ppp.numel() * ppp.element_size():
            ppp = ppp.clone()
        b[-1].update({"p_pred": ppp, "C_pred": ccc, "ir": ir})


def save_results(samples: Sequence[list], directory: str, start_idx: int = 0) -> List[str]:
    """One `<idx>_obj_<obj_id>.pt` per sample (scripts/eval.py:110-113); returns the paths.  `start_idx` continues the
    running index across batches (and across ranks: rank r of a pair-sharded run passes its slice's first index)."""
    os.makedirs(directory, exist_ok=True)
    paths = []
    for off, b in enumerate(samples):
        path = os.path.join(directory, f"{start_idx + off}_obj_{b[-1]['obj_id']}.pt")
        torch.save(b, path)
        paths.append(path)
    return paths


def load_results(directory: str, map_location=None) -> List[list]:
    """Every `.pt` of the directory in `sorted(os.listdir(...))` order (scripts/test_RANSAC.py:331-336)."""
    out = []
    for filename in sorted(os.listdir(directory)):
        if filename.endswith(".pt"):
            out.append(torch.load(os.path.join(directory, filename), map_location=map_location, weights_only=False))
    return out


def correspondences(result: list):
    """`(P_pred [M,2] numpy, CAD vertices, ir, obj_id)` of one loaded sample, as test_RANSAC.py:363-376 reads them."""
    obj = result[-1]
    return obj["p_pred"].T.cpu().detach().numpy(), result[0]["xyz"], obj["ir"], obj["obj_id"]
