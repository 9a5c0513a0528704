"""Host-side grouping of the pipelined batch entry points (no GPU needed)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "6d-pose-estimation-for-unseen-categories_b200"))


def test_groups_cover_the_batch_in_order_with_a_short_first_group():
    from fmap2pointmap_solvers.pipeline import _groups
    for n in (1, 2, 5, 31, 32, 64, 100, 256, 4096):
        for chunks in (1, 2, 4, 16, 1000):
            g = _groups(n, chunks)
            assert g[0][0] == 0 and g[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(g, g[1:])) and all(lo < hi for lo, hi in g)
    assert _groups(256, 4) == [(0, 16), (16, 64), (64, 128), (128, 192), (192, 256)]
    assert _groups(7, 3) == [(0, 3), (3, 6), (6, 7)]            # small groups are not split


def test_groups_respect_the_sharing_alignment():
    from fmap2pointmap_solvers.pipeline import _groups, _share_align
    import torch
    for n, chunks, align in ((512, 4, 16), (256, 4, 16), (96, 5, 16), (64, 4, 64), (48, 3, 8)):
        g = _groups(n, chunks, align)
        assert g[0][0] == 0 and g[-1][1] == n and all(a[1] == b[0] for a, b in zip(g, g[1:]))
        assert all(lo % align == 0 for lo, _ in g) and all(hi % align == 0 or hi == n for _, hi in g)
    assert _share_align(64, torch.zeros(64, 2, 2), torch.zeros(4, 3, 2), torch.zeros(1, 3, 2)) == 16
    assert _share_align(64, torch.zeros(64, 2, 2)) == 1
