"""The allocation-free Qmax / Dmax entry point (acoss_cover_similarity_from_crp_noalloc: static device scratch, no cudaMalloc)
against the existing allocating entry point, through the C-ABI, on two calls in a row (the scratch keeps the previous call's
contents).  Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _mats(rng, shapes, dens):
    out = []
    for (M, N) in shapes:
        B = (rng.random((M, N)) < dens).astype(np.uint8)
        for _ in range(3):
            L = int(rng.integers(2, max(3, min(M, N))))
            i0, j0 = int(rng.integers(0, max(1, M - 1))), int(rng.integers(0, max(1, N - 1)))
            for k in range(L):
                if i0 + k < M and j0 + k < N:
                    B[i0 + k, j0 + k] = 1
        out.append(B)
    return out


def test_noalloc_cover_similarity_matches_allocating_entry_point():
    import torch
    import acoss_b200
    c = acoss_b200.Context(0)
    try:
        rng = np.random.default_rng(2026)
        # two calls in a row with different shapes / pitch classes: the second must not see the first's leftovers
        for call, (shapes, dens) in enumerate([([(40, 37), (300, 512), (3, 9), (257, 190), (64, 64)], 0.15),
                                               ([(700, 1100), (12, 12), (500, 530)], 0.08),
                                               ([(33, 2300), (90, 70)], 0.3)]):
            mats = _mats(rng, shapes, dens)
            flat = torch.from_numpy(np.concatenate([m.ravel() for m in mats])).cuda()
            out = torch.full((len(mats),), -1.0, dtype=torch.float32, device="cuda")
            torch.cuda.synchronize()
            for align in ("serra09", "chen17"):
                exp = c.cover_similarity_from_crp(mats, alignment=align)
                c.cover_similarity_from_crp_noalloc(flat.data_ptr(), [m.shape[0] for m in mats], [m.shape[1] for m in mats],
                                                    out.data_ptr(), alignment=align)
                assert np.array_equal(out.cpu().numpy(), exp), (call, align)
        # limits and input validation: clear errors, nothing allocated
        big = torch.zeros(4096, dtype=torch.uint8, device="cuda")          # never read: the calls below fail before any launch
        o1 = torch.zeros(1, dtype=torch.float32, device="cuda")
        with pytest.raises(acoss_b200.AcossError) as e:
            c.cover_similarity_from_crp_noalloc(big.data_ptr(), [7000], [2500], o1.data_ptr())       # 7000 x 2560 > 16 MiB
        assert e.value.code == -4 and "16 MiB" in str(e.value)
        with pytest.raises(acoss_b200.AcossError) as e:
            c.cover_similarity_from_crp_noalloc(big.data_ptr(), [10], [3000], o1.data_ptr())         # wider than 2560 columns
        assert e.value.code == -4
        with pytest.raises(acoss_b200.AcossError):
            c.cover_similarity_from_crp_noalloc(big.data_ptr(), [2] * 1025, [2] * 1025, o1.data_ptr())
        bad = torch.full((64,), 3, dtype=torch.uint8, device="cuda")
        with pytest.raises(acoss_b200.AcossError) as e:
            c.cover_similarity_from_crp_noalloc(bad.data_ptr(), [8], [8], o1.data_ptr())
        assert e.value.code == -1
        # and the array is still usable after the failed calls
        m = _mats(rng, [(50, 60)], 0.2)
        f = torch.from_numpy(m[0].ravel().copy()).cuda()
        torch.cuda.synchronize()
        c.cover_similarity_from_crp_noalloc(f.data_ptr(), [50], [60], o1.data_ptr())
        assert np.array_equal(o1.cpu().numpy(), c.cover_similarity_from_crp(m))
    finally:
        c.close()
