"""Host-side logic of the triangular-solve schedules (mmadmm_pcg_plan_levels: no GPU needed).  The reference solves an ILU(0)
factor row by row (lib/LASolver/ILU_class.cpp:447-527); the library runs the same rows level by level, a thread solving a group
of `row_block` consecutive rows in sequence.  Checked here on random and mesh-shaped patterns: every dependency on another
group sits in a strictly earlier level, levels are minimal (1 + the highest dependency), row_block = D gives the Method 2
pattern 1/D of the levels, and the block-structure test accepts exactly the D x D block patterns."""
import importlib
import os

import numpy as np
import pytest
import scipy.sparse as sp

mm = importlib.import_module("mm-admm_b200")
pytestmark = pytest.mark.skipif(not os.path.exists(mm.LIB_PATH), reason="libmmadmm_b200.so not built")


def random_pattern(n, density, seed):
    rng = np.random.default_rng(seed)
    A = sp.random(n, n, density=density, random_state=rng, format="csr") + sp.eye(n, format="csr")
    A = sp.csr_matrix(A); A.sort_indices()
    return A.indptr.astype(np.int32), A.indices.astype(np.int32)


def reference_levels(rowptr, col, bs, lower):
    n = len(rowptr) - 1
    ng = (n + bs - 1) // bs
    lev = np.zeros(ng, dtype=np.int64)
    order = range(ng) if lower else range(ng - 1, -1, -1)
    for g in order:
        l = 0
        for i in range(g * bs, min(n, (g + 1) * bs)):
            for k in range(rowptr[i], rowptr[i + 1]):
                c = col[k]
                if (lower and c < g * bs) or (not lower and c >= (g + 1) * bs):
                    l = max(l, lev[c // bs] + 1)
        lev[g] = l
    return lev


@pytest.mark.parametrize("n,density,seed", [(37, 0.08, 1), (240, 0.02, 2), (601, 0.006, 3)])
@pytest.mark.parametrize("bs", [1, 2, 3, 4, 6])
def test_group_levels_are_a_minimal_valid_layering(n, density, seed, bs):
    rowptr, col = random_pattern(n, density, seed)
    levL, levU, info = mm.pcg_plan_levels(rowptr, col, bs)
    assert np.array_equal(levL, reference_levels(rowptr, col, bs, True))
    assert np.array_equal(levU, reference_levels(rowptr, col, bs, False))
    assert info["lower_levels"] == levL.max() + 1 and info["upper_levels"] == levU.max() + 1
    # validity, stated directly: a row's operand in another group is final before the row's group starts
    for i in range(n):
        g = i // bs
        for k in range(rowptr[i], rowptr[i + 1]):
            c = col[k]
            if c < g * bs:
                assert levL[c // bs] < levL[g]
            elif c >= (g + 1) * bs:
                assert levU[c // bs] < levU[g]
    assert info["blockable"] is False or bs in (2, 3)


@pytest.mark.parametrize("D,nx", [(2, 9), (3, 4)])
def test_method2_pattern_blocks_and_level_counts(D, nx):
    """The Method 2 Jacobian has one dense D x D block per pair of adjacent vertices (Mesh::buildEulerJac, src/Mesh.cpp:1111-1261):
    vertex adjacency of the generated mesh (x) ones(D, D)."""
    Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)
    P = Vp.shape[0]
    rows = np.repeat(F, F.shape[1], axis=1).ravel(); cols = np.tile(F, (1, F.shape[1])).ravel()
    adj = sp.csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(P, P))
    A = sp.csr_matrix(sp.kron(adj, np.ones((D, D)))); A.sort_indices()
    rowptr, col = A.indptr.astype(np.int32), A.indices.astype(np.int32)
    l1, u1, i1 = mm.pcg_plan_levels(rowptr, col, 1)
    lD, uD, iD = mm.pcg_plan_levels(rowptr, col, D)
    assert iD["blockable"] and not i1["blockable"]
    assert iD["lower_levels"] * D == i1["lower_levels"] and iD["upper_levels"] * D == i1["upper_levels"]
    # one entry off the block grid, or a missing one, and the block form is refused
    keep = np.ones(len(col), dtype=bool); keep[rowptr[D] + 0] = col[rowptr[D]] == D   # drop row D's first entry unless it is the diagonal
    if not keep.all():
        rp2 = np.concatenate([[0], np.cumsum(np.add.reduceat(keep.astype(np.int32), rowptr[:-1]))]).astype(np.int32)
        assert not mm.pcg_plan_levels(rp2, col[keep], D)[2]["blockable"]
    if (D * P) % 4 == 0 or True:
        l4, u4, i4 = mm.pcg_plan_levels(rowptr, col, 4)      # groups that straddle vertices (and a short last group): still valid
        assert np.array_equal(l4, reference_levels(rowptr, col, 4, True))


def test_plan_rejects_bad_patterns():
    with pytest.raises(mm.MMADMMError):
        mm.pcg_plan_levels(np.array([0, 1, 2]), np.array([1, 0]), 1)      # no stored diagonal
    with pytest.raises(mm.MMADMMError):
        mm.pcg_plan_levels(np.array([0, 2, 3]), np.array([1, 0, 1]), 1)   # columns not ascending
    with pytest.raises(mm.MMADMMError):
        mm.pcg_plan_levels(np.array([0, 1]), np.array([0]), 5)            # unsupported group size
