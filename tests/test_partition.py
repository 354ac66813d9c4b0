"""CPU tests of the multi-GPU host logic: the element-block partition plan and the shared-vertex
partial-sum exchange protocol.  The plan comes from the product library (mmadmm_plan_*, no device needed); the exchange
is emulated with numpy in one process and run for real over torch.distributed `gloo` with world_size 2."""
import importlib
import os
import socket

import numpy as np
import pytest

mm = importlib.import_module("mm-admm_b200")


def mesh(D, nx):
    return mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)


def plans(D, Vp, F, nranks):
    return [mm.partition_plan(D, Vp.shape[0], F.shape[0], F, nranks, r) for r in range(nranks)]


@pytest.mark.parametrize("D,nx,nranks", [(2, 16, 2), (2, 23, 3), (2, 32, 8), (3, 4, 2), (3, 5, 4)])
def test_plan_invariants(D, nx, nranks):
    Vp, F, _ = mesh(D, nx)
    P, E = Vp.shape[0], F.shape[0]
    pls = plans(D, Vp, F, nranks)
    # element blocks tile [0, E) in rank order, boundaries on whole 128-element prox tiles
    assert pls[0]["e0"] == 0 and pls[-1]["e1"] == E
    for a, b in zip(pls[:-1], pls[1:]):
        assert a["e1"] == b["e0"] and a["e1"] % 128 == 0
    val = np.bincount(F.ravel(), minlength=P)
    owned = np.zeros(P, dtype=int)
    for r, pl in enumerate(pls):
        v = pl["verts"]
        assert np.all(np.diff(v) > 0)
        assert np.array_equal(v[pl["Floc"]], F[pl["e0"]:pl["e1"]])          # local connectivity maps back
        assert np.array_equal(pl["valence"], val[v])                            # GLOBAL valence
        owned[v[pl["owner"] == 1]] += 1
        # every neighbour segment pairs up with the mirror segment on the other side, same global vertices, same order
        for j, q in enumerate(pl["nbr_rank"]):
            mine = v[pl["shared"][pl["nbr_list"][pl["nbr_ptr"][j]:pl["nbr_ptr"][j + 1]]]]
            o = pls[q]
            jj = list(o["nbr_rank"]).index(r)
            theirs = o["verts"][o["shared"][o["nbr_list"][o["nbr_ptr"][jj]:o["nbr_ptr"][jj + 1]]]]
            assert np.array_equal(mine, theirs)
    assert np.all(owned == 1), "every vertex has exactly one owner"


def exchange_emulated(pls, partial):
    """partial[r]: (nShared_r, K) partial sums -> totals, following the protocol of csrc/partition.h."""
    send = [p[pl["nbr_list"]] for pl, p in zip(pls, partial)]
    tot = []
    for r, pl in enumerate(pls):
        recv = np.zeros_like(send[r])
        for j, q in enumerate(pl["nbr_rank"]):
            o = pls[q]
            jj = list(o["nbr_rank"]).index(r)
            recv[pl["nbr_ptr"][j]:pl["nbr_ptr"][j + 1]] = send[q][o["nbr_ptr"][jj]:o["nbr_ptr"][jj + 1]]
        t = np.zeros_like(partial[r])
        for i in range(len(pl["shared"])):
            for c in pl["contrib"][pl["contrib_ptr"][i]:pl["contrib_ptr"][i + 1]]:
                t[i] += partial[r][i] if c < 0 else recv[c]
        tot.append(t)
    return tot


@pytest.mark.parametrize("D,nx,nranks", [(2, 20, 4), (3, 4, 3)])
def test_exchange_reproduces_global_scatter_sum(D, nx, nranks):
    Vp, F, _ = mesh(D, nx)
    P, E = Vp.shape[0], F.shape[0]
    rng = np.random.default_rng(3)
    ev = rng.standard_normal((E, D + 1, D))                 # an element-slot quantity, like w(z-u)
    glob = np.zeros((P, D))
    np.add.at(glob, F.ravel(), ev.reshape(-1, D))
    pls = plans(D, Vp, F, nranks)
    loc = []
    for pl in pls:
        s = np.zeros((len(pl["verts"]), D))
        np.add.at(s, pl["Floc"].ravel(), ev[pl["e0"]:pl["e1"]].reshape(-1, D))
        loc.append(s)
    tot = exchange_emulated(pls, [s[pl["shared"]] for s, pl in zip(loc, pls)])
    seen = {}
    for pl, s, t in zip(pls, loc, tot):
        s = s.copy()
        s[pl["shared"]] = t
        assert np.allclose(s, glob[pl["verts"]], rtol=1e-13, atol=1e-13)
        for g, row in zip(pl["verts"][pl["shared"]], t):   # all sharers hold bit-identical totals
            if g in seen:
                assert np.array_equal(seen[g], row)
            seen[g] = row


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, D, nx, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        Vp, F, _ = mesh(D, nx)
        P, E = Vp.shape[0], F.shape[0]
        rng = np.random.default_rng(7)
        ev = rng.standard_normal((E, D + 1, D))
        pl = mm.partition_plan(D, P, E, F, world, rank)
        s = np.zeros((len(pl["verts"]), D))
        np.add.at(s, pl["Floc"].ravel(), ev[pl["e0"]:pl["e1"]].reshape(-1, D))
        part = s[pl["shared"]]
        send = torch.from_numpy(part[pl["nbr_list"]].copy())
        recv = torch.zeros_like(send)
        ops = []
        for j, qr in enumerate(pl["nbr_rank"]):
            a, b = int(pl["nbr_ptr"][j]), int(pl["nbr_ptr"][j + 1])
            ops.append(dist.P2POp(dist.isend, send[a:b], int(qr)))
            ops.append(dist.P2POp(dist.irecv, recv[a:b], int(qr)))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        recv = recv.numpy()
        for i in range(len(pl["shared"])):
            t = np.zeros(D)
            for c in pl["contrib"][pl["contrib_ptr"][i]:pl["contrib_ptr"][i + 1]]:
                t += part[i] if c < 0 else recv[c]
            s[pl["shared"][i]] = t
        glob = np.zeros((P, D))
        np.add.at(glob, F.ravel(), ev.reshape(-1, D))
        err = np.abs(s - glob[pl["verts"]]).max()
        # the 3-scalar reduction: all-gather then sum in rank order (identical on every rank)
        loc = torch.tensor([float(s.sum()), float(rank + 1), 1.0], dtype=torch.float64)
        allv = [torch.zeros(3, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(allv, loc)
        tot = sum(v for v in allv)
        q.put((rank, float(err), len(pl["shared"]), tot.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("D,nx", [(2, 24), (3, 4)])
def test_halo_exchange_over_gloo_world2(D, nx):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, D, nx, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    assert all(r[1] < 1e-12 for r in res), res
    assert res[0][2] == res[1][2] > 0          # both sides share the same interface vertices
    assert res[0][3] == res[1][3]              # identical reduced scalars on both ranks
