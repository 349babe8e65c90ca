"""CPU, world_size 2 (gloo): the host-side logic of the row-partitioned path.  Every rank builds
its Partition, the halo exchange the device does with ncclSend/ncclRecv is replayed with
torch.distributed send/recv from the same tables, the owned rows of the local (oracle) matrices
are applied, inner products are all-reduced -- and a Jacobi-CG on the distributed operator is
compared with the global oracle solve.  Launched by tests/test_dist_gloo.py."""
import os
import sys

import numpy as np
import scipy.sparse.linalg as spl
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import assembly, sem                                  # noqa: E402
from pynama_b200.domain.box_mesh import BoxMesh                   # noqa: E402
from pynama_b200.parallel import Partition                        # noqa: E402
from tests.util import fs_classes                                 # noqa: E402


def halo_exchange(part, tables, x, bs):
    """x: local vector (n_loc*bs); fills the ghost section from the owners (blocking pairs,
    lower rank sends first)"""
    ranks, send_ptr, send_nodes, recv_off, recv_cnt = tables
    xv = x.reshape(-1, bs)
    for i, q in enumerate(ranks):
        sbuf = torch.from_numpy(np.ascontiguousarray(xv[send_nodes[send_ptr[i]:send_ptr[i + 1]]]))
        rbuf = torch.zeros((int(recv_cnt[i]), bs), dtype=torch.float64)
        ops = []
        if sbuf.numel():
            ops.append(dist.P2POp(dist.isend, sbuf, int(q)))
        if rbuf.numel():
            ops.append(dist.P2POp(dist.irecv, rbuf, int(q)))
        for r in dist.batch_isend_irecv(ops):
            r.wait()
        a = part.n_own + int(recv_off[i])
        xv[a:a + int(recv_cnt[i])] = rbuf.numpy()


def allsum(v):
    t = torch.tensor([v], dtype=torch.float64)
    dist.all_reduce(t)
    return float(t.item())


def main():
    dist.init_process_group("gloo")
    rank, size = dist.get_rank(), dist.get_world_size()
    dim, ngl, nelem = 3, 3, [3, 3, 4]
    mesh = BoxMesh(nelem, [0] * dim, [1] * dim, ngl, perturb=0.1)
    tab = sem.SpectralTables(ngl, dim)
    cls = fs_classes(mesh)
    part = Partition(mesh, rank, size)
    tables = part.halo_tables()
    ldofs = (part.l2g[:, None] * dim + np.arange(dim)[None, :]).ravel()
    L = assembly.assemble_kle(tab, part.corner_xyz, part.elem_nodes, part.n_loc, cls[ldofs], "FS")
    K = L["K"].tocsr()[:part.n_own * dim]                       # owned rows only
    nown = part.n_own * dim

    # global reference
    G = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    rng = np.random.default_rng(3)
    xg = rng.uniform(-1, 1, mesh.nnodes * dim)
    # 1. halo exchange delivers exactly the owners' values
    x = np.zeros(part.n_loc * dim)
    x[:nown] = xg[ldofs[:nown]]
    halo_exchange(part, tables, x, dim)
    assert np.array_equal(x, xg[ldofs]), "ghost values differ from the owners' values"
    # 2. distributed SpMV == global SpMV on the owned rows
    y = K @ x
    want = (G["K"] @ xg)[ldofs[:nown]]
    assert np.abs(y - want).max() <= 1e-12 * np.abs(want).max()
    # 3. distributed Jacobi-CG (allreduced dots) == global direct solve
    b_g = G["K"] @ xg
    b = b_g[ldofs[:nown]]
    dinv = 1.0 / G["K"].diagonal()[ldofs[:nown]]
    sol = np.zeros(part.n_loc * dim)
    r = b.copy()
    z = dinv * r
    p = np.zeros(part.n_loc * dim)
    p[:nown] = z
    rz = allsum(r @ z)
    for it in range(2000):
        halo_exchange(part, tables, p, dim)
        w = K @ p
        alpha = rz / allsum(p[:nown] @ w)
        sol[:nown] += alpha * p[:nown]
        r -= alpha * w
        z = dinv * r
        rz_new = allsum(r @ z)
        if np.sqrt(allsum(r @ r)) < 1e-11 * np.linalg.norm(b_g):
            break
        p[:nown] = z + (rz_new / rz) * p[:nown]
        rz = rz_new
    err = np.abs(sol[:nown] - xg[ldofs[:nown]]).max()
    assert err < 1e-8, err
    print(f"[gloo_halo_check] rank {rank}/{size}: neighbours {list(tables[0])}, own {part.n_own} ghost {part.n_ghost}, "
          f"CG its {it + 1}, err {err:.2e} OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
