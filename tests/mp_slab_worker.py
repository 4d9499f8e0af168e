"""Worker for the multi-rank tests (one process per rank).

mode `cpu`  (gloo, no GPU): emulates the z-slab time stepping on the CPU with the oracle's RHS and
            exchanges the stage OUTPUT u planes (what libadvdiff ships) instead of the reference's Lu
            planes; rank 0 checks the merged result bit-for-bit against the oracle's whole-domain run.
mode `gpu`  (nccl): every rank drives its GPU through the C ABI (adv_create with nranks/rank/comm_id),
            rank 0 merges the slabs and checks them bit-for-bit against a 1-rank GPU run and to 1e-12
            against the oracle.
Prints `MP_OK ...` on success.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402


def slab_bounds(nz, P, r):
    nzl = nz // P
    return r * nzl, r * nzl + nzl + 6          # global planes [g0, g1) stored by rank r (Tools.c:181-200)


def exchange_cpu(u, rank, P):
    """send planes [3,6) left and [nzs-6,nzs-3) right; receive into [nzs-3,nzs) and [0,3)."""
    reqs = []
    nzs = u.shape[0]
    if rank < P - 1:
        reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(u[nzs - 6:nzs - 3])), rank + 1, tag=1))
    if rank > 0:
        reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(u[3:6])), rank - 1, tag=5))
    if rank < P - 1:
        buf = torch.empty(u[nzs - 3:].shape, dtype=torch.float64)
        dist.recv(buf, rank + 1, tag=5)
        u[nzs - 3:] = buf.numpy()
    if rank > 0:
        buf = torch.empty(u[:3].shape, dtype=torch.float64)
        dist.recv(buf, rank - 1, tag=1)
        u[:3] = buf.numpy()
    for q in reqs:
        q.wait()


def run_cpu(rank, P, nx, ny, nz, nsteps):
    L = orc.lib()
    L.orc_burgers3d_mgpu_rhs_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    L.orc_burgers3d_mgpu_rhs_f64.restype = None
    g = orc.ic_burgers3d_mgpu(nx, ny, nz)
    dx, dy, dz = 2.0 / (nx - 1), 2.0 / (ny - 1), 2.0 / (nz - 1)
    dt = 0.3 * dx
    g0, g1 = slab_bounds(nz, P, rank)
    u = g[g0:g1].copy()
    nzs = u.shape[0]
    for _ in range(nsteps):
        uo = u.copy()
        Lu = np.zeros_like(u)
        for stage in (1, 2, 3):
            L.orc_burgers3d_mgpu_rhs_f64(u.ctypes.data, Lu.ctypes.data, nx, ny, nzs, dx, dy, dz)
            if stage == 1:
                new = uo + dt * Lu
            elif stage == 2:
                new = 0.75 * uo + 0.25 * (u + dt * Lu)
            else:
                new = (uo + 2 * (u + dt * Lu)) / 3
            # ghost planes that a neighbour owns are not ours to advance: they arrive by exchange
            lo = 0 if rank == 0 else 3
            hi = nzs if rank == P - 1 else nzs - 3
            u[lo:hi] = new[lo:hi]
            exchange_cpu(u, rank, P)
    parts = [None] * P if rank == 0 else None
    dist.gather_object(u, parts, dst=0)
    if rank == 0:
        merged = g.copy()
        nzl = nz // P
        for r in range(P):
            merged[r * nzl + 3: r * nzl + 3 + nzl] = parts[r][3:3 + nzl]
        want, _ = orc.burgers3d_mgpu(g, 1e9, max_steps=nsteps)
        assert np.array_equal(merged, want), float(np.abs(merged - want).max())
        print("MP_OK cpu P=%d bit-exact vs oracle" % P)


def run_cpu2d(rank, P, nx, ny, nsteps):
    """y-slab analogue for MultiGPU/Burgers2d_Baseline: rows instead of planes, same exchange indices."""
    L = orc.lib()
    L.orc_burgers2d_mgpu_rhs_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double]
    L.orc_burgers2d_mgpu_rhs_f64.restype = None
    g = orc.ic_burgers2d_mgpu(nx, ny)
    dx, dy = 2.0 / (nx - 1), 2.0 / (ny - 1)
    dt = 0.4 * dx
    g0, g1 = slab_bounds(ny, P, rank)
    u = g[g0:g1].copy()
    nys = u.shape[0]
    Lu = np.zeros_like(u)                      # zeroed once (main.c:172)
    for _ in range(nsteps):
        uo = u.copy()
        for stage in (1, 2, 3):
            L.orc_burgers2d_mgpu_rhs_f64(u.ctypes.data, Lu.ctypes.data, nx, nys, dx, dy)
            if stage == 1:
                new = uo + dt * Lu
            elif stage == 2:
                new = 0.75 * uo + 0.25 * (u + dt * Lu)
            else:
                new = (uo + 2 * (u + dt * Lu)) / 3
            lo = 0 if rank == 0 else 3
            hi = nys if rank == P - 1 else nys - 3
            u[lo:hi] = new[lo:hi]
            exchange_cpu(u, rank, P)
    parts = [None] * P if rank == 0 else None
    dist.gather_object(u, parts, dst=0)
    if rank == 0:
        merged = g.copy()
        nyl = ny // P
        for r in range(P):
            merged[r * nyl + 3: r * nyl + 3 + nyl] = parts[r][3:3 + nyl]
        want, _ = orc.burgers2d_mgpu(g, 1e9, max_steps=nsteps)
        assert np.array_equal(merged, want), float(np.abs(merged - want).max())
        print("MP_OK cpu2d P=%d bit-exact vs oracle" % P)


def run_cpu_heat2d(rank, P, nx, ny, nsteps):
    """y-slab heat (MultiGPU/Diffusion2d_Baseline): the Laplacian rows of a slab are its owned rows, plus pad row 2
    on the FIRST slab only (Compute_Laplace's jstart-1 offset, Kernels.cu:144) -- the rule libadvdiff applies
    through k_global0 == 0.  numpy arithmetic in the reference's operand order; bit-exact against the oracle."""
    g = orc.ic_heat2d_mgpu(nx, ny) + np.random.default_rng(9).random((ny + 6, nx))
    dx, dy = 2.0 / (nx - 1), 2.0 / (ny - 1)
    dt = 1 / (2 * 1.0 * (1 / dx / dx + 1 / dy / dy)) * 0.8
    kx, ky = 1.0 / (12 * dx * dx), 1.0 / (12 * dy * dy)
    g0, g1 = slab_bounds(ny, P, rank)
    u = g[g0:g1].copy()
    nys = u.shape[0]
    jlo = 3 - (1 if rank == 0 else 0)
    for _ in range(nsteps):
        uo = u.copy()
        for stage in (1, 2, 3):
            Lu = np.zeros_like(u)
            c = u[jlo:nys - 3, 3:nx - 3]
            Lu[jlo:nys - 3, 3:nx - 3] = (
                kx * (-u[jlo:nys - 3, 1:nx - 5] + 16 * u[jlo:nys - 3, 2:nx - 4] - 30 * c + 16 * u[jlo:nys - 3, 4:nx - 2] - u[jlo:nys - 3, 5:nx - 1]) +
                ky * (-u[jlo - 2:nys - 5, 3:nx - 3] + 16 * u[jlo - 1:nys - 4, 3:nx - 3] - 30 * c + 16 * u[jlo + 1:nys - 2, 3:nx - 3] - u[jlo + 2:nys - 1, 3:nx - 3]))
            if stage == 1:
                new = uo + dt * Lu
            elif stage == 2:
                new = 0.75 * uo + 0.25 * (u + dt * Lu)
            else:
                new = (uo + 2 * (u + dt * Lu)) / 3
            lo = 0 if rank == 0 else 3
            hi = nys if rank == P - 1 else nys - 3
            u[lo:hi] = new[lo:hi]
            exchange_cpu(u, rank, P)
    parts = [None] * P if rank == 0 else None
    dist.gather_object(u, parts, dst=0)
    if rank == 0:
        merged = g.copy()
        nyl = ny // P
        for r in range(P):
            merged[r * nyl + 3: r * nyl + 3 + nyl] = parts[r][3:3 + nyl]
        want, wi = orc.heat2d_mgpu(g, nsteps)
        assert wi["it"] == nsteps
        assert np.array_equal(merged, want), float(np.abs(merged - want).max())
        print("MP_OK cpu_heat2d P=%d bit-exact vs oracle" % P)


def run_gpu(rank, P, target, nx, ny, nz, nsteps, halo="nccl"):
    import multigpu_advectiondiffusion_b200 as adv
    torch.cuda.set_device(rank)
    idb = [None]
    if rank == 0:
        buf = (C.c_ubyte * 128)()
        adv._lib.check(adv.lib().adv_comm_unique_id(C.byref(buf)))
        idb = [bytes(buf)]
    dist.broadcast_object_list(idb, 0)
    coef = 0.3 if "burgers" in target else 1.0
    two_d = "2d" in target
    if two_d:       # y-slabs: the split axis is y, arrays are (ny+6, nx)
        g = orc.ic_burgers2d_mgpu(nx, ny) if "burgers" in target else orc.ic_heat2d_mgpu(nx, ny)
        nz = ny
    elif target == "burgers3d_sgpu":
        # SingleGPU/Burgers3d_WENO5 semantics (WENO5-Z + viscous term, frozen frame) on slabs: the program's array has
        # nz + 6 planes of which the outer 3 + 3 are its frozen frame -- they take the place of the pad planes
        g = orc.ic_burgers3d_sgpu(nx, ny, nz + 6)
    else:
        g = orc.ic_burgers3d_mgpu(nx, ny, nz) if "burgers" in target else orc.ic_heat3d_mgpu(nx, ny, nz)
    g0, g1 = slab_bounds(nz, P, rank)
    nz_arg = 1 if two_d else (nz + 6 if target == "burgers3d_sgpu" else nz)
    with adv.Solver(target, nx, ny, nz_arg, coef=coef, device=rank, nranks=P, rank=rank, comm_id=idb[0]) as s:
        if halo == "peer":
            hs = [None] * P
            dist.all_gather_object(hs, s.peer_handle())
            s.peer_connect(hs[rank - 1] if rank > 0 else None, hs[rank + 1] if rank < P - 1 else None)
        assert s.halo_mode == halo, s.halo_mode
        # the library's own IC slicing must equal Init_subdomain of the reference IC
        s.init_reference_ic()
        assert np.array_equal(s.download(), g[g0:g1]), "rank %d: IC slab differs" % rank
        s.step(nsteps)
        s.sync()
        u = s.download()
        # and once more WITHOUT any host barrier between the ranks: what comes back from adv_download -- ghost
        # planes included -- goes straight into adv_upload and is stepped again.  A rank whose stream is idle
        # must therefore already hold its neighbours' last stage in its ghost planes (halo_quiesce in
        # csrc/advdiff.cu); stale ghost planes here would break the comparison with the 1-rank run below.
        s.upload(u)
        s.step(nsteps)
        u_again = s.download()
        # the timed exchange on its own (adv_halo_probe) re-sends the boundary planes that are already in the
        # neighbours' ghost planes: it must leave the field as it is
        ms, sent = s.halo_probe(3)
        assert ms > 0 and sent > 0
        assert np.array_equal(s.download(), u_again), "rank %d: adv_halo_probe changed the field" % rank
    parts = [None] * P if rank == 0 else None
    dist.gather_object(u, parts, dst=0)
    parts_again = [None] * P if rank == 0 else None
    dist.gather_object(u_again, parts_again, dst=0)
    if rank == 0:
        merged, merged_again = g.copy(), g.copy()
        nzl = nz // P
        for r in range(P):
            merged[r * nzl + 3: r * nzl + 3 + nzl] = parts[r][3:3 + nzl]
            merged_again[r * nzl + 3: r * nzl + 3 + nzl] = parts_again[r][3:3 + nzl]
        with adv.Solver(target, nx, ny, nz_arg, coef=coef, device=0) as s1:
            s1.upload(g)
            s1.step(nsteps)
            one = s1.download()
            s1.step(nsteps)
            two = s1.download()
        one[:3], one[-3:] = g[:3], g[-3:]
        two[:3], two[-3:] = g[:3], g[-3:]
        assert np.array_equal(merged, one), "P=%d differs from P=1 on the GPU: %g" % (P, float(np.abs(merged - one).max()))
        assert np.array_equal(merged_again, two), "P=%d, download -> upload -> step with no barrier, differs from P=1: %g" % (
            P, float(np.abs(merged_again - two).max()))
        if target == "burgers2d_mgpu":
            want, _ = orc.burgers2d_mgpu(g, 1e9, CFL=coef, max_steps=nsteps)
        elif target == "heat2d_mgpu":
            want, wi = orc.heat2d_mgpu(g, nsteps)
            assert wi["it"] == nsteps
        elif target == "burgers3d_sgpu":
            want, _ = orc.burgers3d_sgpu(g, 1e9, max_steps=nsteps)
        elif "burgers" in target:
            want, _ = orc.burgers3d_mgpu(g, 1e9, max_steps=nsteps)
        else:
            want, _ = orc.heat3d(g, nsteps, "mgpu", L=2.0, W=2.0, H=2.0)
        err = orc.rel_err(merged, want)
        assert err <= 1e-12, err
        print("MP_OK gpu %s P=%d halo=%s bit-exact vs P=1, rel err vs oracle %.2e" % (target, P, halo, err))


def run_gpu_cfl(rank, P, nx, ny, nz, tEnd, halo):
    """adv_run_until_cfl on P ranks: the max|u| of every step is combined across ranks on the device;
    all ranks must take the same steps, and the merged field must equal the 1-rank run bit for bit."""
    import multigpu_advectiondiffusion_b200 as adv
    torch.cuda.set_device(rank)
    idb = [None]
    if rank == 0:
        buf = (C.c_ubyte * 128)()
        adv._lib.check(adv.lib().adv_comm_unique_id(C.byref(buf)))
        idb = [bytes(buf)]
    dist.broadcast_object_list(idb, 0)
    g = 0.8 * orc.ic_burgers3d_mgpu(nx, ny, nz)
    g0, g1 = slab_bounds(nz, P, rank)
    with adv.Solver("burgers3d_mgpu", nx, ny, nz, coef=0.3, device=rank, nranks=P, rank=rank, comm_id=idb[0]) as s:
        if halo == "peer":
            hs = [None] * P
            dist.all_gather_object(hs, s.peer_handle())
            s.peer_connect(hs[rank - 1] if rank > 0 else None, hs[rank + 1] if rank < P - 1 else None)
        s.upload(g[g0:g1].copy())
        info = s.run_until_cfl(tEnd)
        u = s.download()
    infos = [None] * P if rank == 0 else None
    dist.gather_object(info, infos, dst=0)
    parts = [None] * P if rank == 0 else None
    dist.gather_object(u, parts, dst=0)
    if rank == 0:
        assert all(i == infos[0] for i in infos), infos
        merged = g.copy()
        nzl = nz // P
        for r in range(P):
            merged[r * nzl + 3: r * nzl + 3 + nzl] = parts[r][3:3 + nzl]
        with adv.Solver("burgers3d_mgpu", nx, ny, nz, coef=0.3, device=0) as s1:
            s1.upload(g)
            i1 = s1.run_until_cfl(tEnd)
            one = s1.download()
        one[:3], one[-3:] = g[:3], g[-3:]
        assert i1 == infos[0], (i1, infos[0])
        assert np.array_equal(merged, one), "dynamic CFL: P=%d differs from P=1: %g" % (P, float(np.abs(merged - one).max()))
        want, wi = orc.burgers3d_mgpu_cfl(g, tEnd)
        err = orc.rel_err(merged, want)
        assert wi["it"] == i1["it"] and err <= 1e-12, (wi, i1, err)
        print("MP_OK gpu_cfl burgers3d_mgpu P=%d halo=%s %d dynamic steps, bit-exact vs P=1, rel err vs oracle %.2e" % (P, halo, int(i1["it"]), err))


def main():
    mode = sys.argv[1]
    rank, P = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo" if mode == "cpu" else "nccl",
                            **({} if mode == "cpu" else {"device_id": torch.device("cuda", rank)}))
    try:
        if mode == "cpu":
            run_cpu(rank, P, 20, 22, 24, 2)
            run_cpu2d(rank, P, 30, 48, 3)
            run_cpu_heat2d(rank, P, 34, 48, 4)
        else:
            for halo in ("nccl", "peer"):
                for target, shape, steps in (("burgers3d_mgpu", (40, 36, 48), 3), ("burgers3d_sgpu", (40, 36, 48), 3), ("heat3d_mgpu", (64, 40, 48), 5),
                                             ("burgers2d_mgpu", (72, 96, 1), 3), ("heat2d_mgpu", (70, 96, 1), 5)):
                    run_gpu(rank, P, target, *shape, steps, halo=halo)
                run_gpu_cfl(rank, P, 40, 36, 48, 0.15, halo)
    finally:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
