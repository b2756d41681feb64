"""Slab-sharded stage 3: one grid split along x across ranks, halo exchange every sweep.

Every rank builds the same plan (stages 1-2 + bricks) from the same volume, so brick ids agree everywhere.
Bricks are numbered with the brick plane X (two voxel planes) slowest: rank r owns the contiguous brick range
of its planes [Xa, Xb) and needs, before each sweep, the current values of the neighbouring planes Xa-1 and Xb
(the radius-2 stencil reaches exactly one brick plane).  After each sweep it sends its first and last owned
plane to the neighbours -- 4 SoA sub-ranges per side, packed into one message -- with NCCL send/recv over
NVLink (torch.distributed).  The sweeps themselves are the library's per-range kernel (mrb_qp_plan_sweep).

This is the plain per-sweep NCCL exchange named in BASELINE.json; it is latency-bound (a 256^3 sweep is ~7 us of
GPU work, a send/recv pair tens of us) and exists for correctness and for grids beyond one GPU.  The result is
bit-identical to the single-GPU mrb_optimize for any number of slabs (tested by emulating the ranks on one GPU).
"""
import ctypes as C

import numpy as np

from . import _lib as L
from .api import MrbError


class Plan:
    def __init__(self, ctx, volume, band_size=4, slab=False):
        """slab=True: `volume` is a run of x-planes of a larger grid (mrb_qp_plan_create_slab: the x faces are cuts)."""
        import torch
        self.ctx = ctx
        self.lib = ctx.lib
        keep, pv, shape, like = ctx._in(volume)
        self.shape = tuple(shape)
        self.device = torch.device("cuda", ctx.device)
        h = C.c_void_p()
        create = self.lib.mrb_qp_plan_create_slab if slab else self.lib.mrb_qp_plan_create
        self._check(create(ctx.h, pv, ctx._dims(shape), int(band_size), C.byref(h)))
        self.h = h
        q, nb, nt, nplanes = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.mrb_qp_plan_info(h, C.byref(q), C.byref(nb), C.byref(nt), C.byref(nplanes)))
        self.Q, self.n_band, self.n_tight, self.n_planes = q.value, nb.value, nt.value, nplanes.value
        off = (C.c_int * (self.n_planes + 1))()
        self._check(self.lib.mrb_qp_plan_plane_offsets(h, off))
        self.plane_off = np.array(off[:], dtype=np.int64)

    def _check(self, rc):
        if rc != 0:
            raise MrbError(rc, self.lib.mrb_last_error().decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.mrb_qp_plan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def new_x(self, init=True):
        """Iterate buffer: [4 SoA planes, Q+1 bricks, 4 z] fp32 on the plan's device (slot Q stays zero)."""
        import torch
        x = torch.zeros((4, self.Q + 1, 4), dtype=torch.float32, device=self.device)
        if init:
            self._check(self.lib.mrb_qp_plan_init_x(self.h, C.c_void_p(x.data_ptr())))
        return x

    def sweep(self, x_in, x_out, q0, q1):
        self._check(self.lib.mrb_qp_plan_sweep(self.h, C.c_void_p(x_in.data_ptr()), C.c_void_p(x_out.data_ptr()), int(q0), int(q1)))

    def assemble(self, x):
        import torch
        F = torch.empty(self.shape, dtype=torch.float32, device=self.device)
        self._check(self.lib.mrb_qp_plan_assemble(self.h, C.c_void_p(x.data_ptr()), C.c_void_p(F.data_ptr())))
        return F


def slab_planes(plane_off, world):
    """Splits the brick planes into `world` contiguous groups with balanced brick counts.
    Returns [(Xa, Xb)] per rank (possibly empty groups when there are fewer occupied planes than ranks)."""
    nplanes = len(plane_off) - 1
    Q = int(plane_off[-1])
    cuts = [0]
    for r in range(1, world):
        target = Q * r / world
        X = int(np.searchsorted(plane_off, target, side="left"))
        X = min(max(X, cuts[-1]), nplanes)
        cuts.append(X)
    cuts.append(nplanes)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def _halo_ranges(plane_off, Xa, Xb):
    """(send_low, send_high, recv_low, recv_high) brick ranges of a rank owning planes [Xa, Xb)."""
    n = len(plane_off) - 1
    po = plane_off
    send_low = (int(po[Xa]), int(po[min(Xa + 1, Xb)])) if Xb > Xa else (0, 0)
    send_high = (int(po[max(Xb - 1, Xa)]), int(po[Xb])) if Xb > Xa else (0, 0)
    recv_low = (int(po[Xa - 1]), int(po[Xa])) if Xa > 0 else (0, 0)
    recv_high = (int(po[Xb]), int(po[Xb + 1])) if Xb < n else (0, 0)
    return send_low, send_high, recv_low, recv_high


def solve_emulated(ctx, volume, nranks, iters=500, band_size=4):
    """All ranks on ONE GPU, each with its own iterate buffers and explicit halo copies between them:
    exercises exactly the ownership / halo logic of the distributed run.  Returns F (torch tensor)."""
    plan = Plan(ctx, volume, band_size)
    groups = slab_planes(plan.plane_off, nranks)
    xin = [plan.new_x() for _ in range(nranks)]
    xout = [plan.new_x(init=False) for _ in range(nranks)]
    halos = [_halo_ranges(plan.plane_off, Xa, Xb) for (Xa, Xb) in groups]
    owned = [(int(plan.plane_off[Xa]), int(plan.plane_off[Xb])) for (Xa, Xb) in groups]
    for _ in range(iters):
        for r in range(nranks):
            plan.sweep(xin[r], xout[r], owned[r][0], owned[r][1])
        for r in range(nranks):                     # "send" = neighbour copies from this rank's fresh output
            sl, sh, rl, rh = halos[r]
            if rl[1] > rl[0]:                       # plane Xa-1 lives on the nearest lower rank that owns it
                src = max(s for s in range(r) if owned[s][1] > owned[s][0] and owned[s][1] == rl[1])
                xout[r][:, rl[0]:rl[1], :] = xout[src][:, rl[0]:rl[1], :]
            if rh[1] > rh[0]:
                src = min(s for s in range(r + 1, nranks) if owned[s][1] > owned[s][0] and owned[s][0] == rh[0])
                xout[r][:, rh[0]:rh[1], :] = xout[src][:, rh[0]:rh[1], :]
        xin, xout = xout, xin
    import torch
    x = torch.zeros_like(xin[0])
    for r in range(nranks):
        x[:, owned[r][0]:owned[r][1], :] = xin[r][:, owned[r][0]:owned[r][1], :]
    torch.cuda.synchronize()
    F = plan.assemble(x)
    plan.close()
    return F


def solve_distributed(ctx, volume, iters=500, band_size=4, timings=None):
    """One rank per GPU (torch.distributed, NCCL).  Every rank returns the full F."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    plan = Plan(ctx, volume, band_size)
    groups = slab_planes(plan.plane_off, world)
    owned = [(int(plan.plane_off[Xa]), int(plan.plane_off[Xb])) for (Xa, Xb) in groups]
    nonempty = [r for r in range(world) if owned[r][1] > owned[r][0]]
    Xa, Xb = groups[rank]
    sl, sh, rl, rh = _halo_ranges(plan.plane_off, Xa, Xb)
    lower = max([r for r in nonempty if r < rank], default=None) if rank in nonempty else None
    upper = min([r for r in nonempty if r > rank], default=None) if rank in nonempty else None
    xin, xout = plan.new_x(), plan.new_x(init=False)
    ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ev0.record()
    for _ in range(iters):
        plan.sweep(xin, xout, owned[rank][0], owned[rank][1])
        ops, recvs = [], []
        if lower is not None:
            ops.append(dist.P2POp(dist.isend, xout[:, sl[0]:sl[1], :].contiguous(), lower))
            buf = torch.empty((4, rl[1] - rl[0], 4), dtype=torch.float32, device=xout.device)
            ops.append(dist.P2POp(dist.irecv, buf, lower)); recvs.append((rl, buf))
        if upper is not None:
            ops.append(dist.P2POp(dist.isend, xout[:, sh[0]:sh[1], :].contiguous(), upper))
            buf = torch.empty((4, rh[1] - rh[0], 4), dtype=torch.float32, device=xout.device)
            ops.append(dist.P2POp(dist.irecv, buf, upper)); recvs.append((rh, buf))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            for (rg, buf) in recvs:
                xout[:, rg[0]:rg[1], :] = buf
            torch.cuda.current_stream().synchronize()
        xin, xout = xout, xin
    ev1.record(); torch.cuda.synchronize()
    if timings is not None:
        timings["solve_ms"] = ev0.elapsed_time(ev1)
        timings["n_band"] = plan.n_band
    x = torch.zeros_like(xin)
    x[:, owned[rank][0]:owned[rank][1], :] = xin[:, owned[rank][0]:owned[rank][1], :]
    dist.all_reduce(x)                               # disjoint supports: x + 0 is exact
    torch.cuda.synchronize()
    F = plan.assemble(x)
    plan.close()
    return F


def balanced_cuts(plane_off, world):
    """world+1 brick-plane boundaries with balanced brick counts and at least one occupied plane per rank."""
    groups = slab_planes(plane_off, world)
    cuts = [g[0] for g in groups] + [groups[-1][1]]
    occupied = [X for X in range(len(plane_off) - 1) if plane_off[X + 1] > plane_off[X]]
    if len(occupied) < world:
        raise ValueError("fewer occupied brick planes (%d) than ranks (%d)" % (len(occupied), world))
    # repair empty groups by moving cuts onto occupied planes
    for r in range(1, world):
        lo = occupied[r - 1] + 1
        hi = occupied[len(occupied) - (world - r)]
        cuts[r] = min(max(cuts[r], lo, cuts[r - 1] + 1), hi)
    return cuts


class FusedSolver:
    """Slab-sharded stage 3 with the halo exchange INSIDE the persistent kernel.  Every rank's two iterate buffers and a
    small flag block live in torch symmetric memory (peer-mapped over NVLink); mrb_qp_plan_solve_sharded stores the
    boundary brick planes straight into the neighbours' buffers (plain stores, no TMA) and synchronises with per-sweep
    flags.  The symmetric buffers are allocated once for `max_bricks` bricks and reused by every solve (the rendezvous
    costs milliseconds).  Works with world size 1 (no process group needed)."""

    def __init__(self, ctx, max_bricks):
        import torch
        import torch.distributed as dist
        self.ctx = ctx
        self.device = torch.device("cuda", ctx.device)
        self.distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        self.world = dist.get_world_size() if self.distributed else 1
        self.rank = dist.get_rank() if self.distributed else 0
        self.max_bricks = int(max_bricks)
        nfloat = 16 * (self.max_bricks + 9)          # slab-local layout: plane stride rounded up to 8 bricks
        if self.distributed:
            import torch.distributed._symmetric_memory as symm_mem
            group = dist.group.WORLD.group_name
            self.xa = symm_mem.empty(nfloat, dtype=torch.float32, device=self.device)
            self.xb = symm_mem.empty(nfloat, dtype=torch.float32, device=self.device)
            self.flags = symm_mem.empty(512, dtype=torch.int32, device=self.device)
            ha, hb, hf = (symm_mem.rendezvous(t, group) for t in (self.xa, self.xb, self.flags))
            self._handles = (ha, hb, hf)
            self.pa, self.pb, self.pf = list(ha.buffer_ptrs), list(hb.buffer_ptrs), list(hf.buffer_ptrs)
        else:
            self.xa = torch.empty(nfloat, dtype=torch.float32, device=self.device)
            self.xb = torch.empty(nfloat, dtype=torch.float32, device=self.device)
            self.flags = torch.empty(512, dtype=torch.int32, device=self.device)
            self.pa, self.pb, self.pf = [self.xa.data_ptr()], [self.xb.data_ptr()], [self.flags.data_ptr()]

    def solve_plan(self, plan, iters=500, timings=None):
        """Runs the sharded solve on an existing plan; returns the full iterate x [4, Q+1, 4] on every rank."""
        import torch
        import torch.distributed as dist
        if plan.Q > self.max_bricks:
            raise ValueError("plan has %d bricks, the solver's buffers hold %d" % (plan.Q, self.max_bricks))
        world, rank = self.world, self.rank
        cuts = balanced_cuts(plan.plane_off, world)
        self.flags.zero_()
        plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(self.xa.data_ptr())))
        plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(self.xb.data_ptr())))
        torch.cuda.current_stream().synchronize()
        if self.distributed:
            dist.barrier()                           # every rank's flags are zero and its buffers initialised
            torch.cuda.current_stream().synchronize()
        VP = C.c_void_p * world
        ccuts = (C.c_int * (world + 1))(*cuts)
        plan._check(plan.lib.mrb_qp_plan_solve_sharded(plan.h, rank, world, VP(*self.pa), VP(*self.pb), VP(*self.pf),
                                                       ccuts, int(iters)))
        if timings is not None:
            timings["kernel_ms"] = float(plan.lib.mrb_qp_plan_last_kernel_ms(plan.h))
            timings["n_band"] = plan.n_band
            timings["cuts"] = cuts
        res = (self.xb if iters % 2 else self.xa)[:16 * (plan.Q + 1)].view(4, plan.Q + 1, 4)
        q0, q1 = int(plan.plane_off[cuts[rank]]), int(plan.plane_off[cuts[rank + 1]])
        x = torch.zeros((4, plan.Q + 1, 4), dtype=torch.float32, device=self.device)
        x[:, q0:q1, :] = res[:, q0:q1, :]
        if self.distributed:
            dist.barrier()                           # nobody reuses peer-mapped buffers while a neighbour still runs
            dist.all_reduce(x)                       # disjoint supports: x + 0 is exact
        torch.cuda.current_stream().synchronize()    # the library works on its own stream: x must be complete
        return x

    def reconstruct_sharded(self, volume, iters=500, band_size=4, isolevel=0.0, timings=None):
        """The whole path for ONE grid on several GPUs with stages 3 AND 4 slab-sharded: after the fused
        solve every rank assembles F only on its own voxel planes plus one (the bricks of that extra plane are read from
        the upper neighbour's peer-mapped buffer), runs marching cubes on that slab in whole-grid coordinates, and the
        slabs' meshes are gathered and merged (shared-plane vertices belong to the lower rank, ids = local id + exclusive
        prefix of the ranks' new-vertex counts).  No all-reduce of the iterate, no full-grid F.  Stages 1-2 still run on
        the whole grid on every rank.  Returns ((verts, faces) of the whole grid on every rank, F_slab, (x0, x1))."""
        import time
        import torch
        import torch.distributed as dist
        prof = timings.get("phases") if timings is not None and isinstance(timings.get("phases"), dict) else None
        tlast = [time.perf_counter()]

        def tick(name):
            if prof is not None:
                torch.cuda.synchronize()
                now = time.perf_counter()
                prof[name] = prof.get(name, 0.0) + (now - tlast[0]) * 1e3
                tlast[0] = now
        plan = Plan(self.ctx, volume, band_size)
        tick("plan")
        try:
            world, rank = self.world, self.rank
            if plan.Q > self.max_bricks:
                raise ValueError("plan has %d bricks, the solver's buffers hold %d" % (plan.Q, self.max_bricks))
            cuts = balanced_cuts(plan.plane_off, world)
            self.flags.zero_()
            plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(self.xa.data_ptr())))
            plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(self.xb.data_ptr())))
            torch.cuda.current_stream().synchronize()
            if self.distributed:
                dist.barrier()
                torch.cuda.current_stream().synchronize()
            tick("init+barrier")
            VP = C.c_void_p * world
            ccuts = (C.c_int * (world + 1))(*cuts)
            plan._check(plan.lib.mrb_qp_plan_solve_sharded(plan.h, rank, world, VP(*self.pa), VP(*self.pb), VP(*self.pf),
                                                           ccuts, int(iters)))
            if timings is not None:
                timings["kernel_ms"] = float(plan.lib.mrb_qp_plan_last_kernel_ms(plan.h))
                timings["n_band"] = plan.n_band
                timings["cuts"] = cuts
            tick("solve")
            nx = plan.shape[0]
            x0 = 2 * cuts[rank] if rank > 0 else 0
            x1 = 2 * cuts[rank + 1] if rank + 1 < world else nx
            hi = min(x1 + 1, nx)                                   # cells x0 .. x1-1 need plane x1 too
            n16 = 16 * (plan.Q + 1)
            mine = (self.xb if iters % 2 else self.xa)[:n16].view(4, plan.Q + 1, 4)
            x = mine
            if self.distributed:
                dist.barrier()                                      # every rank's result is in its buffer
                if rank + 1 < world:
                    # the first brick plane of the upper neighbour, out of its peer-mapped buffer (plan layout on return)
                    qa, qb = int(plan.plane_off[cuts[rank + 1]]), int(plan.plane_off[cuts[rank + 1] + 1])
                    h = self._handles[1 if iters % 2 else 0]
                    peer = h.get_buffer(rank + 1, (n16,), torch.float32).view(4, plan.Q + 1, 4)
                    x = mine.clone()
                    x[:, qa:qb, :] = peer[:, qa:qb, :]
                torch.cuda.current_stream().synchronize()
            tick("halo plane")
            F_slab = torch.empty((hi - x0,) + tuple(plan.shape[1:]), dtype=torch.float32, device=self.device)
            plan._check(plan.lib.mrb_qp_plan_assemble_slab(plan.h, C.c_void_p(x.data_ptr()), int(x0), int(hi),
                                                           C.c_void_p(F_slab.data_ptr())))
            if self.distributed:
                dist.barrier()                                      # peers have read this rank's buffer
            tick("assemble slab")
            v, f = self.ctx.mcubes_slab(F_slab, x0, isolevel)
            tick("mcubes slab")
            parts = self._gather_parts(v, f, x0, x1)
            tick("gather")
            mesh = merge_slab_meshes(parts, plan.shape)
            tick("merge")
            return mesh, F_slab, (x0, hi)
        finally:
            plan.close()

    def reconstruct_local(self, volume, iters=500, band_size=4, isolevel=0.0, timings=None):
        """The whole path for ONE grid on several GPUs with EVERY stage slab-sharded:
          stages 1-2  each rank builds its plan from its own x-planes plus a margin of band_size + 5 voxel planes (and the two
                      brick planes the QP halo needs) -- no exchange: everything that far from a cut equals the whole grid's;
          stage 3     mrb_qp_plan_solve_sharded_sub (fused NVLink kernel); the ranks exchange three brick counts first;
          stage 4     as reconstruct_sharded: F on the rank's planes + 1, marching cubes per slab, merged mesh.
        The x-range that holds the shape is cut into `world` equal runs of brick planes (one pass over the volume finds it).
        `volume` is the whole grid on every rank (torch CUDA tensor).  Returns ((verts, faces), F_slab, (x0, hi))."""
        import time
        import torch
        import torch.distributed as dist
        prof = timings.get("phases") if timings is not None and isinstance(timings.get("phases"), dict) else None
        tlast = [time.perf_counter()]

        def tick(name):
            if prof is not None:
                torch.cuda.synchronize()
                now = time.perf_counter()
                prof[name] = prof.get(name, 0.0) + (now - tlast[0]) * 1e3
                tlast[0] = now
        world, rank = self.world, self.rank
        nx = int(volume.shape[0])
        bx = (nx + 1) // 2
        # occupied x-range -> equal runs of brick planes
        occ = (volume.reshape(nx, -1).amax(dim=1) > 0).nonzero()
        if occ.numel() == 0:
            raise ValueError("empty volume")
        xmin, xmax = int(occ[0]), int(occ[-1])
        reach = band_size + 2                                           # the band reaches band_size + 1 voxels beyond the shape
        b0, b1 = max(0, (xmin - reach) // 2), min(bx, (xmax + reach) // 2 + 1)
        if b1 - b0 < world:
            raise ValueError("fewer occupied brick planes (%d) than ranks (%d)" % (b1 - b0, world))
        cuts = [0] + [b0 + (b1 - b0) * r // world for r in range(1, world)] + [bx]
        Xa, Xb = cuts[rank], cuts[rank + 1]
        margin = (band_size + 5 + 1) // 2 + 1                           # brick planes: stage 1-2 margin + the QP halo plane
        s0, s1 = max(0, Xa - margin), min(bx, Xb + margin)              # brick planes of this rank's sub-grid
        v0, v1 = 2 * s0, min(nx, 2 * s1)
        sub = volume[v0:v1]
        tick("cuts")
        plan = Plan(self.ctx, sub, band_size, slab=True)
        tick("plan")
        try:
            if plan.Q > self.max_bricks:
                raise ValueError("plan has %d bricks, the solver's buffers hold %d" % (plan.Q, self.max_bricks))
            po = plan.plane_off
            la, lb = Xa - s0, Xb - s0                                    # owned planes in the sub-grid's numbering
            Qo = int(po[lb] - po[la])
            nlo = int(po[la] - po[la - 1]) if rank > 0 else 0
            nhi = int(po[lb + 1] - po[lb]) if rank + 1 < world else 0
            first = (int(po[la]), int(po[la + 1]))                       # this rank's first owned brick plane, in ITS ids
            if Qo == 0:
                raise ValueError("rank %d owns no brick" % rank)
            info = torch.tensor([Qo, nlo, nhi, first[0], first[1], plan.Q], dtype=torch.int64, device=self.device)
            infos = torch.empty((world, 6), dtype=torch.int64, device=self.device)
            self.flags.zero_()
            plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(self.xa.data_ptr())))
            plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(self.xb.data_ptr())))
            torch.cuda.current_stream().synchronize()
            if self.distributed:
                dist.all_gather_into_tensor(infos, info)                 # also the barrier: every rank's flags are zero
                infos = infos.cpu()
            else:
                infos = info.cpu().view(1, 6)
            lo = infos[rank - 1] if rank > 0 else infos[rank]
            hi_i = infos[rank + 1] if rank + 1 < world else infos[rank]
            nbl = (C.c_int * 6)(int(lo[0]), int(lo[1]), int(lo[2]), int(hi_i[0]), int(hi_i[1]), int(hi_i[2]))
            tick("init+exchange")
            VP = C.c_void_p * world
            plan._check(plan.lib.mrb_qp_plan_solve_sharded_sub(plan.h, rank, world, VP(*self.pa), VP(*self.pb), VP(*self.pf),
                                                               int(la), int(lb), nbl, int(iters)))
            if timings is not None:
                timings["kernel_ms"] = float(plan.lib.mrb_qp_plan_last_kernel_ms(plan.h))
                timings["cuts"] = cuts
            tick("solve")
            x0 = 2 * Xa if rank > 0 else 0
            x1 = 2 * Xb if rank + 1 < world else nx
            hi = min(x1 + 1, nx)
            n16 = 16 * (plan.Q + 1)
            mine = (self.xb if iters % 2 else self.xa)[:n16].view(4, plan.Q + 1, 4)
            x = mine
            if self.distributed:
                dist.barrier()                                           # every rank's result is in its buffer
                if rank + 1 < world:
                    # the upper neighbour's first brick plane: same bricks, same (Y,Z) order, other ids
                    pa_, pb_ = int(hi_i[3]), int(hi_i[4])
                    qa, qb = int(po[lb]), int(po[lb + 1])
                    if pb_ - pa_ != qb - qa:
                        raise RuntimeError("interface plane holds %d bricks here, %d on rank %d" % (qb - qa, pb_ - pa_, rank + 1))
                    h = self._handles[1 if iters % 2 else 0]
                    qp = int(hi_i[5])                                        # the peer's plan has its own brick count
                    peer = h.get_buffer(rank + 1, (16 * (qp + 1),), torch.float32).view(4, qp + 1, 4)
                    x = mine.clone()
                    x[:, qa:qb, :] = peer[:, pa_:pb_, :]
                torch.cuda.current_stream().synchronize()
            tick("halo plane")
            F_slab = torch.empty((hi - x0,) + tuple(plan.shape[1:]), dtype=torch.float32, device=self.device)
            plan._check(plan.lib.mrb_qp_plan_assemble_slab(plan.h, C.c_void_p(x.data_ptr()), int(x0 - v0), int(hi - v0),
                                                           C.c_void_p(F_slab.data_ptr())))
            if self.distributed:
                dist.barrier()                                           # peers have read this rank's buffer
            tick("assemble slab")
            v, f = self.ctx.mcubes_slab(F_slab, x0, isolevel)
            tick("mcubes slab")
            parts = self._gather_parts(v, f, x0, x1)
            tick("gather")
            mesh = merge_slab_meshes(parts, (nx,) + tuple(plan.shape[1:]))
            tick("merge")
            return mesh, F_slab, (x0, hi)
        finally:
            plan.close()

    def _gather_parts(self, v, f, x0, x1):
        """All ranks' (verts, faces, x0, x1): one all-gather of the counts, two of the padded arrays."""
        import torch
        import torch.distributed as dist
        if not self.distributed:
            return [(v, f, x0, x1)]
        world = self.world
        meta = torch.tensor([v.shape[0], f.shape[0], x0, x1], dtype=torch.int64, device=self.device)
        metas = torch.empty((world, 4), dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(metas, meta)
        metas = metas.cpu()
        mv, mf = int(metas[:, 0].max()), int(metas[:, 1].max())
        pv = torch.zeros((max(mv, 1), 3), dtype=torch.float32, device=self.device)
        pf = torch.zeros((max(mf, 1), 3), dtype=torch.int32, device=self.device)
        pv[:v.shape[0]] = v
        pf[:f.shape[0]] = f
        av = torch.empty((world,) + tuple(pv.shape), dtype=torch.float32, device=self.device)
        af = torch.empty((world,) + tuple(pf.shape), dtype=torch.int32, device=self.device)
        dist.all_gather_into_tensor(av, pv)
        dist.all_gather_into_tensor(af, pf)
        return [(av[r, :int(metas[r, 0])], af[r, :int(metas[r, 1])], int(metas[r, 2]), int(metas[r, 3])) for r in range(world)]

    def solve(self, volume, iters=500, band_size=4, timings=None):
        """Every rank returns the full F."""
        plan = Plan(self.ctx, volume, band_size)
        try:
            x = self.solve_plan(plan, iters, timings)
            return plan.assemble(x)
        finally:
            plan.close()

    def reconstruct(self, volume, iters=500, band_size=4, isolevel=0.0, timings=None):
        """The whole path for ONE grid on several GPUs: stages 1-2 on every rank, stage 3 slab-sharded in the fused
        NVLink kernel, then marching cubes + vertex de-dup of the assembled embedding on every rank's copy of F."""
        F = self.solve(volume, iters=iters, band_size=band_size, timings=timings)
        return self.ctx.mcubes(F, isolevel), F


def hash_keys(verts, shape):
    """pointHashmap::hashXYZ (hasher.h:99-104) for a [V, 3] float32 torch tensor: the same fp64 products, left to right,
    truncated toward zero, summed as integers."""
    import torch
    ny, nz = int(shape[1]), int(shape[2])
    v = verts.to(torch.float64)
    kx = (v[:, 0] * 1000000000.0 * float(ny * nz)).to(torch.int64)
    ky = (v[:, 1] * 1000000.0 * float(nz)).to(torch.int64)
    kz = (v[:, 2] * 1000.0).to(torch.int64)
    return kx + ky + kz


def merge_slab_meshes(parts, shape):
    """Meshes of consecutive x-slabs -> the mesh of the whole grid, exactly as mcubes (mcubes.cpp:592-662) numbers it.
    parts: [(verts [V,3] float32 in whole-grid coordinates, faces [F,3] slab-local ids, x0, x1)] in slab order (torch tensors,
    any device); slab r holds the cells x0 <= i < x1, so its triangles follow those of slab r-1 in the reference's cell order
    and its vertex list (first-appearance order of its own triangles) follows slab r-1's.  pointHashmap (hasher.cpp:13-23)
    gives a corner the id of the FIRST vertex with its hash key; inside a slab keys are already unique, so the whole-grid
    numbering is: concatenate the slabs' vertex lists, keep the first vertex of every key (a vertex on a shared plane belongs
    to the lower slab), number the kept ones in order.
    Returns (verts, faces) with global ids."""
    import torch
    parts = [p for p in parts]
    if not parts:
        return torch.zeros((0, 3)), torch.zeros((0, 3), dtype=torch.int32)
    dev = parts[0][0].device
    nvs = [int(p[0].shape[0]) for p in parts]
    V = torch.cat([p[0] for p in parts]) if sum(nvs) else torch.zeros((0, 3), dtype=torch.float32, device=dev)
    voff, acc = [], 0
    for n in nvs:
        voff.append(acc)
        acc += n
    Fc = [p[1].to(torch.int64) + o for p, o in zip(parts, voff) if p[1].shape[0]]
    Fa = torch.cat(Fc) if Fc else torch.zeros((0, 3), dtype=torch.int64, device=dev)
    if V.shape[0] == 0:
        return V, Fa.to(torch.int32)
    # only vertices on a plane shared by two slabs can repeat a key: sort those few, not the whole list
    idx = torch.arange(V.shape[0], device=dev)
    owner = idx.clone()
    planes = sorted({float(p[2]) for p in parts[1:]})
    if planes:
        cand_mask = torch.zeros(V.shape[0], dtype=torch.bool, device=dev)
        for xc in planes:
            cand_mask |= V[:, 0] == xc
        ci = cand_mask.nonzero().squeeze(1)
        if ci.numel():
            ks, perm = torch.sort(hash_keys(V[ci], shape), stable=True)     # equal keys keep their order: the first one owns the key
            start = torch.ones_like(ks, dtype=torch.bool)
            start[1:] = ks[1:] != ks[:-1]
            j = torch.arange(ks.shape[0], device=dev)
            first = torch.cummax(torch.where(start, j, torch.zeros_like(j)), 0).values
            owner[ci[perm]] = ci[perm[first]]
    is_new = owner == idx
    new_id = torch.cumsum(is_new.to(torch.int64), 0) - 1          # id of a kept vertex = how many kept ones precede it
    gid = new_id[owner]
    return V[is_new], gid[Fa].to(torch.int32)


def max_bricks_for(shape):
    """Upper bound of the brick count of a grid (every 2x2x4 brick active), for sizing a FusedSolver."""
    nx, ny, nz = shape
    return ((nx + 1) // 2) * ((ny + 1) // 2) * ((nz + 3) // 4)


def solve_fused(ctx, volume, iters=500, band_size=4, timings=None):
    """One-shot convenience wrapper around FusedSolver (allocates the symmetric buffers for this one solve)."""
    plan = Plan(ctx, volume, band_size)
    try:
        solver = FusedSolver(ctx, plan.Q)
        x = solver.solve_plan(plan, iters, timings)
        return plan.assemble(x)
    finally:
        plan.close()


def reconstruct_fused(ctx, volume, iters=500, band_size=4, isolevel=0.0, timings=None):
    """One-shot whole path on several GPUs (see FusedSolver.reconstruct); every rank returns the same (verts, faces)."""
    F = solve_fused(ctx, volume, iters=iters, band_size=band_size, timings=timings)
    return ctx.mcubes(F, isolevel)


def check_against_single_gpu(ctx, solver, volume, iters=500, band_size=4, reps=3):
    """Runs the slab-sharded whole path `reps` times and compares it, in the run, with this rank's own single-GPU
    mrb_reconstruct / mrb_optimize of the same volume (F, vertices and faces, bit for bit; AND over the ranks).
    Times are device times (CUDA events on the context's stream, which is made torch's current stream so that the
    NCCL gather of the iterate is ordered on it too), max over ranks.  Returns a dict for bench.py's `slab` record."""
    import torch
    import torch.distributed as dist
    from .dist import max_over_ranks
    stream = torch.cuda.ExternalStream(ctx.stream, device=solver.device)
    host = lambda a: a.cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    def same(mesh, F_slab, rng):
        (v2, f2), (xa_, xb_) = mesh, rng
        return bool(torch.equal(F1[xa_:xb_], F_slab) and np.array_equal(host(f1), host(f2)) and
                    np.array_equal(host(v1).view(np.uint32), host(v2).view(np.uint32)))
    import os
    mode = os.environ.get("MRB_SLAB_MODE", "local")      # local: every stage per slab; sharded: stages 1-2 on the whole grid
    run = solver.reconstruct_local if mode == "local" else solver.reconstruct_sharded
    with torch.cuda.stream(stream):
        F1, st = ctx.optimize(volume, band_size=band_size, iters=iters, return_stats=True)
        v1, f1 = ctx.reconstruct(volume, band_size=band_size, iters=iters)
        t = {}
        mesh, F_slab, rng = run(volume, iters=iters, band_size=band_size, timings=t)      # warm-up + check
        ok = same(mesh, F_slab, rng)
        mesh, F_slab, rng = run(volume, iters=iters, band_size=band_size, timings=t)      # second warm-up: the plan arena has its final size now
        ok = ok and same(mesh, F_slab, rng)
        step_ms, solve_ms = [], []
        for _ in range(reps):
            torch.cuda.synchronize()
            if solver.distributed:
                dist.barrier()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            mesh, F_slab, rng = run(volume, iters=iters, band_size=band_size, timings=t)
            e1.record(stream)
            torch.cuda.synchronize()
            ok = ok and same(mesh, F_slab, rng)
            step_ms.append(e0.elapsed_time(e1))
            solve_ms.append(t["kernel_ms"])
        # single-GPU step for the speed-up
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.reconstruct(volume, band_size=band_size, iters=iters)
        e1.record(stream)
        torch.cuda.synchronize()
        single_ms = e0.elapsed_time(e1)
    flag = torch.tensor([1 if ok else 0], device=solver.device)
    if solver.distributed:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    step, solve, single = max_over_ranks([float(np.mean(step_ms)), float(np.mean(solve_ms)), single_ms], device=solver.device)
    B = int(st["n_band"])
    return {"grid": list(volume.shape), "band_voxels": B, "sweeps": int(iters), "world": solver.world,
            "bit_identical": bool(int(flag)), "compared": "this rank's slab of F, and the merged vertices and faces, vs this rank's single-GPU mrb_optimize / mrb_reconstruct; every rank",
            "solve_ms": solve, "step_ms": step, "single_gpu_step_ms": single, "single_gpu_solve_ms": float(st["ms_qp_kernel"]),
            "voxel_updates_per_s": B * iters / (step / 1000.0), "speedup_vs_single_gpu_step": single / step,
            "plane_cuts": t.get("cuts"), "reps": reps, "mode": mode}
