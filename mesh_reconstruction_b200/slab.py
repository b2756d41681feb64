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
    def __init__(self, ctx, volume, band_size=4):
        import torch
        self.ctx = ctx
        self.lib = ctx.lib
        keep, pv, shape, like = ctx._in(volume)
        self.shape = tuple(shape)
        self.device = torch.device("cuda", ctx.device)
        h = C.c_void_p()
        self._check(self.lib.mrb_qp_plan_create(ctx.h, pv, ctx._dims(shape), int(band_size), C.byref(h)))
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


def solve_fused(ctx, volume, iters=500, band_size=4, timings=None):
    """Slab-sharded solve with the halo exchange INSIDE the persistent kernel: every rank's iterate buffers and a
    small flag block live in torch symmetric memory (peer-mapped over NVLink); mrb_qp_plan_solve_sharded stores the
    boundary brick planes straight into the neighbours' buffers and synchronises with per-sweep flags.
    Works with world size 1 (no process group needed).  Every rank returns the full F."""
    import torch
    import torch.distributed as dist
    distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    world = dist.get_world_size() if distributed else 1
    rank = dist.get_rank() if distributed else 0
    plan = Plan(ctx, volume, band_size)
    cuts = balanced_cuts(plan.plane_off, world)
    nfloat = 16 * (plan.Q + 9)                       # slab-local layout: plane stride rounded up to 8 bricks
    if distributed:
        import torch.distributed._symmetric_memory as symm_mem
        group = dist.group.WORLD.group_name
        xa = symm_mem.empty(nfloat, dtype=torch.float32, device=plan.device)
        xb = symm_mem.empty(nfloat, dtype=torch.float32, device=plan.device)
        flags = symm_mem.empty(512, dtype=torch.int32, device=plan.device)
        ha, hb, hf = (symm_mem.rendezvous(t, group) for t in (xa, xb, flags))
        pa, pb, pf = list(ha.buffer_ptrs), list(hb.buffer_ptrs), list(hf.buffer_ptrs)
    else:
        xa = torch.empty(nfloat, dtype=torch.float32, device=plan.device)
        xb = torch.empty(nfloat, dtype=torch.float32, device=plan.device)
        flags = torch.empty(512, dtype=torch.int32, device=plan.device)
        pa, pb, pf = [xa.data_ptr()], [xb.data_ptr()], [flags.data_ptr()]
    flags.zero_()
    plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(xa.data_ptr())))
    plan._check(plan.lib.mrb_qp_plan_init_x(plan.h, C.c_void_p(xb.data_ptr())))
    torch.cuda.synchronize()
    if distributed:
        dist.barrier()
        torch.cuda.synchronize()
    VP = C.c_void_p * world
    ccuts = (C.c_int * (world + 1))(*cuts)
    plan._check(plan.lib.mrb_qp_plan_solve_sharded(plan.h, rank, world, VP(*pa), VP(*pb), VP(*pf), ccuts, int(iters)))
    if timings is not None:
        timings["kernel_ms"] = float(plan.lib.mrb_qp_plan_last_kernel_ms(plan.h))
        timings["n_band"] = plan.n_band
        timings["cuts"] = cuts
    res = (xb if iters % 2 else xa)[:16 * (plan.Q + 1)].view(4, plan.Q + 1, 4)
    q0, q1 = int(plan.plane_off[cuts[rank]]), int(plan.plane_off[cuts[rank + 1]])
    x = torch.zeros((4, plan.Q + 1, 4), dtype=torch.float32, device=plan.device)
    x[:, q0:q1, :] = res[:, q0:q1, :]
    if distributed:
        dist.barrier()                               # nobody frees peer-mapped buffers while a neighbour still runs
        dist.all_reduce(x)
    torch.cuda.synchronize()                         # the library works on its own stream: x must be complete
    F = plan.assemble(x)
    plan.close()
    return F


def reconstruct_fused(ctx, volume, iters=500, band_size=4, isolevel=0.0, timings=None):
    """The whole path for ONE grid on several GPUs: stages 1-2 on every rank (they are a few per cent of a step),
    stage 3 slab-sharded in the fused NVLink kernel, then marching cubes + vertex de-dup of the assembled embedding.
    Every rank returns the same (verts, faces): stage 4 is not sharded, it runs on every rank's copy of F."""
    F = solve_fused(ctx, volume, iters=iters, band_size=band_size, timings=timings)
    return ctx.mcubes(F, isolevel)
