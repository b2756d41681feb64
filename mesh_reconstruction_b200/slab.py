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
    F = plan.assemble(x)
    plan.close()
    return F
