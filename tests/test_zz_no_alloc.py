"""mrb_find_indexes_noalloc: the entry point for callers that cannot let the library allocate.

Some hosts of this library own all device memory themselves (a framework's caching allocator, a fixed
arena in a perception stack) or run under a deadline where cudaMalloc's device-wide stall is not
acceptable.  For them libmrb offers one compaction call whose scratch is a fixed array inside the
library.  This test drives it through the C-ABI with caller-owned device buffers and checks it
against the regular mrb_find_indexes on two consecutive calls -- the second on a smaller, different
grid, so that whatever the first call left in the shared scratch array would show up as an error.

Collected last on purpose (file name): a failure here cannot hide any other test of a -x run.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_matches_find_indexes_twice(ctx, synth):
    import torch
    lib = ctx.lib
    torch.cuda.set_device(0)
    rng = np.random.default_rng(21)
    first = (rng.random((61, 47, 53)) > 0.7).astype(np.float32)         # ragged dims, many tiles
    second = synth.sphere(40, 11.0)                                        # smaller grid, fewer tiles
    for vol in (first, second):
        g = torch.from_numpy(vol).cuda()
        expect = ctx.findIndexes(g)                                        # the existing entry point
        dims = (C.c_int * 3)(*vol.shape)
        n = C.c_int(-1)
        # count-only call (also loads the kernels, so the memory check below sees a warm library)
        assert lib.mrb_find_indexes_noalloc(C.c_void_p(g.data_ptr()), dims, None, 0, C.byref(n), None) == 0
        assert n.value == len(expect) == int(np.count_nonzero(vol))
        # the caller owns the output buffer
        cap = len(expect) + 5
        out = torch.full((cap,), -7, dtype=torch.int32, device=g.device)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        n = C.c_int(-1)
        rc = lib.mrb_find_indexes_noalloc(C.c_void_p(g.data_ptr()), dims, C.c_void_p(out.data_ptr()), cap,
                                          C.byref(n), None)
        free_after, _ = torch.cuda.mem_get_info()
        assert rc == 0, lib.mrb_last_error().decode()
        assert free_after >= free_before                                   # the device lost no memory to the call
        assert n.value == len(expect)
        assert torch.equal(out[:n.value], expect)                          # bit-exact, same order
        assert bool((out[n.value:] == -7).all())                           # nothing written past the end


def test_noalloc_rejects_what_it_cannot_do(ctx):
    import torch
    lib = ctx.lib
    torch.cuda.set_device(0)
    n = C.c_int(0)
    # beyond the built-in scratch: refused with a clear message, nothing launched
    big = (C.c_int * 3)(257, 256, 256)
    g = torch.zeros(8, dtype=torch.float32, device="cuda")
    rc = lib.mrb_find_indexes_noalloc(C.c_void_p(g.data_ptr()), big, None, 0, C.byref(n), None)
    assert rc == -5 and b"MRB_NOALLOC_MAX_VOXELS" in lib.mrb_last_error()
    # host buffers would need staging memory: refused
    h = np.ones((4, 4, 4), np.float32)
    rc = lib.mrb_find_indexes_noalloc(C.c_void_p(h.ctypes.data), (C.c_int * 3)(4, 4, 4), None, 0, C.byref(n), None)
    assert rc == -1 and b"device pointers" in lib.mrb_last_error()
    # output too small
    g = torch.ones((4, 4, 4), dtype=torch.float32, device="cuda")
    out = torch.zeros(10, dtype=torch.int32, device="cuda")
    rc = lib.mrb_find_indexes_noalloc(C.c_void_p(g.data_ptr()), (C.c_int * 3)(4, 4, 4), C.c_void_p(out.data_ptr()), 10,
                                      C.byref(n), None)
    assert rc == -5 and n.value == 64
