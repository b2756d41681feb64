"""binvox files (the shape-completion network's output, read_binvox in binvoxToPCL/binvoxToPcl.cpp:15-103): host I/O only.

read() returns what mrb_binvox_to_points / Context.front_end take: the decoded voxel bytes in file order
(index = x * width * height + z * width + y), dwh = (depth, width, height) and xform = (scale, tx, ty, tz)."""
import numpy as np


def read(path):
    with open(path, "rb") as f:
        data = f.read()
    head, sep, body = data.partition(b"data\n")
    if not sep:
        head, sep, body = data.partition(b"data\r\n")
    tok = head.split()
    if not tok or tok[0] != b"#binvox" or not sep:
        raise ValueError("%s: not a binvox file" % path)
    dims, tr, scale = None, (0.0, 0.0, 0.0), 1.0
    i = 2
    while i < len(tok):
        if tok[i] == b"dim":
            dims = tuple(int(t) for t in tok[i + 1:i + 4]); i += 4         # depth, height, width (the reference's reading order)
        elif tok[i] == b"translate":
            tr = tuple(float(t) for t in tok[i + 1:i + 4]); i += 4
        elif tok[i] == b"scale":
            scale = float(tok[i + 1]); i += 2
        else:
            i += 1
    if dims is None:
        raise ValueError("%s: missing dimensions in header" % path)
    depth, height, width = dims
    runs = np.frombuffer(body[:len(body) // 2 * 2], np.uint8).reshape(-1, 2)
    vox = np.repeat(runs[:, 0], runs[:, 1].astype(np.int64))
    size = depth * height * width
    if vox.size < size:
        vox = np.concatenate([vox, np.zeros(size - vox.size, np.uint8)])
    return np.ascontiguousarray(vox[:size]), (depth, width, height), (np.float32(scale), np.float32(tr[0]), np.float32(tr[1]), np.float32(tr[2]))


def write(path, voxels, dwh, xform=(1.0, 0.0, 0.0, 0.0)):
    """voxels: bytes in file order (or an array shaped [depth][height][width] = [x][z][y])."""
    d, w, h = dwh
    flat = np.ascontiguousarray(voxels, np.uint8).ravel()
    assert flat.size == d * w * h
    change = np.flatnonzero(np.diff(flat)) + 1
    starts = np.concatenate([[0], change])
    lens = np.diff(np.concatenate([starts, [flat.size]]))
    out = bytearray()
    for s, n in zip(starts.tolist(), lens.tolist()):
        v = int(flat[s])
        while n > 0:
            c = min(n, 255)
            out += bytes((v, c))
            n -= c
    with open(path, "wb") as f:
        f.write(b"#binvox 1\ndim %d %d %d\ntranslate %r %r %r\nscale %r\ndata\n" % (d, h, w, float(xform[1]), float(xform[2]), float(xform[3]), float(xform[0])))
        f.write(bytes(out))
