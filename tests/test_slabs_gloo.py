"""world_size-2 gloo test (CPU) of the multi-GPU host logic: slab partition, size + stream all-gather, container framing.
The per-slab compressor is a CPU stand-in (the oracle) injected through compress_fn -- test infrastructure only."""
import os
import socket

import numpy as np
import pytest

import szref

pytestmark = pytest.mark.skipif(not szref.have_oracle(), reason="oracle not built")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, shape, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sz_b200 import slabs
    full = szref.smooth_field(shape, np.float32, span=4 * np.pi)
    lo, hi = slabs.slab_bounds(shape[0], world)[rank]
    slab = np.ascontiguousarray(full[lo:hi])
    orc = szref.Oracle()
    orc.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))

    def compress_fn(x, buf):
        s = orc.compress(x.numpy(), szref.REL, 1e-4, 1e-3)
        buf[:len(s)] = torch.frombuffer(bytearray(s), dtype=torch.uint8)
        return len(s)

    buf = torch.zeros(slab.nbytes + 4096, dtype=torch.uint8)
    sizes, parts = slabs.compress_slab_and_gather(torch.from_numpy(slab), buf, szref.REL, 1e-4, 1e-3, compress_fn=compress_fn)
    container = slabs.container_from(sizes, parts)
    # every rank holds the same container; decode every slab and check the bound slab by slab
    streams = slabs.unframe(container)
    ok = len(streams) == world
    for r, (l2, h2) in enumerate(slabs.slab_bounds(shape[0], world)):
        ref_slab = np.ascontiguousarray(full[l2:h2])
        d = orc.decompress(streams[r], np.float32, ref_slab.shape)
        eb = 1e-3 * float(ref_slab.max() - ref_slab.min())
        ok = ok and float(np.max(np.abs(d.astype(np.float64) - ref_slab))) <= eb * (1 + 1e-6)
        ok = ok and streams[r] == orc.compress(ref_slab, szref.REL, 1e-4, 1e-3)
    q.put((rank, ok, sizes))
    dist.destroy_process_group()


def test_slab_gather_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    shape = (45, 32, 40)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, shape, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] for r in res), res
    assert res[0][2] == res[1][2]


def test_slab_bounds_and_framing():
    from sz_b200 import slabs
    for n0, g in ((512, 8), (5500, 8), (100, 3), (13, 4)):
        b = slabs.slab_bounds(n0, g)
        assert b[0][0] == 0 and b[-1][1] == n0 and all(b[i][1] == b[i + 1][0] for i in range(g - 1))
        sizes = [h - l for l, h in b]
        assert max(sizes) - min(sizes) <= 1
    parts = [b"abc", b"", b"\x00\x01" * 50]
    assert slabs.unframe(slabs.frame(parts)) == parts
