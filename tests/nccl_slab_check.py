"""Worker of tests/test_gpu_nccl.py (launched with torch.distributed.run, one rank per GPU): every rank compresses ITS slab of one
field on ITS GPU through the C-ABI, the slab streams are gathered over NCCL (sz_b200.slabs), and every rank checks every gathered
slab stream byte for byte against the CPU oracle run on that slab (the checker), then decodes all of them on its GPU.
Exit code 0 = all good."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def main():
    import torch
    import torch.distributed as dist
    import szref
    import sz_b200
    from sz_b200 import slabs
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sz_b200.set_device(local)
    cfg = dict(errorBoundMode="REL", relBoundRatio=1e-3)
    sz_b200.init(cfg)
    bad = []
    for dtype, shape in ((np.float32, (97, 64, 80)), (np.float64, (61, 50, 70))):
        full = szref.smooth_field(shape, dtype, span=4 * np.pi)
        bounds = slabs.slab_bounds(shape[0], world)
        lo, hi = bounds[rank]
        x = torch.from_numpy(np.ascontiguousarray(full[lo:hi])).cuda()
        buf = torch.empty(sz_b200.compress_bound(dtype, tuple(x.shape)), dtype=torch.uint8, device="cuda")
        for rep in range(2):            # twice: the second call reuses stream_buf while nothing may still read it
            sizes, parts = slabs.compress_slab_and_gather(x, buf, sz_b200.REL, 1e-4, 1e-3)
        container = slabs.container_from(sizes, parts)
        streams = slabs.unframe(container)
        orc = szref.Oracle()
        orc.init(cfg)
        for r, (l2, h2) in enumerate(bounds):
            ref_slab = np.ascontiguousarray(full[l2:h2])
            want = orc.compress(ref_slab, szref.REL, 1e-4, 1e-3)
            if streams[r] != want:
                bad.append("rank %d: slab %d of %s differs from the oracle stream (%d vs %d bytes)" % (rank, r, np.dtype(dtype).name, len(streams[r]), len(want)))
                continue
            out = torch.empty(ref_slab.shape, dtype=x.dtype, device="cuda")
            sz_b200.decompress_device(parts[r], sizes[r], out)
            if out.cpu().numpy().tobytes() != orc.decompress(want, dtype, ref_slab.shape).tobytes():
                bad.append("rank %d: slab %d of %s decodes differently" % (rank, r, np.dtype(dtype).name))
    flag = torch.tensor([len(bad)], dtype=torch.int64, device="cuda")
    dist.all_reduce(flag)
    for b in bad:
        print(b, flush=True)
    if rank == 0:
        print("nccl slab check: world %d, %s" % (world, "ok" if int(flag.item()) == 0 else "FAILED"), flush=True)
    sz_b200.finalize()
    dist.destroy_process_group()
    return 1 if int(flag.item()) else 0


if __name__ == "__main__":
    sys.exit(main())
