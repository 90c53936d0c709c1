"""Multi-GPU slab sharding: the slowest dimension is cut into G contiguous, independent SZ slabs, one per
rank / GPU, exactly like the reference's slab-parallel example (example/test_threadsafe.c:40-58).  Each slab is a complete
SZ stream (bit-exact to the reference called on that slab, decodable by the reference).  The only exchange step is an
all-gather of the stream sizes followed by an all-gather of the (max-padded) streams; the container

    [G : u32 LE][size_g : u64 LE x G][stream_0 ... stream_{G-1}]

is ours -- the reference has no multi-slab container.  Under REL / ABS_AND_REL every slab derives its bound from ITS OWN
value range, as in the reference's example, so the concatenation is not byte-equal to one whole-field call.
"""
import struct

import numpy as np


def slab_bounds(n0, world):
    """contiguous [lo, hi) ranges of the slowest dimension, sizes floor(n0/G) or +1 (first n0 % G slabs get the +1)."""
    base, extra = divmod(int(n0), int(world))
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def frame(streams):
    head = struct.pack("<I", len(streams)) + b"".join(struct.pack("<Q", len(s)) for s in streams)
    return head + b"".join(streams)


def unframe(buf):
    (g,) = struct.unpack_from("<I", buf, 0)
    sizes = struct.unpack_from("<%dQ" % g, buf, 4)
    off = 4 + 8 * g
    out = []
    for s in sizes:
        out.append(bytes(buf[off:off + s]))
        off += s
    return out


def gather_streams(local_stream, local_size, group=None):
    """All-gather of per-rank streams.  local_stream: 1-D uint8 torch tensor (CUDA with NCCL, CPU with gloo) holding at least
    local_size valid bytes.  Returns (sizes list, list of per-rank uint8 tensors trimmed to their sizes)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dev = local_stream.device
    sz = torch.tensor([int(local_size)], dtype=torch.int64, device=dev)
    all_sz = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(all_sz, sz, group=group)                       # 8 bytes per rank
    sizes = [int(t.item()) for t in all_sz]
    cap = max(sizes)
    if local_stream.numel() < cap:
        pad = torch.zeros(cap, dtype=torch.uint8, device=dev)
        pad[:local_size] = local_stream[:local_size]
        send = pad
    else:
        send = local_stream[:cap]
    bufs = [torch.empty(cap, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(bufs, send.contiguous(), group=group)          # the concatenated stream (about 1/10 of the input)
    return sizes, [b[:s] for b, s in zip(bufs, sizes)]


def compress_slab_and_gather(x_slab, stream_buf, mode, abs_eb, rel, group=None, compress_fn=None):
    """Rank-local compress of one slab followed by the gather.  compress_fn(x_slab, stream_buf) -> nbytes defaults to the
    CUDA path; tests on CPU (gloo) inject a stand-in."""
    if compress_fn is None:
        from . import api
        n = api.compress_device(x_slab, stream_buf, mode, abs_eb, rel)
    else:
        n = compress_fn(x_slab, stream_buf)
    return gather_streams(stream_buf, n, group)


def container_from(sizes, parts):
    return frame([bytes(p.cpu().numpy().tobytes()) for p in parts])
