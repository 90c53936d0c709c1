"""Multi-GPU slab sharding: the slowest dimension is cut into G contiguous, independent SZ slabs, one per
rank / GPU, exactly like the reference's slab-parallel example (example/test_threadsafe.c:40-58).  Each slab is a complete
SZ stream (bit-exact to the reference called on that slab, decodable by the reference).  The only exchange step is an
all-gather of the stream sizes followed by the streams themselves at their exact lengths (one broadcast per rank into
one buffer); the container

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


def gather_streams(local_stream, local_size, group=None, out=None):
    """All-gather of per-rank streams of different lengths.  local_stream: 1-D uint8 torch tensor (CUDA with NCCL, CPU with gloo)
    holding at least local_size valid bytes.  One all-gather of the eight-byte sizes into one tensor, read back with ONE host
    synchronisation; then every rank's stream travels at its exact length into its place in one buffer (`out`, allocated here if
    not given or too small): one broadcast per rank, all enqueued before the first is waited for.  No padding to the longest
    stream, no list outputs.  Returns (sizes list, list of per-rank views into the buffer)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    me = dist.get_rank(group)
    dev = local_stream.device
    mine = torch.tensor([int(local_size)], dtype=torch.int64, device=dev)
    all_sz = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(all_sz, mine, group=group)         # 8 bytes per rank
    sizes = [int(v) for v in all_sz.tolist()]
    offs = [0]
    for v in sizes:
        offs.append(offs[-1] + v)
    if out is None or out.numel() < offs[-1] or out.device != dev:
        out = torch.empty(max(offs[-1], 1), dtype=torch.uint8, device=dev)
    parts = [out[offs[r]:offs[r + 1]] for r in range(world)]
    parts[me].copy_(local_stream[:sizes[me]])
    works = []
    for r in range(world):
        if sizes[r]:
            src = dist.get_global_rank(group, r) if group is not None else r
            works.append(dist.broadcast(parts[r], src=src, group=group, async_op=True))
    for w in works:
        w.wait()
    return sizes, parts


def compress_slab_and_gather(x_slab, stream_buf, mode, abs_eb, rel, group=None, compress_fn=None, out=None):
    """Rank-local compress of one slab followed by the gather.  compress_fn(x_slab, stream_buf) -> nbytes defaults to the
    CUDA path; tests on CPU (gloo) inject a stand-in.  (The CUDA path orders itself after torch's current stream, api.py, so a
    collective still reading stream_buf from the previous call is waited for on the device.)"""
    if compress_fn is None:
        from . import api
        n = api.compress_device(x_slab, stream_buf, mode, abs_eb, rel)
    else:
        n = compress_fn(x_slab, stream_buf)
    return gather_streams(stream_buf, n, group, out)


def container_from(sizes, parts):
    return frame([bytes(p.cpu().numpy().tobytes()) for p in parts])
