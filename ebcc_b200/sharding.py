"""Frame-wise sharding of a batch over the GPUs of one box.

Frames are independent (reference chunk loop, src/ebcc_codec.c:1007-1046, carries no state), so rank g of G
compresses the contiguous range [g*ceil(F/G), (g+1)*ceil(F/G)) on its own GPU.  The only exchange is a host-side
gather of the per-frame stream sizes, turned into EBCK container offsets (u64 length prefix + stream,
src/ebcc_codec.c:1037-1038) by an exclusive prefix sum.  No collective touches frame data."""
import struct

EBCK_HEADER_BYTES = 80


def frame_range(nframes, rank, world):
    per = -(-nframes // world)
    lo = min(nframes, rank * per)
    return lo, min(nframes, lo + per)


def container_offsets(sizes):
    """byte offset of every chunk's length prefix inside the EBCK container, and the container size"""
    offs, o = [], EBCK_HEADER_BYTES
    for s in sizes:
        offs.append(o)
        o += 8 + int(s)
    return offs, o


def gather_sizes(local_sizes, group=None):
    """all ranks' per-frame sizes in frame order (torch.distributed object gather; works on gloo and nccl)"""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return list(local_sizes)
    parts = [None] * dist.get_world_size(group)
    dist.all_gather_object(parts, list(local_sizes), group=group)
    return [s for p in parts for s in p]


def ebck_header(dims, chunk_dims, num_chunks, chunk_size):
    return struct.pack("<4sIII3Q3QQQ", b"EBCK", 1, 3, 0, *dims, *chunk_dims, num_chunks, chunk_size)
