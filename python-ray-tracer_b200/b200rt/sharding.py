"""Host-side partitioning of the render path over GPUs; pure Python / numpy, no CUDA.

A frame is cut along array axis 1 (`x`) into strips of `strip` columns; strip s belongs to rank s % world
(interleaved for load balance: per-pixel cost is not uniform).  Per channel a strip is one contiguous byte range of
the (3,W,H) uint8 framebuffer, so a gather is three contiguous copies per strip.  The same rule is implemented in
libb200rt (`strip_width/strip_rank/strip_world` of b200rt_args); tests/test_sharding.py checks the two agree on a GPU
and checks this module alone with world_size-2 gloo processes on CPU.
"""
import numpy as np

DEFAULT_STRIP = 16


def strip_ranges(W, rank, world, strip=DEFAULT_STRIP, x_begin=0, x_end=None):
    """[(x0, x1)] column ranges of `rank` inside the tile [x_begin, x_end)."""
    x_end = W if x_end is None else x_end
    if not (0 <= rank < world) or strip <= 0:
        raise ValueError("bad rank/world/strip")
    out = []
    for s, x0 in enumerate(range(x_begin, x_end, strip)):
        if s % world == rank:
            out.append((x0, min(x_end, x0 + strip)))
    return out


def columns_of_rank(W, rank, world, strip=DEFAULT_STRIP):
    r = strip_ranges(W, rank, world, strip)
    return np.concatenate([np.arange(a, b) for a, b in r]) if r else np.zeros(0, dtype=np.int64)


def frames_of_rank(n_frames, rank, world):
    """Camera-pose batch sharding: frame k renders on rank k % world."""
    return list(range(rank, n_frames, world))


def pack_strips(frame, rank, world, strip=DEFAULT_STRIP):
    """(3,W,H) array -> this rank's columns as a contiguous (3, ncols, H) array (the all-gather payload)."""
    return np.ascontiguousarray(frame[:, columns_of_rank(frame.shape[1], rank, world, strip), :])


def unpack_strips(parts, W, strip=DEFAULT_STRIP, out=None):
    """Inverse of pack_strips over all ranks: parts[r] is rank r's (3, ncols_r, H) payload (extra padding columns allowed)."""
    world = len(parts)
    H = parts[0].shape[2]
    if out is None:
        out = np.zeros((3, W, H), dtype=parts[0].dtype)
    for r, p in enumerate(parts):
        cols = columns_of_rank(W, r, world, strip)
        out[:, cols, :] = p[:, :len(cols), :]
    return out


def allgather_frame(result, rank, world, strip=DEFAULT_STRIP, group=None):
    """torch.distributed all-gather of the ranks' strips into the full frame on every rank.

    `result` is this rank's (3,W,H) uint8 torch tensor (CUDA with NCCL, CPU with gloo) whose own strips are rendered.
    Payloads are padded to the largest rank's column count so that one fixed-size all_gather suffices."""
    import torch
    import torch.distributed as dist
    W = result.shape[1]
    cols = [torch.as_tensor(columns_of_rank(W, r, world, strip), device=result.device) for r in range(world)]
    nmax = max(len(c) for c in cols)
    mine = torch.zeros((3, nmax, result.shape[2]), dtype=result.dtype, device=result.device)
    mine[:, :len(cols[rank]), :] = result.index_select(1, cols[rank])
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    for r in range(world):
        if r != rank:
            result.index_copy_(1, cols[r], parts[r][:, :len(cols[r]), :])
    return result


class SharedFrame:
    """One (3,W,H) uint8 host frame shared by the ranks of a multi-GPU job: the direct device-to-host gather of SURVEY
    §8(e) across PROCESSES.  Every rank maps the same /dev/shm file and pins the mapping (b200rt_host_register); its
    b200rt_render_host call then copies its strips straight into the frame, so after a barrier the creating rank (and
    every other) holds the assembled frame — the `result.copy_to_host()` of reference src/main.py:51 for N GPUs, with no
    extra copy and no collective.

        fr = SharedFrame(name, W, H, create=(rank == 0)); barrier()         # rank 0 creates, then everyone attaches
        render_host(..., out=fr.array, strips=(16, rank, world)); barrier()  # fr.array is now the whole frame
    """

    def __init__(self, name, W, H, create, pin=True, directory="/dev/shm"):
        import os
        self.path = os.path.join(directory, name)
        self.created = bool(create)
        n = 3 * int(W) * int(H)
        if create:
            with open(self.path, "wb") as f:
                f.truncate(n)
        self.array = np.memmap(self.path, dtype=np.uint8, mode="r+", shape=(3, int(W), int(H)))
        self.pinned = False
        if pin:
            from . import _lib
            _lib.check(_lib.lib().b200rt_host_register(self.array.ctypes.data, n))
            self.pinned = True

    def close(self):
        import os
        if self.pinned:
            from . import _lib
            _lib.lib().b200rt_host_unregister(self.array.ctypes.data)
            self.pinned = False
        self.array = None
        if self.created and os.path.exists(self.path):
            os.unlink(self.path)
