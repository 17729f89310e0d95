"""Sharding one frame over the GPUs of a box: one process per GPU (torch.distributed).

Pixels are independent (camera.zig:116-121: each pixel only reads the immutable world), so the frame
partitions with no collective during rendering.  Rank r renders the interleaved `tile_rows`-row tiles
r, r+N, r+2N, ... (load balance: neighbouring rows cost about the same).  The only exchange of the path
is the framebuffer, assembled on rank 0 in one of two ways:

  PeerFrame    rank 0's canvas is symmetric memory mapped into every rank's address space over NVLink;
               each rank's render kernel stores its pixels straight into it (rtc_render_device_frame), so
               the "collective" is fused into the kernel's own epilogue and costs one device barrier.
  GatherFrame  each rank renders packed rows locally, one gather (NCCL on GPUs, gloo on CPU tests) brings
               them to rank 0, and an index copy de-interleaves them.  Portable fallback, and the
               host-side logic the CPU tests exercise.

Neither class renders anything itself: both only own buffers and the exchange.

The scene is parsed ONCE: broadcast_scene sends rank 0's flattened blob and camera to every rank (NCCL broadcast on
GPUs, gloo on the CPU tests) -- no rank but 0 reads a JSON, OBJ or PNG file.  With PeerFrame the ranks do not keep fixed
row partitions either: TileShare gives every rank's scene the peer-mapped cost buffers of all ranks
(rtc_scene_share_tiles), and the tiles of each frame are dealt by what they cost in the frame before.
"""
import torch
import torch.distributed as dist


def partition_rows(vsize, tile_rows, part, n_parts):
    """The y's of partition `part`, increasing -- Python mirror of rtc_partition_rows (rtc_api.cu)."""
    return [y for y in range(vsize) if (y // tile_rows) % n_parts == part]


class GatherFrame:
    """Packed partition rows -> whole frame on rank 0 through one torch.distributed gather."""

    def __init__(self, width, height, tile_rows, rank, world, device, dtype=torch.float64, channels=3):
        self.rank, self.world, self.device = rank, world, device
        self.rows = [partition_rows(height, tile_rows, r, world) for r in range(world)]
        self.max_rows = max(len(r) for r in self.rows)
        # every rank sends the same shape (gather needs equal sizes); short partitions are padded
        self.local = torch.zeros((self.max_rows, width, channels), dtype=dtype, device=device)
        self.frame = torch.zeros((height, width, channels), dtype=dtype, device=device) if rank == 0 else None
        self.parts = [torch.zeros_like(self.local) for _ in range(world)] if (rank == 0 and world > 1) else None
        self.index = [torch.tensor(r, dtype=torch.long, device=device) for r in self.rows] if rank == 0 else None

    def local_ptr(self):
        return self.local.data_ptr()

    def assemble(self):
        """Call on every rank after rendering into `local`.  Returns the frame on rank 0, None elsewhere."""
        if self.world == 1:
            return self.local[: len(self.rows[0])]  # one partition holding every row in order: already the frame
        dist.gather(self.local, self.parts, dst=0)
        if self.rank != 0:
            return None
        src = self.parts
        for r in range(self.world):
            n = len(self.rows[r])
            if n:
                self.frame.index_copy_(0, self.index[r], src[r][:n])
        return self.frame


class PeerFrame:
    """Rank 0's frame as symmetric memory: every rank gets a device pointer to it (NVLink peer mapping)."""

    def __init__(self, width, height, rank, world, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.rank, self.world = rank, world
        n = height * width * 3
        self.buf = symm_mem.empty(n, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.frame = self.buf.view(height, width, 3) if rank == 0 else None
        self.root_ptr = int(self.handle.buffer_ptrs[0])  # rank 0's buffer as THIS device addresses it

    def finish(self):
        """Device-side barrier on the current stream: when it returns on rank 0's stream, every rank's stores
        into the frame have been issued and completed."""
        self.handle.barrier()
        return self.frame


def broadcast_scene(blob, cam_struct, rank, device, group=None):
    """rank 0's (blob bytes, ctypes rtc_camera) -> the same pair on every rank.  The blob travels as a uint8 tensor on
    `device` (NCCL over NVLink on GPUs), the 168-byte camera with it."""
    import ctypes
    head = torch.zeros(2, dtype=torch.int64, device=device)
    if rank == 0:
        cam_bytes = ctypes.string_at(ctypes.byref(cam_struct), ctypes.sizeof(cam_struct))
        head[0], head[1] = len(blob), len(cam_bytes)
    dist.broadcast(head, src=0, group=group)
    n_blob, n_cam = int(head[0]), int(head[1])
    buf = torch.empty(n_blob + n_cam, dtype=torch.uint8, device=device)
    if rank == 0:
        buf.copy_(torch.frombuffer(bytearray(blob + cam_bytes), dtype=torch.uint8))
    dist.broadcast(buf, src=0, group=group)
    if rank == 0:
        return blob, cam_struct
    host = buf.cpu().numpy().tobytes()
    import importlib
    cam = importlib.import_module(__package__).rtc_camera()
    assert n_cam == ctypes.sizeof(cam)
    ctypes.memmove(ctypes.byref(cam), host[n_blob:], n_cam)
    return host[:n_blob], cam


class TileShare:
    """Two symmetric cost buffers per rank, every rank's mapped into every rank (rtc_scene_share_tiles).  Collective."""

    def __init__(self, world_obj, width, height, rank, world, device, pkg, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.words = int(pkg.lib().rtc_tile_share_words(width, height))
        self.buf = symm_mem.empty(2 * self.words, dtype=torch.int32, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        ptrs = [[int(self.handle.buffer_ptrs[r]) + b * self.words * 4 for r in range(world)] for b in range(2)]
        world_obj.share_tiles(rank, world, self.words, ptrs)
        torch.cuda.synchronize(device)
        dist.barrier(group=group)  # every rank's buffers are clean before anyone renders


class SharedHostFrame:
    """The caller's HOST canvas shared by all ranks of the box: one POSIX shared-memory segment, mapped by every
    process and (on GPUs) page-locked + device-mapped with cudaHostRegister.  Each rank then calls the public
    Camera.render(world, partition=..., out=frame.array): rtc_render recognises the mapped memory and its kernel
    stores the rank's rows straight into the host frame over that GPU's OWN PCIe link.  No assembly on rank 0,
    no device->host copy of the whole frame through one link: the end-to-end exchange is a barrier."""

    def __init__(self, width, height, rank, world, register=True, group=None):
        import numpy as np
        from multiprocessing import shared_memory
        self.rank, self.world = rank, world
        nbytes = height * width * 3 * 8
        name = [None]
        if rank == 0:
            self.shm = shared_memory.SharedMemory(create=True, size=nbytes + 4096)  # + one arrival counter per rank
            name[0] = self.shm.name
        if world > 1:
            dist.broadcast_object_list(name, src=0, group=group)
        if rank != 0:
            self.shm = shared_memory.SharedMemory(name=name[0])
            try:  # the segment belongs to rank 0: keep this process's resource tracker from unlinking it at exit
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:  # noqa: BLE001
                pass
        self.array = np.ndarray((height, width, 3), dtype=np.float64, buffer=self.shm.buf)
        self.flags = np.ndarray((world,), dtype=np.int64, buffer=self.shm.buf, offset=nbytes)
        self.round = 0
        if rank == 0:
            self.array[:] = 0.0
            self.flags[:] = 0
        self.registered = False
        if register:
            rt = torch.cuda.cudart()
            # cudaHostRegisterPortable (1) | cudaHostRegisterMapped (2)
            err = rt.cudaHostRegister(self.array.ctypes.data, nbytes, 3)
            if int(err) != 0:
                raise RuntimeError("cudaHostRegister failed: %s" % err)
            self.registered = True
        if world > 1:
            dist.barrier(group=group)

    def barrier(self):
        """Host-side barrier of the ranks sharing the canvas: every rank publishes how many frames it has delivered in
        its own slot of the segment and waits until all slots have caught up (microseconds; an NCCL barrier is a kernel
        launch and a stream synchronisation per rank).  Call after Camera.render has returned (its frame rows are in)."""
        self.round += 1
        self.flags[self.rank] = self.round
        while int(self.flags.min()) < self.round:
            pass

    def close(self):
        if self.registered:
            torch.cuda.cudart().cudaHostUnregister(self.array.ctypes.data)
            self.registered = False
        self.array = None
        self.flags = None
        self.shm.close()
        if self.rank == 0:
            self.shm.unlink()
