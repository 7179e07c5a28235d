"""Multi-GPU host logic: replicated cloud, query shards by Morton range.

One process per GPU.  Rank 0 owns the scan; `broadcast_cloud` replicates it (NCCL over NVLink on the GPU box,
gloo in the CPU tests), every rank builds the identical Morton order (the device radix sort is deterministic),
answers the contiguous range `shard_range(n, rank, world)` of the SORTED order, and `gather_rows` all-gathers
the per-point rows; `unpermute` (ccg_unpermute on the GPU) scatters them back to original order.  There is no
other exchange on the path: queries are independent given read-only access to the whole cloud.

This module holds no compute: the per-range work is a callable so that the CPU tests (gloo, world_size 2) can
drive the same logic with the oracle standing in for the GPU.
"""
import torch
import torch.distributed as dist

LEAF = 32   # ccg ranges must begin on a leaf boundary (include/ccgpu.h, ccg_*_range)


def shard_range(n, rank, world):
    """-> (begin, end, per): sorted positions [begin, end) of `rank`; per = rows every rank contributes to the
    all-gather (a multiple of 32; the last ranks may own fewer or no points)."""
    nleaf = (n + LEAF - 1) // LEAF
    per = (nleaf + world - 1) // world * LEAF
    if rank * per >= n:                      # more ranks than leaves: an empty range on a leaf boundary
        b = n // LEAF * LEAF
        return b, b, per
    return rank * per, min(n, (rank + 1) * per), per


def broadcast_cloud(points, src=0):
    """replicate the [n, S] float32 cloud from `src` (in place on the other ranks)"""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(points, src)
    return points


def gather_rows(local_rows, per, world):
    """all-gather equal-sized [per, C] row blocks -> [world*per, C] (sorted order, tail rows beyond n unused)"""
    if world == 1:
        return local_rows
    out = torch.empty((world * per,) + tuple(local_rows.shape[1:]), dtype=local_rows.dtype, device=local_rows.device)
    dist.all_gather_into_tensor(out, local_rows.contiguous())
    return out


def sharded_rows(n, compute_range, width, dtype=torch.float32, device="cpu"):
    """Run `compute_range(begin, end, out_rows)` on this rank's shard and return the gathered [world*per, width]
    rows in sorted order.  `out_rows` is a [per, width] buffer whose first end-begin rows must be filled."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    begin, end, per = shard_range(n, rank, world)
    local = torch.zeros((per, width), dtype=dtype, device=device)
    if end > begin:
        compute_range(begin, end, local)
    return gather_rows(local, per, world)


# ---- fused query + scatter: the collector's output arrays opened as NVLink peer memory on every rank -----------
class DevBuffer:
    """A device array at a raw address (own cudaMalloc on the collector, IPC-mapped peer memory elsewhere)."""

    def __init__(self, addr, shape, owner):
        self.addr, self.shape, self.owner = addr, tuple(shape), owner
        self.__cuda_array_interface__ = {"shape": self.shape, "typestr": "<f4", "data": (addr, False), "version": 3, "strides": None}

    def tensor(self, device):
        """torch view (collector only: the memory lives on its GPU)"""
        return torch.as_tensor(self, device=device)


class PeerRows:
    """Row arrays [n, w] (float32) that live on rank `collector` and are writable from every local rank.

    The collector allocates them through the library (plain cudaMalloc: an IPC handle names an allocation base),
    exports one 64-byte handle per array, the handles travel through torch.distributed (any backend), and the other
    ranks map them with ccg_ipc_open.  ccg_pca_knn_features_scatter then stores finished rows straight into the
    collector's memory over NVLink from inside the query kernel."""

    def __init__(self, ctx, n, widths, collector=0):
        self.ctx, self.collector = ctx, collector
        rank = dist.get_rank() if dist.is_initialized() else 0
        world = dist.get_world_size() if dist.is_initialized() else 1
        self.is_collector = rank == collector
        self.bufs = []
        handles = []
        if self.is_collector:
            for w in widths:
                addr = ctx.dev_alloc(max(1, n * w * 4))
                self.bufs.append(DevBuffer(addr, (n, w), True))
                handles.append(ctx.ipc_export(addr) if world > 1 else b"")
        if world > 1:
            box = [handles]
            dist.broadcast_object_list(box, src=collector)
            if not self.is_collector:
                for w, h in zip(widths, box[0]):
                    self.bufs.append(DevBuffer(ctx.ipc_open(h), (n, w), False))

    def close(self):
        for b in self.bufs:
            if b.owner:
                self.ctx.dev_free(b.addr)
            else:
                self.ctx.ipc_close(b.addr)
        self.bufs = []


class PeerSlices:
    """Row arrays partitioned by ORIGINAL index: every rank owns the [rows_per_rank, w] slice of each array and maps all
    the other ranks' slices as NVLink peer memory (ccg_ipc_export / ccg_ipc_open), so that ccg_pca_knn_features_route can
    store every finished row with its owner.  tables[a][r] = device address of array a's slice on rank r."""

    def __init__(self, ctx, rows_per_rank, widths):
        self.ctx = ctx
        rank = dist.get_rank() if dist.is_initialized() else 0
        world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank, self.world, self.rows = rank, world, rows_per_rank
        self.own = [DevBuffer(ctx.dev_alloc(max(1, rows_per_rank * w * 4)), (rows_per_rank, w), True) for w in widths]
        self.tables = [[None] * world for _ in widths]
        self._opened = []
        handles = [ctx.ipc_export(b.addr) if world > 1 else b"" for b in self.own]
        every = [None] * world
        if world > 1:
            dist.all_gather_object(every, handles)
        for a in range(len(widths)):
            for r in range(world):
                if r == rank:
                    self.tables[a][r] = self.own[a].addr
                else:
                    addr = ctx.ipc_open(every[r][a])
                    self._opened.append(addr)
                    self.tables[a][r] = addr

    def close(self):
        for addr in self._opened:
            self.ctx.ipc_close(addr)
        for b in self.own:
            self.ctx.dev_free(b.addr)
        self._opened, self.own = [], []


class SharedPinned:
    """One host array shared by the local ranks (POSIX shared memory under /dev/shm) and page-locked in every process
    (cudaHostRegister), so that each rank moves ITS slice of the input / output through its own PCIe link."""

    def __init__(self, name, shape, dtype=torch.float32, create=False, register=True):
        import os
        self.path = f"/dev/shm/{name}"
        nelem = 1
        for d in shape:
            nelem *= d
        self.nbytes = nelem * torch.empty((), dtype=dtype).element_size()
        if create:
            with open(self.path, "wb") as f:
                f.truncate(max(self.nbytes, 1))
        self.created = create
        self.t = torch.from_file(self.path, shared=True, size=nelem, dtype=dtype).view(shape)
        self.registered = False
        if register:
            self.register()

    def register(self):
        """page-lock the mapping in this process.  Pages that were never touched are allocated HERE, on the caller's NUMA
        node: let every rank write its own slice first (register=False, touch, barrier, register) so that each GPU's copies
        go to memory next to its own PCIe root."""
        if not self.registered and torch.cuda.is_available() and self.nbytes:
            rc = torch.cuda.cudart().cudaHostRegister(self.t.data_ptr(), self.nbytes, 0)
            self.registered = int(rc) == 0
        return self.registered

    def close(self):
        import os
        if self.registered:
            torch.cuda.cudart().cudaHostUnregister(self.t.data_ptr())
            self.registered = False
        self.t = None
        if self.created and os.path.exists(self.path):
            os.unlink(self.path)
