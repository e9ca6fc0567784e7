"""Multi-GPU plumbing for the shift-and-stack search: one process per GPU, start pixels sharded in row bands,
psi/phi replicated, per-GPU sorted results merged by an NCCL all-gather + global sort (in the C library:
``kb_merge_results_allgather``).  The reference has no multi-GPU code.

torch.distributed is used only as the control plane (rendezvous, barrier, max-reduction of timings and the
128-byte NCCL unique id); the data path is the library's own NCCL communicator.
"""
import ctypes as C
import os

import numpy as np

from ._lib import TRAJECTORY_DTYPE, check, lib


def env_rank():
    """(rank, world_size, local_rank) from the torchrun environment (defaults: single process)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def set_device(local_rank):
    check(lib.kb_set_device(C.c_int(local_rank)))


def parse_cpulist(text):
    """The CPU numbers of a sysfs cpulist string ("0-31,64-95")."""
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_to_device_cpus():
    """Restrict this process to the CPUs of its GPU's own socket (sysfs ``local_cpulist`` of the current device).

    Call it after ``set_device`` and before the first stack is built: the threads the library starts later inherit
    the mask, so the host copy of the stack (first touch) and the pinned staging buffers land in the memory the
    GPU's PCIe root reads without crossing sockets.  With one process per GPU every socket then serves its own GPUs.
    Returns the sorted CPU list it bound to, or None where the platform reports nothing narrower than the current
    mask (virtual machines, single-socket hosts).  No reference counterpart (the reference is single-GPU).
    """
    buf = C.create_string_buffer(32)
    check(lib.kb_device_pci_bus_id(buf, C.c_int(32)))
    path = f"/sys/bus/pci/devices/{buf.value.decode().lower()}/local_cpulist"
    try:
        with open(path) as f:
            local = parse_cpulist(f.read())
        now = os.sched_getaffinity(0)
        want = local & now
        if not want or want == now:
            return None
        os.sched_setaffinity(0, want)
        return sorted(want)
    except (OSError, ValueError, AttributeError):
        return None


def row_bands(y_min, y_max, world):
    """Contiguous start-pixel row bands [y_k, y_{k+1}) covering [y_min, y_max), as even as possible."""
    n = y_max - y_min
    edges = [y_min + (n * k) // world for k in range(world + 1)]
    return [(edges[k], edges[k + 1]) for k in range(world)]


def band_rows(band, vy, zeroed_times, height, margin=2):
    """Image rows [lo, hi) a rank needs psi/phi for when it searches the start rows ``band`` = (y0, y1) with candidate
    y-velocities ``vy``: the band plus the candidates' reach max|vy| * max|t| (``StackSearch(..., rows=...)`` adds the
    PSF radius for the convolution's inputs itself)."""
    vmax = float(np.max(np.abs(np.asarray(vy, dtype=np.float32)))) if len(vy) else 0.0
    tabs = float(np.max(np.abs(np.asarray(zeroed_times, dtype=np.float64)))) if len(zeroed_times) else 0.0
    reach = int(np.ceil(vmax * tabs)) + 1 + margin
    return max(0, band[0] - reach), min(int(height), band[1] + reach)


def merge_sorted_lists(lists, top_n=0):
    """Host statement of the merge rule (used by the CPU tests and as documentation of the device merge):
    concatenate the per-rank lists in rank order and stable-sort by lh descending; keep the first top_n."""
    cat = np.concatenate([np.asarray(a, dtype=TRAJECTORY_DTYPE) for a in lists]) if lists else np.zeros(0, TRAJECTORY_DTYPE)
    order = np.argsort(-cat["lh"].astype(np.float64), kind="stable")
    out = cat[order]
    return out[:top_n] if top_n else out


def merge_sorted_lists_device(lists, top_n=0):
    """The same merge on the device (``kb_merge_sorted_lists``: the kernel the NCCL merge runs on the gathered
    blocks), from host arrays."""
    from ._lib import ptr, ptr_array
    arrs = [np.ascontiguousarray(a, dtype=TRAJECTORY_DTYPE) for a in lists]
    counts = np.array([a.size for a in arrs], dtype=np.uint64)
    total = int(counts.sum())
    keep = min(total, top_n) if top_n else total
    out = np.zeros(max(keep, 1), dtype=TRAJECTORY_DTYPE)
    n = C.c_uint64(0)
    check(lib.kb_merge_sorted_lists(ptr_array(arrs), ptr(counts), C.c_int(len(arrs)), C.c_uint64(top_n), ptr(out),
                                    C.byref(n)))
    return out[: n.value]


class Comm:
    """Control-plane communicator (gloo) + the library's NCCL communicator for the data path."""

    def __init__(self, rank, world, init_nccl=True, backend="gloo"):
        import torch
        import torch.distributed as td
        self.torch, self.td, self.rank, self.world = torch, td, rank, world
        if not td.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            td.init_process_group(backend, rank=rank, world_size=world)
        self.nccl = False
        if init_nccl:
            uid = torch.zeros(128, dtype=torch.uint8)
            if rank == 0:
                buf = (C.c_uint8 * 128)()
                check(lib.kb_nccl_unique_id(buf))
                uid = torch.tensor(list(buf), dtype=torch.uint8)
            td.broadcast(uid, src=0)
            buf = (C.c_uint8 * 128)(*uid.tolist())
            # NCCL may print its version banner on stdout (NCCL_DEBUG=VERSION); callers such as bench.py promise
            # exactly one JSON line there, so route fd 1 to stderr while the communicator comes up.
            import sys
            sys.stdout.flush()
            saved = os.dup(1)
            try:
                os.dup2(2, 1)
                check(lib.kb_nccl_init(buf, C.c_int(rank), C.c_int(world)))
            finally:
                os.dup2(saved, 1)
                os.close(saved)
            self.nccl = True

    def barrier(self):
        self.td.barrier()

    def max_float(self, v):
        t = self.torch.tensor([float(v)], dtype=self.torch.float64)
        self.td.all_reduce(t, op=self.td.ReduceOp.MAX)
        return float(t.item())

    def all_gather_arrays(self, arr):
        """Gather one structured numpy array per rank (control-plane path, used by the CPU tests)."""
        out = [None] * self.world
        self.td.all_gather_object(out, np.asarray(arr))
        return out

    def finalize(self):
        if self.nccl:
            check(lib.kb_nccl_finalize())
            self.nccl = False
        if self.td.is_initialized():
            self.td.destroy_process_group()


def merge_results(search, top_n=0):
    """All-gather every rank's sorted result list over NCCL and leave the global best `top_n` (0 = all) as the
    result list of `search` on every rank.  Returns the merged count."""
    n = C.c_uint64(0)
    check(lib.kb_merge_results_allgather(search._ppa._require(), C.c_uint64(top_n), C.byref(n)))
    return n.value


def make_banded_search(sci_imgs, var_imgs, psf_kernels, zeroed_times, candidates_vy, rank, world, num_bytes=-1,
                       y_bounds=None):
    """The StackSearch of one rank of a multi-GPU search: start rows ``y_bounds`` (default: the whole image) are
    split into ``world`` contiguous bands, this rank's band becomes the search's y start bounds, and -- for float32
    stacks -- psi/phi is built only for the image rows that band can reach with the given candidate y-velocities.
    Returns ``(search, band)``; run ``search.search_all(candidates, True)`` and then ``merge_results(search)``."""
    from .search import StackSearch
    height = int(np.shape(sci_imgs[0])[0])
    y0, y1 = (0, height) if y_bounds is None else y_bounds
    band = row_bands(y0, y1, world)[rank]
    rows = None
    if num_bytes not in (1, 2) and band[0] < band[1]:   # the scale of an encoded stack depends on every row
        rows = band_rows(band, candidates_vy, zeroed_times, height)
    s = StackSearch(sci_imgs, var_imgs, psf_kernels, zeroed_times, num_bytes, rows=rows)
    if band[0] < band[1]:
        s.set_start_bounds_y(band[0], band[1])
    return s, band


def sharded_search(search, candidates, comm, top_n=0):
    """search_all over the full start bounds of `search`, sharded in row bands over comm.world ranks."""
    p = search.params
    y0, y1 = p.y_start_min, p.y_start_max
    band = row_bands(y0, y1, comm.world)[comm.rank]
    try:
        if band[0] < band[1]:
            search.set_start_bounds_y(band[0], band[1])
            search.search_all(candidates, True)
        else:
            search.clear_results()
    finally:
        p.y_start_min, p.y_start_max = y0, y1
    return merge_results(search, top_n)
