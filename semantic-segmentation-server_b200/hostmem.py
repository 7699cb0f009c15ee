"""Placement of pinned host buffers on NUMA nodes (Linux).  The host->device copies of the host-buffer
pipelines run at whatever rate the source pages can be read over the GPU's PCIe root, so on a box with
several GPUs per socket it matters which node the pages sit on.  No libnuma here: the three system calls
(mbind, set_mempolicy, move_pages) go through ctypes, and topology comes from sysfs.

    gpu_numa_node(i)            node of GPU i's PCIe function (/sys/bus/pci/devices/<bdf>/numa_node)
    allowed_nodes()             nodes this process may allocate on (Mems_allowed_list)
    PlacedPinned(nbytes, ...)   page-aligned anonymous mapping bound to nodes, touched, cudaHostRegister'ed
"""
import ctypes
import mmap
import os

import numpy as np

_libc = ctypes.CDLL(None, use_errno=True)
_SYS_MBIND, _SYS_SET_MEMPOLICY, _SYS_MOVE_PAGES = 237, 238, 279      # x86_64
MPOL_DEFAULT, MPOL_PREFERRED, MPOL_BIND, MPOL_INTERLEAVE = 0, 1, 2, 3


def _parse_list(text):
    out = []
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        out.extend(range(int(a), int(b or a) + 1))
    return out


def allowed_nodes():
    with open("/proc/self/status") as f:
        for line in f:
            if line.startswith("Mems_allowed_list:"):
                return _parse_list(line.split(":", 1)[1])
    return [0]


def node_cpus(node):
    try:
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            return set(_parse_list(f.read()))
    except OSError:
        return set()


def gpu_pci_bdf(device_index):
    import torch
    p = torch.cuda.get_device_properties(device_index)
    return "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)


def gpu_numa_node(device_index):
    """NUMA node of the GPU's PCIe function, or None when sysfs does not say (-1 on single-node boxes)."""
    try:
        with open("/sys/bus/pci/devices/%s/numa_node" % gpu_pci_bdf(device_index)) as f:
            n = int(f.read().strip())
        return n if n >= 0 else None
    except (OSError, ValueError):
        return None


def _nodemask(nodes):
    mask = 0
    for n in nodes:
        mask |= 1 << n
    words = max(1, (max(nodes) + 64) // 64)
    arr = (ctypes.c_ulong * words)(*[(mask >> (64 * i)) & (2 ** 64 - 1) for i in range(words)])
    return arr, words * 64 + 1


def pages_on_nodes(addr, nbytes, samples=64):
    """{node: count} for `samples` pages spread over [addr, addr+nbytes) (move_pages in query mode)."""
    page = mmap.PAGESIZE
    n_pages = max(1, nbytes // page)
    idx = sorted({int(i * (n_pages - 1) / max(1, samples - 1)) for i in range(min(samples, n_pages))})
    ptrs = (ctypes.c_void_p * len(idx))(*[addr + i * page for i in idx])
    status = (ctypes.c_int * len(idx))()
    rc = _libc.syscall(_SYS_MOVE_PAGES, 0, ctypes.c_ulong(len(idx)), ptrs, None, status, 0)
    if rc != 0:
        return {}
    out = {}
    for s in status:
        out[int(s)] = out.get(int(s), 0) + 1
    return out


class PlacedPinned:
    """`nbytes` of page-locked host memory whose pages were placed by `policy` before pinning:
    "bind" (all pages on `nodes`), "interleave" (round-robin over `nodes`), or "default" (first touch by
    the calling thread).  `.array` is a uint8 numpy view; `.close()` unregisters and unmaps."""

    def __init__(self, nbytes, policy="default", nodes=None):
        import torch
        self.nbytes = (nbytes + mmap.PAGESIZE - 1) // mmap.PAGESIZE * mmap.PAGESIZE
        self._map = mmap.mmap(-1, self.nbytes, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
        self.array = np.frombuffer(self._map, dtype=np.uint8)
        self.addr = self.array.ctypes.data
        self.policy_applied = "default"
        if policy in ("bind", "interleave") and nodes:
            mask, maxnode = _nodemask(nodes)
            mode = MPOL_BIND if policy == "bind" else MPOL_INTERLEAVE
            rc = _libc.syscall(_SYS_MBIND, ctypes.c_void_p(self.addr), ctypes.c_ulong(self.nbytes), mode, mask,
                               ctypes.c_ulong(maxnode), 0)
            self.policy_applied = policy if rc == 0 else "default (mbind failed: errno %d)" % ctypes.get_errno()
        # touch every page (this is where they get their node), then lock them for DMA
        self.array[::mmap.PAGESIZE] = 0
        self._registered = False
        rc = torch.cuda.cudart().cudaHostRegister(self.addr, self.nbytes, 0)
        if int(rc) != 0:
            raise RuntimeError("cudaHostRegister failed: %s" % rc)
        self._registered = True

    def nodes_of_pages(self):
        return pages_on_nodes(self.addr, self.nbytes)

    def close(self):
        import torch
        if self._registered:
            torch.cuda.cudart().cudaHostUnregister(self.addr)
            self._registered = False
        self.array = None
        try:
            self._map.close()
        except BufferError:
            pass

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
