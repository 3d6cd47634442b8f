"""Mini-batch blocks on the device: `sample_neighbors`, `in_subgraph`,
`to_block` with the reference's meaning (python/dgl/sampling/neighbor.py, python/dgl/transform.py
in_subgraph / to_block; used by examples/pytorch/graphsage/train_sampling_multi_gpu.py:22-38),
so the sampled trainers run entirely on the GPU: sampling is one kernel over the in-CSR
(csrc/sample.cu), compaction is one hash-table pass (csrc/block.cu), and the block is a bipartite
`dgl_hack_b200.graph.Graph` consumed by the same kernels as the full graph.

    frontier = sample_neighbors(g, seeds, fanout, replace=True)
    block = to_block(frontier, seeds)            # block.srcdata[NID], dstdata[NID], edata[EID]
    h = layer(block, (x[block.srcdata[NID]], x[block.dstdata[NID]]))

Differences from the reference are confined to the random stream (a counter-based hash instead of
the host RNG): which neighbours are drawn differs, the distribution and every structural rule do
not (fanout picks per seed with replacement, min(deg, fanout) distinct ones without, nothing for
zero in-degree; block sources = destinations first, then new sources by first appearance).
"""
import ctypes

import torch

from ._lib import LIB, DGLError, check_call
from .graph import Graph

NID = "_ID"
EID = "_ID"

__all__ = ["NID", "EID", "Frontier", "sample_neighbors", "in_subgraph", "to_block",
           "compact_block"]


class Frontier(object):
    """Sampled edges in parent ids, grouped by seed: src, dst, eid (int64 device tensors)."""

    def __init__(self, src, dst, eid, num_nodes, seeds=None, dst_local=None):
        self.src, self.dst, self.eid, self.num_nodes = src, dst, eid, num_nodes
        # set by sample_neighbors: the seeds and, per edge, the index of its seed -- lets to_block
        # skip the destination lookup when it is asked for a block over the same seeds
        self.seeds, self.dst_local = seeds, dst_local

    def number_of_edges(self):
        return int(self.src.numel())

    def all_edges(self):
        return self.src, self.dst


_counter = [0]


def _gidx(g):
    return g.gidx if hasattr(g, "gidx") else g


def sample_neighbors(g, seeds, fanout, replace=False, rng_seed=None):
    """`fanout` in-neighbours of every seed, uniformly (dgl.sampling.sample_neighbors with
    edge_dir='in', prob=None). fanout = -1 selects every in-edge. `rng_seed` makes a call
    reproducible; by default a per-process counter seeded from torch's generator is used."""
    gidx = _gidx(g)
    dev = gidx.device
    seeds = torch.as_tensor(seeds, dtype=torch.int64, device=dev).contiguous()
    if seeds.dim() != 1:
        raise DGLError("seeds must be a 1-D tensor of node ids")
    if fanout == 0:
        raise DGLError("fanout must be positive (or -1 for all in-edges)")
    n = int(seeds.numel())
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    offsets = torch.empty(n + 2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        check_call(LIB.DGLB200SampleCounts(
            gidx.handle, ctypes.c_void_p(seeds.data_ptr()), n, int(fanout), int(bool(replace)),
            ctypes.c_void_p(offsets.data_ptr()), stream))
    total, bad = offsets[n:].tolist()                                       # the one host sync
    if bad & 1:
        raise DGLError("seed id out of range")
    if bad & 2:
        raise DGLError("sample_neighbors: more than 2^31 - 1 sampled edges in one call")
    out = torch.empty((4, max(total, 1)), dtype=torch.int64, device=dev)
    if rng_seed is None:
        if _counter[0] == 0:
            _counter[0] = int(torch.randint(1, 2 ** 62, (1,)).item())
        _counter[0] += 1
        rng_seed = _counter[0]
    with torch.cuda.device(dev):
        check_call(LIB.DGLB200SampleNeighbors(
            gidx.handle, ctypes.c_void_p(seeds.data_ptr()), n, int(fanout),
            int(bool(replace)), int(rng_seed) & (2 ** 64 - 1), ctypes.c_void_p(offsets.data_ptr()),
            ctypes.c_void_p(out[0].data_ptr()), ctypes.c_void_p(out[1].data_ptr()),
            ctypes.c_void_p(out[2].data_ptr()), ctypes.c_void_p(out[3].data_ptr()), stream))
    out = out[:, :total]
    return Frontier(out[0], out[1], out[2], gidx.num_dst, seeds, out[3])


def in_subgraph(g, nodes):
    """All in-edges of `nodes` (dgl.in_subgraph): grouped by node, edge-id order inside a node."""
    return sample_neighbors(g, nodes, -1)


def compact_block(src, dst, dst_nodes, dst_local=None):
    """Relabelling of dgl.to_block on plain tensors (any device): returns (src_nid, src_local,
    dst_local) -- block source nodes are `dst_nodes` first, then the other sources by first
    appearance in `src`; edges keep their order. `dst_local` (index into dst_nodes per edge), when
    already known, skips the destination lookup and its checks."""
    dev = src.device
    dst_nodes = torch.as_tensor(dst_nodes, dtype=torch.int64, device=dev)
    n_dst = int(dst_nodes.numel())
    if dev.type == "cuda":
        return _compact_block_device(src, dst, dst_nodes, dst_local)
    # host tensors (partition tooling, CPU tests): the same relabelling with torch ops
    allv = torch.cat([dst_nodes, src])
    uniq, inv = torch.unique(allv, return_inverse=True)
    if dst_local is None and n_dst and int(torch.unique(dst_nodes).numel()) != n_dst:
        raise DGLError("to_block: destination nodes must be unique")
    first = torch.full((uniq.numel(),), allv.numel(), dtype=torch.int64, device=dev)
    first.scatter_reduce_(0, inv, torch.arange(allv.numel(), device=dev), "amin")
    order = torch.argsort(first)                      # unique values by first appearance
    rank = torch.empty_like(order)
    rank[order] = torch.arange(order.numel(), device=dev)
    src_nid = uniq[order]
    src_local = rank[inv[n_dst:]]
    if dst_local is not None:
        return src_nid, src_local, dst_local
    pos = torch.searchsorted(uniq, dst).clamp(max=max(int(uniq.numel()) - 1, 0))
    if dst.numel() and not torch.equal(uniq[pos], dst):
        raise DGLError("to_block: an edge ends in a node that is not a destination node")
    dst_local = rank[pos] if dst.numel() else dst
    if dst_local.numel() and int(dst_local.max()) >= n_dst:
        raise DGLError("to_block: an edge ends in a node that is not a destination node")
    return src_nid, src_local, dst_local


def _compact_block_device(src, dst, dst_nodes, dst_local):
    """compact_block on device tensors: one hash-table pass (csrc/block.cu), one host sync for the
    number of source nodes and the status bits."""
    dev = src.device
    src = src.to(torch.int64).contiguous()
    dst_nodes = dst_nodes.contiguous()
    n_dst, n_e = int(dst_nodes.numel()), int(src.numel())
    look_up = dst_local is None
    if look_up:
        dst = torch.as_tensor(dst, dtype=torch.int64, device=dev).contiguous()
        if dst.numel() != n_e:
            raise DGLError("to_block: src and dst differ in length")
    buf = torch.empty(max(n_dst + 3 * n_e, 1) + 2, dtype=torch.int64, device=dev)
    counts = buf[:2]
    src_nid = buf[2:2 + n_dst + n_e]
    src_local = buf[2 + n_dst + n_e:2 + n_dst + 2 * n_e]
    out_dst = buf[2 + n_dst + 2 * n_e:2 + n_dst + 3 * n_e]
    ptr = lambda t: ctypes.c_void_p(t.data_ptr() if t.numel() else 0)  # noqa: E731
    with torch.cuda.device(dev):
        check_call(LIB.DGLB200CompactBlock(
            ptr(dst_nodes), n_dst, ptr(src), ptr(dst) if look_up else ctypes.c_void_p(0), n_e,
            ptr(src_nid), ptr(src_local), ptr(out_dst) if look_up else ctypes.c_void_p(0),
            ctypes.c_void_p(counts.data_ptr()),
            ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    n_src, status = counts.tolist()                     # the one host sync
    if status & 4:
        raise DGLError("to_block: negative node id")
    if status & 1:
        raise DGLError("to_block: destination nodes must be unique")
    if status & 2:
        raise DGLError("to_block: an edge ends in a node that is not a destination node")
    return src_nid[:n_src], src_local, (out_dst if look_up else dst_local)


def to_block(frontier, dst_nodes, device=None):
    """Bipartite block of a frontier (dgl.to_block): destination nodes = `dst_nodes` (unique), source
    nodes = the destinations first, then every other source in order of first appearance; edges
    keep the frontier's order. Parent ids are stored as block.srcdata[NID], dstdata[NID],
    edata[EID]."""
    dev = frontier.src.device if device is None else torch.device(device)
    dst_nodes = torch.as_tensor(dst_nodes, dtype=torch.int64, device=dev)
    known = None
    if frontier.dst_local is not None and frontier.seeds is not None and (
            frontier.seeds is dst_nodes or (frontier.seeds.shape == dst_nodes.shape
                                            and frontier.seeds.data_ptr() == dst_nodes.data_ptr())):
        known = frontier.dst_local      # same seeds: the sampler already knows every edge's slot
    src_nid, src_local, dst_local = compact_block(frontier.src, frontier.dst, dst_nodes, known)
    block = Graph(src_local, dst_local, num_src=int(src_nid.numel()),
                  num_dst=int(dst_nodes.numel()), device=dev)
    block.srcdata[NID] = src_nid
    block.dstdata[NID] = dst_nodes
    block.edata[EID] = frontier.eid
    return block
