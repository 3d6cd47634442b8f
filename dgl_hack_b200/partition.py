"""Destination-node partitioning with 1-hop halo exchange (multi-GPU message passing).

Host-side logic of SURVEY 8(e), mirroring the reference's partitioning semantics
(`dgl.transform.partition_graph_with_halo`, python/dgl/transform.py:551-640 ->
GraphOp::GetSubgraphWithHalo, src/graph/graph_op.cc:403-509): rank r owns a set of destination
nodes ("inner" nodes, `inner_node` mask) together with all their in-edges; the sources of those
edges that live on other ranks are its 1-hop halo. Local ids: inner nodes first (ascending
global id, as the reference's subgraphs preserve parent order), then halo nodes grouped by owner.

The reference has no GPU collective at all (its multi-GPU story is DDP over sampled
mini-batches); here each layer does ONE exchange of halo source rows over NCCL (grouped
send/recv = all-to-all-v over NVLink/NVSwitch), overlapped with the aggregation over the edges
whose sources are local:

    forward :  post exchange(X_inner -> X_halo) | SpMM(inner-source edges)  ; wait ; SpMM(halo edges)
    backward:  transposed SpMM(halo edges) ; post reverse exchange | transposed SpMM(inner edges) ;
               wait ; add the received halo gradients into their owners' rows (fixed peer order)

Everything here is torch + torch.distributed and device agnostic, so the exchange logic is tested
on CPU with gloo (world_size 2); the aggregation itself is the CUDA kernel library.
"""
import os

import torch
import torch.distributed as dist

SLOT_SHIFT = 27                 # csrc/spmm_kernels.cuh kSlotShift: code = (slot << 27) | row
MAX_SLOTS = 16


def range_partition(num_nodes, dst, k):
    """Contiguous node ranges balanced by in-edge count (the default when the node numbering has
    community structure; METIS assignments from the reference's `metis_partition` can be passed to
    `Partition.build` instead -- any node -> part vector works)."""
    deg = torch.bincount(dst, minlength=num_nodes).to(torch.float64)
    cum = torch.cumsum(deg + 1e-3, 0)  # +eps: isolated nodes still spread
    bounds = cum[-1] * torch.arange(1, k, dtype=torch.float64, device=dst.device) / k
    cuts = torch.searchsorted(cum, bounds)
    part = torch.bucketize(torch.arange(num_nodes, device=dst.device), cuts, right=True)
    return part


def symmetric_csr(num_nodes, src, dst):
    """Host CSR (xadj, adjncy int64) of the symmetrised simple graph -- what the reference feeds
    METIS: `to_bidirected(g)` then the in-CSR (python/dgl/transform.py:625-626,
    src/graph/metis_partition.cc:25-31). Self loops and parallel edges are dropped."""
    src = torch.as_tensor(src, dtype=torch.int64)
    dst = torch.as_tensor(dst, dtype=torch.int64).to(src.device)
    if not src.is_cuda:  # host arrays: counting sort + per-row sort / unique in the library
        import ctypes
        from ._lib import LIB, check_call
        src, dst = src.contiguous(), dst.contiguous()
        xadj = torch.empty(num_nodes + 1, dtype=torch.int64)
        adjncy = torch.empty(max(2 * src.numel(), 1), dtype=torch.int64)
        nnz = ctypes.c_int64()
        check_call(LIB.DGLB200SymmetrizeCOO(
            int(num_nodes), int(src.numel()), ctypes.c_void_p(src.data_ptr()),
            ctypes.c_void_p(dst.data_ptr()), ctypes.c_void_p(xadj.data_ptr()),
            ctypes.c_void_p(adjncy.data_ptr()), ctypes.byref(nnz)))
        return xadj, adjncy[:nnz.value].clone()
    keep = src != dst
    u = torch.cat([src[keep], dst[keep]])
    v = torch.cat([dst[keep], src[keep]])
    key = torch.unique(u * num_nodes + v)  # sorted: rows ascending, neighbours ascending
    rows = torch.div(key, num_nodes, rounding_mode="floor")
    adjncy = (key - rows * num_nodes).cpu().contiguous()
    xadj = torch.zeros(num_nodes + 1, dtype=torch.int64)
    xadj[1:] = torch.cumsum(torch.bincount(rows, minlength=num_nodes), 0).cpu()
    return xadj, adjncy


def metis_partition_assignment(num_nodes, src, dst, k, balance="nodes", seed=0,
                               return_cut=False):
    """Node -> part vector (int64, CPU) from the library's multilevel k-way partitioner
    (DGLB200MetisPartition, csrc/kway.cu), the counterpart of
    `_CAPI_DGLMetisPartition(to_bidirected(g), k)` (transform.py:623-629). balance="nodes"
    equalises node counts as the reference does (vwgt = NULL); "in_edges" equalises
    in-degree + 1, i.e. the aggregation work each GPU gets."""
    import ctypes
    from ._lib import LIB, check_call
    xadj, adjncy = symmetric_csr(num_nodes, src, dst)
    vw = None
    if balance == "in_edges":
        d = torch.as_tensor(dst, dtype=torch.int64)
        vw = (torch.bincount(d, minlength=num_nodes) + 1).cpu().contiguous()
    elif balance != "nodes":
        raise ValueError("balance must be 'nodes' or 'in_edges'")
    part = torch.empty(num_nodes, dtype=torch.int64)
    cut = ctypes.c_int64()
    check_call(LIB.DGLB200MetisPartition(
        int(num_nodes), ctypes.c_void_p(xadj.data_ptr()), ctypes.c_void_p(adjncy.data_ptr()),
        ctypes.c_void_p(vw.data_ptr() if vw is not None else None), int(k), int(seed),
        ctypes.c_void_p(part.data_ptr()), ctypes.byref(cut)))
    return (part, int(cut.value)) if return_cut else part


def kway_partition_on_rank0(num_nodes, src, dst, world, rank, device, balance="in_edges"):
    """Rank 0 runs the k-way partitioner (host code) and broadcasts the node -> part vector, the
    way tools/partition.py partitions once and every trainer loads its part. Returns
    (part on `device`, seconds rank 0 spent partitioning)."""
    import time
    part = torch.empty(num_nodes, dtype=torch.int64, device=device)
    secs = torch.zeros(1, dtype=torch.float64, device=device)
    if rank == 0:
        t0 = time.time()
        part.copy_(metis_partition_assignment(num_nodes, src, dst, world, balance=balance))
        secs[0] = time.time() - t0
    if world > 1:
        dist.broadcast(part, 0)
        dist.broadcast(secs, 0)
    return part, float(secs.item())


def partition_graph_with_halo(num_nodes, src, dst, node_part, num_parts=None):
    """1-hop form of `dgl.transform.partition_graph_with_halo(g, node_part, 1)`
    (transform.py:551-586 -> GraphOp::GetSubgraphWithHalo, src/graph/graph_op.cc:403-509) on COO
    arrays. For every part: `induced_nodes` (parent ids, ascending), `induced_edges` (parent ids of
    the in-edges of the owned nodes, ordered by (destination, edge id) as `InEdges(nodes)` returns
    them), `inner_node` (1 = owned, 0 = halo), `inner_edge` (1 = the source is owned too, 0 = cut
    edge, graph_op.cc:427-433) and the local COO in `induced_edges` order."""
    src = torch.as_tensor(src, dtype=torch.int64)
    dst = torch.as_tensor(dst, dtype=torch.int64)
    node_part = torch.as_tensor(node_part, dtype=torch.int64)
    assert node_part.numel() == num_nodes
    k = int(num_parts if num_parts is not None else node_part.max().item() + 1)
    out = {}
    for p in range(k):
        eids = torch.nonzero(node_part[dst] == p).flatten()
        eids = eids[torch.sort(dst[eids], stable=True).indices]
        es, ed = src[eids], dst[eids]
        owned = node_part == p
        mark = owned.clone()
        mark[es] = True
        nodes = torch.nonzero(mark).flatten()
        g2l = torch.full((num_nodes,), -1, dtype=torch.int64)
        g2l[nodes] = torch.arange(nodes.numel())
        out[p] = {"induced_nodes": nodes, "induced_edges": eids,
                  "inner_node": owned[nodes].to(torch.int32),
                  "inner_edge": owned[es].to(torch.int32),
                  "src": g2l[es], "dst": g2l[ed], "part_id": node_part[nodes]}
    return out


class Partition(object):
    """One rank's share of a graph: inner nodes, halo nodes, local COO split by source locality."""

    def __init__(self):
        self.rank = self.world = 0
        self.inner_global = self.halo_global = None
        self.halo_counts = self.send_idx = None
        self.edges_inner = self.edges_halo = None  # (src_local, dst_local) pairs
        self.n_inner = self.n_halo = 0
        self.cut_edges = self.local_edges = 0
        # direct peer-gather plan (None when the plan was loaded from per-part files)
        self.fwd_coo = self.bwd_coo = self.fwd_code = self.bwd_code = None
        self.n_ohalo = 0
        self.rows_max = 0
        self.push = None       # [(rows, codes)] forward / backward push lists (push mode)
        self.ext_max = 0

    @classmethod
    def build(cls, num_nodes, src, dst, part, rank, world):
        """src/dst: global COO (int64, any device), part: owner of every node. Deterministic and
        communication free: every rank derives its send lists from the same global COO."""
        p = cls()
        p.rank, p.world = rank, world
        dev = src.device
        part = part.to(dev)
        own_dst = part[dst] == rank
        e_src, e_dst = src[own_dst], dst[own_dst]
        p.inner_global = torch.nonzero(part == rank).flatten()
        p.n_inner = int(p.inner_global.numel())
        g2l = torch.full((num_nodes,), -1, dtype=torch.int64, device=dev)
        g2l[p.inner_global] = torch.arange(p.n_inner, device=dev)
        remote = part[e_src] != rank
        halo = torch.unique(e_src[remote])
        owner = part[halo]
        order = torch.argsort(owner * num_nodes + halo)
        p.halo_global = halo[order]
        p.n_halo = int(halo.numel())
        p.halo_counts = torch.bincount(part[p.halo_global], minlength=world).tolist()
        g2l[p.halo_global] = p.n_inner + torch.arange(p.n_halo, device=dev)
        ls, ld = g2l[e_src], g2l[e_dst]
        # edges keep their global order inside each class (edge id order = reduction order)
        p.edges_inner = (ls[~remote], ld[~remote])
        p.edges_halo = (ls[remote] - p.n_inner, ld[remote])
        p.local_edges = int(e_src.numel())
        p.cut_edges = int(remote.sum())
        # what each peer needs from me: sources I own of edges whose destination the peer owns
        mine_src = part[src] == rank
        p.send_idx = []
        for q in range(world):
            if q == rank:
                p.send_idx.append(torch.empty(0, dtype=torch.int64, device=dev))
                continue
            need = torch.unique(src[mine_src & (part[dst] == q)])  # ascending = peer's halo order
            p.send_idx.append(g2l[need])
        p.send_counts = [int(t.numel()) for t in p.send_idx]
        p.send_cat = torch.cat(p.send_idx) if world > 1 else p.send_idx[0]
        p._build_direct_plan(num_nodes, src, dst, part, e_src, e_dst, ls, ld, mine_src)
        return p

    def _build_direct_plan(self, num_nodes, src, dst, part, e_src, e_dst, ls, ld, mine_src):
        """Plan of the single-pass aggregation with peer loads (SURVEY 8e option iii): every node
        is addressed as (owner rank, row inside the owner's block) = code (owner << 27) | row,
        row = rank of the node among its owner's nodes in ascending global id (the order of
        `inner_global` on the owner). Forward graph: ALL in-edges of the inner nodes, sources
        numbered inner-first then halo (`fwd_code` translates them); backward graph: ALL out-edges
        of the inner nodes, destinations numbered inner-first then the remote destinations
        (`bwd_code`), so the gradient is one transposed walk that reads grad_out rows from the
        peers -- no pushed halo gradients and no reduction pass at the owner."""
        rank, world, dev = self.rank, self.world, src.device
        counts = torch.bincount(part, minlength=world)
        self.rows_max = int(counts.max()) if num_nodes else 0
        if world > MAX_SLOTS or self.rows_max >= (1 << SLOT_SHIFT):
            return
        order = torch.argsort(part, stable=True)
        starts = torch.cumsum(counts, 0) - counts
        code = torch.empty(num_nodes, dtype=torch.int64, device=dev)
        code[order] = (torch.arange(num_nodes, device=dev) - starts[part[order]]) + \
            (part[order] << SLOT_SHIFT)
        del order
        self.fwd_coo = (ls, ld)                      # global edge-id order
        self.fwd_code = torch.cat([code[self.inner_global], code[self.halo_global]]).to(torch.int32)
        o_src, o_dst = src[mine_src], dst[mine_src]
        away = part[o_dst] != rank
        ohalo = torch.unique(o_dst[away])
        ohalo = ohalo[torch.argsort(part[ohalo] * num_nodes + ohalo)]   # grouped by owner, like halo
        self.n_ohalo = int(ohalo.numel())
        g2c = torch.full((num_nodes,), -1, dtype=torch.int64, device=dev)
        g2c[self.inner_global] = torch.arange(self.n_inner, device=dev)
        lsrc = g2c[o_src]
        g2c[ohalo] = self.n_inner + torch.arange(self.n_ohalo, device=dev)
        self.bwd_coo = (lsrc, g2c[o_dst])
        self.bwd_code = torch.cat([code[self.inner_global], code[ohalo]]).to(torch.int32)
        self._build_push_plan(num_nodes, src, dst, part, e_src, e_dst)

    def _build_push_plan(self, num_nodes, src, dst, part, e_src, e_dst):
        """Push lists of the "push" reduction mode (PeerGather): which of this rank's rows every
        peer needs and where they land in the peer's receive buffer. Forward: the peer's halo
        (sources grouped by owner, ascending id: this rank's block starts after the blocks of the
        owners below it); backward: the peer's remote destinations, same layout. Every rank
        derives all offsets from the global COO -- no communication. `push[which]` =
        (rows int32, codes int32): row rows[i] goes to row (codes[i] & mask) of rank
        (codes[i] >> 27)'s buffer; the list is interleaved over the destination ranks starting at
        rank + 1, so no GPU's ingress takes all senders at once."""
        rank, world, dev = self.rank, self.world, src.device
        cross = part[src] != part[dst]
        cs, cd = src[cross], dst[cross]
        g2l = torch.full((num_nodes,), -1, dtype=torch.int64, device=dev)
        g2l[self.inner_global] = torch.arange(self.n_inner, device=dev)
        self.push = []
        # forward: (node = source, consumer = owner of the destination);
        # backward: (node = destination, consumer = owner of the source)
        for node, consumer in ((cs, part[cd]), (cd, part[cs])):
            uk = torch.unique(node * world + consumer)
            un, uc = uk // world, uk % world            # ascending node id inside a consumer
            owner = part[un]
            need = torch.bincount(uc * world + owner, minlength=world * world).view(world, world)
            mine = owner == rank
            un, uc = un[mine], uc[mine]
            order = torch.argsort(uc * num_nodes + un)  # grouped by consumer, ascending id
            un, uc = un[order], uc[order]
            cnt = torch.bincount(uc, minlength=world)
            first = torch.cumsum(cnt, 0) - cnt
            pos = torch.arange(un.numel(), device=dev) - first[uc]
            base = need[:, :rank].sum(1)                # my block's first row in each consumer
            codes = (uc << SLOT_SHIFT) + base[uc] + pos
            rr = torch.argsort(pos * world + (uc - rank - 1) % world)
            self.push.append((g2l[un][rr].to(torch.int32).contiguous(),
                              codes[rr].to(torch.int32).contiguous()))
            assert int(need[rank].sum()) == (self.n_halo if len(self.push) == 1 else self.n_ohalo)
            # the largest receive buffer any rank needs (rows): the same number on every rank
            self.ext_max = max(self.ext_max, int(need.sum(1).max()))

    @classmethod
    def from_files(cls, directory, rank, world, device="cpu"):
        """The same plan as `build`, from the per-part files partition_tool.py writes (the way the
        reference's trainers load `<part_id>.dgl`): this rank's file gives its nodes and edges, the
        other parts' `induced_nodes` / `inner_node` / `part_id` give what they need from here."""
        import numpy as np
        files = [np.load(os.path.join(directory, "%d.npz" % q)) for q in range(world)]
        me = files[rank]
        assert int(me["num_parts"]) == world, "files were written for another number of parts"
        p = cls()
        p.rank, p.world = rank, world
        dev = torch.device(device)
        t = lambda a: torch.from_numpy(np.asarray(a)).to(dev)  # noqa: E731
        nodes, inner, owner = t(me["induced_nodes"]), t(me["inner_node"]).bool(), t(me["part_id"])
        p.inner_global = nodes[inner]
        p.n_inner = int(p.inner_global.numel())
        halo, halo_owner = nodes[~inner], owner[~inner]
        order = torch.argsort(halo_owner * int(me["num_nodes"]) + halo)
        p.halo_global = halo[order]
        p.n_halo = int(halo.numel())
        p.halo_counts = torch.bincount(halo_owner, minlength=world).tolist()
        # local ids of the file (parent order) -> plan ids (inner first, halo grouped by owner)
        remap = torch.empty(nodes.numel(), dtype=torch.int64, device=dev)
        remap[torch.nonzero(inner).flatten()] = torch.arange(p.n_inner, device=dev)
        remap[torch.nonzero(~inner).flatten()[order]] = p.n_inner + torch.arange(p.n_halo, device=dev)
        # edges back in parent edge-id order (= the order `build` keeps)
        eorder = torch.argsort(t(me["induced_edges"]))
        ls, ld = remap[t(me["src"])[eorder]], remap[t(me["dst"])[eorder]]
        remote = ~t(me["inner_edge"]).bool()[eorder]
        p.edges_inner = (ls[~remote], ld[~remote])
        p.edges_halo = (ls[remote] - p.n_inner, ld[remote])
        p.local_edges = int(ls.numel())
        p.cut_edges = int(remote.sum())
        g2l = {int(g): i for i, g in enumerate(p.inner_global.tolist())} if world > 1 else {}
        p.send_idx = []
        for q in range(world):
            if q == rank:
                p.send_idx.append(torch.empty(0, dtype=torch.int64, device=dev))
                continue
            fq = files[q]
            need = np.sort(fq["induced_nodes"][(fq["inner_node"] == 0) & (fq["part_id"] == rank)])
            p.send_idx.append(torch.tensor([g2l[int(g)] for g in need], dtype=torch.int64,
                                           device=dev))
        p.send_counts = [int(x.numel()) for x in p.send_idx]
        p.send_cat = torch.cat(p.send_idx) if world > 1 else p.send_idx[0]
        p.node_feats = {k[5:]: t(me[k])[torch.nonzero(inner).flatten()] for k in me.files
                        if k.startswith("feat/")}
        return p

    def _a2a(self):
        """One all-to-all-v instead of 2(k-1) grouped sends/recvs where the backend has it (NCCL)."""
        return self.world > 1 and dist.is_initialized() and dist.get_backend() == "nccl"

    # ---- exchanges (grouped isend/irecv: works on gloo and NCCL)
    def start_forward_exchange(self, x_inner):
        """Posts the halo exchange; returns (requests, x_halo, keepalive)."""
        d = x_inner.shape[1:]
        peer = getattr(self, "peer", None)
        if peer is not None and peer.usable(x_inner):
            work, x_halo = peer.forward(x_inner)
            return [work], x_halo, [x_inner]
        x_halo = x_inner.new_empty((self.n_halo,) + tuple(d))
        if self._a2a():
            buf = x_inner.index_select(0, self.send_cat)  # one pack kernel for all peers
            work = dist.all_to_all_single(x_halo, buf, output_split_sizes=self.halo_counts,
                                          input_split_sizes=self.send_counts, async_op=True)
            return [work], x_halo, [buf]
        ops, keep, off = [], [], 0
        for q in range(self.world):
            cnt = self.halo_counts[q]
            if q != self.rank and cnt > 0:
                ops.append(dist.P2POp(dist.irecv, x_halo[off:off + cnt], q))
            off += cnt
        for q in range(self.world):
            if q != self.rank and self.send_idx[q].numel() > 0:
                buf = x_inner.index_select(0, self.send_idx[q])
                keep.append(buf)
                ops.append(dist.P2POp(dist.isend, buf, q))
        reqs = dist.batch_isend_irecv(ops) if ops else []
        return reqs, x_halo, keep

    def start_backward_exchange(self, g_halo):
        """Sends halo gradients back to their owners; returns (requests, per-peer recv buffers)."""
        d = g_halo.shape[1:]
        peer = getattr(self, "peer", None)
        if peer is not None and peer.usable(g_halo):
            work, buf = peer.backward(g_halo)
            return [work], buf
        if self._a2a():
            buf = g_halo.new_empty((int(self.send_cat.numel()),) + tuple(d))
            work = dist.all_to_all_single(buf, g_halo.contiguous(),
                                          output_split_sizes=self.send_counts,
                                          input_split_sizes=self.halo_counts, async_op=True)
            return [work], buf
        ops, recv, off = [], [], 0
        for q in range(self.world):
            n = self.send_idx[q].numel()
            buf = g_halo.new_empty((n,) + tuple(d)) if (q != self.rank and n > 0) else None
            recv.append(buf)
            if buf is not None:
                ops.append(dist.P2POp(dist.irecv, buf, q))
        for q in range(self.world):
            cnt = self.halo_counts[q]
            if q != self.rank and cnt > 0:
                ops.append(dist.P2POp(dist.isend, g_halo[off:off + cnt].contiguous(), q))
            off += cnt
        reqs = dist.batch_isend_irecv(ops) if ops else []
        return reqs, recv

    def reduce_received(self, g_inner, recv):
        """Adds the halo gradients received from all peers (rows in send_cat order) into their
        owners' rows. Overridden by PartitionedGraph with a deterministic segment-sum kernel."""
        off = 0
        for q in range(self.world):
            n = self.send_counts[q]
            if n:
                g_inner.index_add_(0, self.send_idx[q], recv[off:off + n])
            off += n
        return g_inner

    def exchange_bytes(self, row_bytes):
        sent = sum(int(s.numel()) for s in self.send_idx) * row_bytes
        return {"send": sent, "recv": self.n_halo * row_bytes}


class _HaloCopySum(torch.autograd.Function):
    """out[v] = sum over in-edges of X[src] for the rank's inner destinations, sources local or
    halo. `agg(which, x)` / `agg_t(which, grad_out)` are the aggregation and its transpose for
    the inner-source ("inner") and halo-source ("halo") edge sets."""

    @staticmethod
    def forward(ctx, part, agg, agg_t, x_inner):
        reqs, x_halo, keep = part.start_forward_exchange(x_inner)
        out = agg("inner", x_inner)            # overlaps the exchange
        for r in reqs:
            r.wait()
        if part.n_halo > 0:
            out = out + agg("halo", x_halo)
        ctx.part, ctx.agg_t = part, agg_t
        del keep
        return out

    @staticmethod
    def backward(ctx, grad_out):
        part, agg_t = ctx.part, ctx.agg_t
        grad_out = grad_out.contiguous()
        g_halo = agg_t("halo", grad_out) if part.n_halo > 0 else \
            grad_out.new_empty((0,) + tuple(grad_out.shape[1:]))
        reqs, recv = part.start_backward_exchange(g_halo)
        g_inner = agg_t("inner", grad_out)     # overlaps the exchange
        for r in reqs:
            r.wait()
        if torch.is_tensor(recv):              # all-to-all path: one buffer in send_cat order
            g_inner = part.reduce_received(g_inner, recv)
        else:
            for q in range(part.world):        # fixed peer order; ids unique within a peer
                if recv[q] is not None:
                    g_inner.index_add_(0, part.send_idx[q], recv[q])
        return None, None, None, g_inner


class PartitionedGraph(object):
    """A Partition bound to the CUDA kernels: two GraphIndex objects (inner-source and
    halo-source edges, destinations = inner nodes) and the halo-exchanging aggregation."""

    def __init__(self, part, device, halo="auto", max_feat=256, blocks=None):
        """halo: "direct" = ONE aggregation kernel per direction whose gathers read the peers'
        staged feature blocks over NVLink (PeerGather; needs all ranks on one node), "peer" =
        peer-memory pushes into halo buffers + two-pass aggregation (PeerExchange), "nccl" =
        all-to-all-v + two-pass aggregation, "auto" = direct when the environment allows it
        (B200_HALO overrides). `blocks`: an already-built block provider (LocalBlocks for
        one-process tests / all-local timing); default = CUDA-IPC blocks agreed on by all ranks."""
        from .graph_index import GraphIndex
        self.part = part
        self.device = torch.device(device)
        halo = os.environ.get("B200_HALO", halo)
        part.peer = None
        self.gather = None
        self.g_inner = self.g_halo = self.g_rev = None
        nccl = part.world > 1 and dist.is_initialized() and dist.get_backend() == "nccl"
        if part.world == 1:
            # one rank: the whole graph, no staging, no codes
            si, di = part.edges_inner
            self.g_inner = GraphIndex.from_coo(part.n_inner, part.n_inner, si, di, device)
            self.halo_path = "none"
            return
        if halo in ("auto", "direct") and part.fwd_coo is not None and \
                (blocks is not None or nccl):
            if blocks is None:
                blocks = IpcBlocks.create(part, self.device, max_feat, required=(halo == "direct"))
            if blocks is not None:
                self.gather = PeerGather(part, self.device, blocks)
                self.halo_path = "direct"
                return
        if halo == "direct":
            raise RuntimeError("halo='direct' needs NCCL ranks on one node (or a block provider)")
        if nccl and halo in ("auto", "peer"):
            ok = torch.ones(1, device=self.device)
            try:
                px = PeerExchange(part, self.device, max_feat)
            except Exception:
                if halo == "peer":
                    raise
                px, ok = None, torch.zeros(1, device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            part.peer = px if ok.item() > 0 else None
        self.halo_path = "peer" if part.peer is not None else "nccl"
        si, di = part.edges_inner
        sh, dh = part.edges_halo
        self.g_inner = GraphIndex.from_coo(part.n_inner, part.n_inner, si, di, device)
        self.g_halo = GraphIndex.from_coo(max(part.n_halo, 1), part.n_inner, sh, dh, device)
        # received halo gradients -> owners: row i of the receive buffer belongs to inner node
        # send_cat[i]; as a graph (edge i -> node send_cat[i]) the reduction is copy_e + sum, one
        # deterministic kernel instead of k-1 index_add_ launches with atomics
        if part.send_cat.numel() > 0:
            sc = part.send_cat.to(device)
            self.g_rev = GraphIndex.from_coo(part.n_inner, part.n_inner, sc, sc, device)
            part.reduce_received = self._reduce_received

    def _agg(self, which, x):
        from . import kernel as K
        g = self.g_inner if which == "inner" else self.g_halo
        out = x.new_empty((self.part.n_inner,) + tuple(x.shape[1:]))
        K.copy_reduce("sum", g, 0, x.contiguous(), out)
        return out

    def _agg_t(self, which, grad_out):
        from . import kernel as K
        g = self.g_inner if which == "inner" else self.g_halo
        n_src = self.part.n_inner if which == "inner" else self.part.n_halo
        rows = max(n_src, 1) if which == "halo" else n_src
        grad = grad_out.new_empty((rows,) + tuple(grad_out.shape[1:]))
        K.backward_copy_reduce("sum", g, 0, grad, grad_out, grad_out, grad)
        return grad[:n_src]

    def _reduce_received(self, g_inner, recv):
        from . import kernel as K
        add = g_inner.new_empty(g_inner.shape)
        K.copy_reduce("sum", self.g_rev, 2, recv, add)
        return g_inner.add_(add)

    def copy_src_sum(self, x_inner, src_scale=None, dst_scale=None):
        """Halo-exchanging `update_all(copy_src, sum)` for this rank's inner nodes. With scales
        (float vectors over the inner nodes, constants of the graph such as GraphConv's degree
        norms): out[v] = dst_scale[v] * sum_{u->v} src_scale[u] * x[u], the scaling done inside the
        aggregation kernels (backend.scaled_copy_sum); the halo nodes' src_scale values are
        fetched from their owners once and cached."""
        if self.part.world == 1:
            from . import backend as F
            return F.scaled_copy_sum(self.g_inner, x_inner, src_scale, dst_scale)
        if self.gather is not None:
            cs = src_scale.detach().reshape(-1).contiguous() if src_scale is not None else None
            rs = dst_scale.detach().reshape(-1).contiguous() if dst_scale is not None else None
            return _PeerCopySum.apply(self.gather, x_inner.contiguous(), cs, rs)
        if src_scale is None and dst_scale is None:
            return _HaloCopySum.apply(self.part, self._agg, self._agg_t, x_inner)
        from . import kernel as K
        part = self.part
        cs_in = src_scale.detach().reshape(-1).contiguous() if src_scale is not None else None
        rs = dst_scale.detach().reshape(-1).contiguous() if dst_scale is not None else None
        cs_halo = None
        if cs_in is not None:
            # keyed on the caller's tensor, which the cache keeps alive (an address alone could be
            # recycled for another tensor with other contents)
            cache = self.__dict__.setdefault("_halo_scale", {})
            key = (id(src_scale), src_scale._version)
            if cache.get("owner") is not src_scale or key not in cache:
                cache.clear()
                cache["owner"] = src_scale
                h = cs_in.new_ones((max(part.n_halo, 1),))
                if part.world > 1:
                    peer, part.peer = part.peer, None      # width 1: not a peer-push shape
                    reqs, got, keep = part.start_forward_exchange(cs_in.reshape(-1, 1))
                    for r in reqs:
                        r.wait()
                    part.peer = peer
                    h[:part.n_halo] = got.reshape(-1)
                cache[key] = h
            cs_halo = cache[key]

        def agg(which, x):
            g, cs = (self.g_inner, cs_in) if which == "inner" else (self.g_halo, cs_halo)
            out = x.new_empty((part.n_inner,) + tuple(x.shape[1:]))
            K.scaled_copy_sum(g, False, x.contiguous(), cs, rs, out)
            return out

        def agg_t(which, grad_out):
            g, cs = (self.g_inner, cs_in) if which == "inner" else (self.g_halo, cs_halo)
            n_src = part.n_inner if which == "inner" else part.n_halo
            rows = max(n_src, 1) if which == "halo" else n_src
            grad = grad_out.new_empty((rows,) + tuple(grad_out.shape[1:]))
            K.scaled_copy_sum(g, True, grad_out, rs, cs, grad)
            return grad[:n_src]
        return _HaloCopySum.apply(part, agg, agg_t, x_inner)


class _PeerCopySum(torch.autograd.Function):
    """out[v] = rs[v] * sum_{u -> v} cs[u] * x[u] over ALL in-edges of the rank's inner nodes in
    one kernel: every rank stages cs * x in its block, one barrier, then the SpMM gathers each
    source row from the block of its owner (own HBM or a peer's over NVLink). The gradient is the
    same thing on the out-edges with grad_out staged (scaled by rs) and cs applied per row."""

    @staticmethod
    def forward(ctx, gather, x, cs, rs):
        ctx.gather, ctx.scales = gather, (cs, rs)
        return gather.run(0, x, cs, rs)

    @staticmethod
    def backward(ctx, grad_out):
        cs, rs = ctx.scales
        return None, ctx.gather.run(1, grad_out.contiguous(), rs, cs), None, None


class PeerGather(object):
    """Partitioned aggregation over staged feature blocks. `blocks` provides, per
    rank, two forward and two backward staging blocks of rows_max x max_feat floats that every
    rank of the node can read, plus a barrier. Blocks alternate, so a block is rewritten only
    after every rank has passed the barrier that follows its last readers (two barriers lie
    between two writes of the same block). Four ways to reduce (B200_HALO_MODE):

    "push" (default): every rank STORES the rows its peers need straight into their receive
        buffers (one streaming kernel on a high-priority side stream, fused with the degree
        scaling; peer stores are posted, so the link is not bounded by outstanding remote reads
        the way a pull is) while the SpMM over the inner-source edges runs on x itself -- no
        staging copy at all; after the barrier a second SpMM adds the halo-source edges from the
        receive buffer to the same rows (`addend`). Backward is the same thing on the out-edges:
        no reduction of received gradients.
    "split": the de-duplicated halo rows are pulled into a local buffer by ONE streaming
        kernel on a side stream (all of its loads are remote; rows listed round-robin over their
        owners) WHILE the SpMM over the inner-source edges runs; a second SpMM then adds the
        halo-source edges to the same rows (`addend`, no add kernel, no third pass).
        Measured at 8 GPUs the pull is the largest item of the GCN epoch, so it is the thing to
        hide; remote bytes are the de-duplicated ones (0.7x of the per-edge reads).
    "pull": pull first, then ONE SpMM over all edges gathering from two local slots.
    "direct": ONE SpMM whose gathers read the peers' blocks in place (SURVEY 8e option iii;
        per-edge remote reads, no halo buffer)."""

    def __init__(self, part, device, blocks, mode=None):
        from .graph_index import GraphIndex
        self.part, self.device, self.blocks = part, torch.device(device), blocks
        if mode is None:
            mode = os.environ.get("B200_HALO_MODE", "push")
        if mode not in ("push", "split", "pull", "direct"):
            raise ValueError("B200_HALO_MODE must be push, split, pull or direct (got %r)" % (mode,))
        if mode == "push" and (part.push is None or not getattr(blocks, "has_recv", False)):
            mode = "split"           # no receive buffers in this block provider
        self.mode = mode
        n_in = part.n_inner
        n_ext = (part.n_halo, part.n_ohalo)
        dev = self.device
        self.code = (part.fwd_code.to(dev).contiguous(), part.bwd_code.to(dev).contiguous())
        self._turn = [0, 0]
        self._halo_buf = {}
        if mode in ("pull", "direct"):
            fs, fd = part.fwd_coo
            bs, bd = part.bwd_coo
            self.g_f = GraphIndex.from_coo(n_in + part.n_halo, n_in, fs, fd, dev)
            self.g_b = GraphIndex.from_coo(n_in, n_in + part.n_ohalo, bs, bd, dev)
            self.local_code = tuple(torch.cat([
                torch.arange(n_in, dtype=torch.int32, device=dev),
                (1 << SLOT_SHIFT) + torch.arange(n, dtype=torch.int32, device=dev)]).contiguous()
                for n in n_ext)
        else:
            # inner -> inner edges serve both directions (in-CSR forward, out-CSR backward)
            si, di = part.edges_inner
            self.g_ii = GraphIndex.from_coo(n_in, n_in, si, di, dev)
            sh, dh = part.edges_halo
            self.g_fh = GraphIndex.from_coo(max(part.n_halo, 1), n_in, sh, dh, dev)
            bs, bd = part.bwd_coo
            away = bd >= n_in
            self.g_bh = GraphIndex.from_coo(n_in, max(part.n_ohalo, 1), bs[away], bd[away] - n_in,
                                            dev)
            # high priority: the pull's few CTAs get onto the SMs as soon as CTAs of the SpMM
            # that was launched first retire, instead of queueing behind its whole grid
            self._side = torch.cuda.Stream(device=dev, priority=-1)
            # CTAs per SM of the transfer kernel beside the SpMM (8 GPUs, GCN epoch: push 9.3 /
            # 9.5 / 10.0 ms at 1 / 2 / 3; the pull needs more loads in flight: 13.5 / 11.4 / 11.6
            # at 1 / 2 / 4)
            ctas = os.environ.get("B200_PULL_CTAS_PER_SM", "1" if mode == "push" else "2")
            self._pull_blocks = int(ctas) * \
                torch.cuda.get_device_properties(dev).multi_processor_count
            if mode == "push":
                self._push_lists = [(r.to(dev), c.to(dev)) for r, c in part.push]
                self._pending = [None, None]
        # pull lists: the remote rows of each direction, round-robin over their owners starting
        # at rank + 1 (every GPU's egress serves all its readers at once instead of one GPU
        # serving everybody in turn); dst_row = position in the halo buffer
        self.pull_codes, self.pull_rows = [], []
        for c, n in zip(self.code, n_ext):
            ext = c[n_in:].to(torch.int64)
            owner = ext >> SLOT_SHIFT
            cnt = torch.bincount(owner, minlength=part.world)
            first = torch.cumsum(cnt, 0) - cnt
            pos = torch.arange(n, device=dev) - first[owner]      # index inside the owner's group
            key = pos * part.world + (owner - part.rank - 1) % part.world
            order = torch.argsort(key)
            self.pull_codes.append(ext[order].to(torch.int32).contiguous())
            self.pull_rows.append(order.to(torch.int32).contiguous())

    def stage(self, which, x, scale):
        """Writes scale * x into this rank's next block of direction `which`; returns its index.
        Push mode: stores the rows the peers need straight into their receive buffers instead
        (side stream), and keeps x / scale for the inner-source pass of reduce()."""
        b = 2 * which + self._turn[which]
        self._turn[which] ^= 1
        n, d = x.shape
        if self.mode == "push":
            import ctypes
            from ._lib import LIB, check_call
            if d > self.blocks.max_feat:
                raise ValueError("feature width %d exceeds the receive buffers (max_feat=%d)"
                                 % (d, self.blocks.max_feat))
            rows, codes = self._push_lists[which]
            cur = torch.cuda.current_stream(self.device)
            self._side.wait_stream(cur)                       # x is complete
            if int(rows.numel()):
                bases = self.blocks.rbases(b)
                arr = (ctypes.c_void_p * len(bases))(*[int(p) for p in bases])
                with torch.cuda.stream(self._side):
                    check_call(LIB.DGLB200HaloScatter(
                        len(bases), arr, ctypes.c_void_p(codes.data_ptr()),
                        ctypes.c_void_p(rows.data_ptr()),
                        ctypes.c_void_p(scale.data_ptr()) if scale is not None else None,
                        int(rows.numel()), d, ctypes.c_void_p(x.data_ptr()), self._pull_blocks,
                        self.device.index or 0, ctypes.c_void_p(self._side.cuda_stream)))
            # no record_stream: reduce() makes the current stream wait for the side stream before
            # x can be released, and record_stream would keep freed blocks out of the allocator's
            # pool until the side stream is polled (new cudaMallocs: +4 ms of host time per epoch)
            self._pending[which] = (x, scale)
            return b
        if d > self.blocks.max_feat:
            raise ValueError("feature width %d exceeds the staging blocks (max_feat=%d)"
                             % (d, self.blocks.max_feat))
        view = self.blocks.own_view(b)[: n * d].view(n, d)
        if scale is None:
            view.copy_(x)
        else:
            torch.mul(x, scale.view(-1, 1), out=view)
        return b

    def _pulled(self, which, b, d, like, max_blocks=0):
        """Local copy of the rows this rank needs from its peers (one streaming pull kernel on
        the current stream)."""
        import ctypes
        from ._lib import LIB, check_call
        n = int(self.pull_codes[which].numel())
        key = (which, d)
        buf = self._halo_buf.get(key)
        if buf is None or buf.dtype != like.dtype:
            buf = self._halo_buf[key] = like.new_empty((max(n, 1), d))
        if n:
            bases = self.blocks.bases(b)
            arr = (ctypes.c_void_p * len(bases))(*[int(p) for p in bases])
            check_call(LIB.DGLB200HaloPull(
                len(bases), arr, ctypes.c_void_p(self.pull_codes[which].data_ptr()),
                ctypes.c_void_p(self.pull_rows[which].data_ptr()), n, d,
                ctypes.c_void_p(buf.data_ptr()), int(max_blocks), self.device.index or 0,
                ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return buf

    def _own_base(self, b):
        bases = self.blocks.bases(b)
        return bases[self.part.rank if len(bases) > self.part.rank else 0]

    def reduce(self, which, b, d, row_scale, like, barrier=False):
        """The aggregation of direction `which` over the blocks staged as index b. `barrier`:
        issue the blocks' barrier first (run() does; tests and phase timings call it themselves)."""
        from . import kernel as K
        n_in = self.part.n_inner
        out = like.new_empty((n_in, d))
        if self.mode == "push":
            n_ext = self.part.n_halo if which == 0 else self.part.n_ohalo
            x, scale = self._pending[which]
            cur = torch.cuda.current_stream(self.device)
            if barrier:
                with torch.cuda.stream(self._side):           # after this rank's pushes
                    self.blocks.barrier()
            # inner-source edges (own rows, scaled per gathered row) while the halo rows land
            K.scaled_copy_sum(self.g_ii, which == 1, x, scale, None if n_ext else row_scale, out)
            cur.wait_stream(self._side)
            if n_ext:
                g = self.g_fh if which == 0 else self.g_bh
                halo = self.blocks.recv_view(b)[: n_ext * d].view(n_ext, d)
                K.scaled_copy_sum(g, which == 1, halo, None, row_scale, out, addend=out)
            return out
        if self.mode == "split":
            n_ext = self.part.n_halo if which == 0 else self.part.n_ohalo
            cur = torch.cuda.current_stream(self.device)
            self._side.wait_stream(cur)                       # the staged block is complete
            with torch.cuda.stream(self._side):
                if barrier:
                    self.blocks.barrier()
                halo = self._pulled(which, b, d, like, self._pull_blocks)
            own = self.blocks.own_view(b)[: n_in * d].view(n_in, d)
            # inner-source edges while the halo rows arrive
            K.scaled_copy_sum(self.g_ii, which == 1, own, None, None if n_ext else row_scale, out)
            cur.wait_stream(self._side)
            if n_ext:
                g = self.g_fh if which == 0 else self.g_bh
                K.scaled_copy_sum(g, which == 1, halo[:n_ext], None, row_scale, out, addend=out)
            return out
        if barrier:
            self.blocks.barrier()
        g = self.g_f if which == 0 else self.g_b
        if self.mode == "pull":
            halo = self._pulled(which, b, d, like)
            K.peer_copy_sum(g, which == 1, [self._own_base(b), halo.data_ptr()],
                            self.local_code[which], row_scale, out)
        else:
            K.peer_copy_sum(g, which == 1, self.blocks.bases(b), self.code[which], row_scale, out)
        return out

    def run(self, which, x, stage_scale, row_scale):
        shape = x.shape
        x2 = x.reshape(shape[0], -1)
        b = self.stage(which, x2, stage_scale)
        out = self.reduce(which, b, x2.shape[1], row_scale, x2, barrier=True)
        return out.view((self.part.n_inner,) + tuple(shape[1:]))


class LocalBlocks(object):
    """Staging blocks of several ranks inside ONE process / one GPU: the in-process stand-in for
    IpcBlocks (tests on a one-GPU box; `alias=True` maps every slot to this rank's own block --
    same kernel, same indices, no NVLink traffic -- which is how bench.py measures the exposed
    communication time of the real run)."""

    def __init__(self, group, rank):
        self._group, self.rank = group, rank
        self.max_feat = group["max_feat"]

    @classmethod
    def group(cls, world, rows_max, max_feat, device, alias_rank=None, ext_max=None):
        """ext_max (Partition.ext_max) also allocates the receive buffers of the push mode."""
        k = world if alias_rank is None else 1
        bufs = [[torch.zeros(max(rows_max, 1) * max_feat, device=device) for _ in range(4)]
                for _ in range(k)]
        g = {"bufs": bufs, "max_feat": int(max_feat), "world": world, "alias": alias_rank}
        if ext_max is not None:
            g["rbufs"] = [[torch.zeros(max(ext_max, 1) * max_feat, device=device)
                           for _ in range(4)] for _ in range(k)]
        return [cls(g, r) for r in range(world)]

    def _rmine(self):
        return self._group["rbufs"][0 if self._group["alias"] is not None else self.rank]

    def recv_view(self, b):
        return self._rmine()[b]

    def rbases(self, b):
        if self._group["alias"] is not None:
            return [self._rmine()[b].data_ptr()] * self._group["world"]
        return [bufs[b].data_ptr() for bufs in self._group["rbufs"]]

    @property
    def has_recv(self):
        return "rbufs" in self._group

    def _mine(self):
        return self._group["bufs"][0 if self._group["alias"] is not None else self.rank]

    def own_view(self, b):
        return self._mine()[b]

    def bases(self, b):
        if self._group["alias"] is not None:
            return [self._mine()[b].data_ptr()] * self._group["world"]
        return [bufs[b].data_ptr() for bufs in self._group["bufs"]]

    def barrier(self):
        # one process: every rank's kernels are on this device's streams (aliased blocks = the
        # all-local timing of ONE rank: nothing to wait for)
        if self._group["alias"] is None:
            torch.cuda.synchronize()

    def close(self):
        pass


class IpcBlocks(object):
    """Staging blocks in cudaMalloc'ed memory mapped into every rank of the node with CUDA IPC
    (csrc/peer.cu DGLB200PeerAlloc / PeerOpen); the barrier is one stream-ordered NCCL
    all-reduce of a flag."""

    def __init__(self):
        self._own, self._mapped, self._ptr = [], [], None

    @classmethod
    def create(cls, part, device, max_feat, required=False):
        """Collective over all ranks. Returns None (on every rank) when any rank cannot allocate
        or map the blocks; raises instead when `required`."""
        import ctypes
        from ._lib import LIB, check_call
        self = cls()
        self._lib, self._ct = LIB, ctypes
        self.max_feat, self.device = int(max_feat), torch.device(device)
        k, me = part.world, part.rank
        dev_index = self.device.index or 0
        n = max(part.rows_max, 1) * self.max_feat
        # buffers 0..3: staging blocks of this rank's rows (pull / split / direct modes);
        # 4..7: receive buffers of the push mode (largest halo of any rank)
        self.has_recv = part.push is not None
        sizes = [n] * 4 + ([max(part.ext_max, 1) * self.max_feat] * 4 if self.has_recv else [])
        nb = len(sizes)
        handles = torch.zeros(nb, 64, dtype=torch.uint8)
        ok, err = 1.0, None
        try:
            for i in range(nb):
                ptr, h = ctypes.c_void_p(), (ctypes.c_ubyte * 64)()
                check_call(LIB.DGLB200PeerAlloc(sizes[i] * 4, dev_index, ctypes.byref(ptr), h))
                self._own.append(ptr.value)
                handles[i] = torch.tensor(list(h), dtype=torch.uint8)
        except Exception as e:  # agreed on below: every rank takes the same path
            ok, err = 0.0, e
        if not self._agree(ok):
            self.close()
            if required:
                raise RuntimeError("IpcBlocks: allocation failed on some rank: %r" % (err,))
            return None
        hd = handles.to(self.device)
        allh = [torch.empty_like(hd) for _ in range(k)]
        dist.all_gather(allh, hd)
        self._ptr = [[0] * k for _ in range(nb)]
        try:
            for q in range(k):
                hq = allh[q].cpu()
                for i in range(nb):
                    if q == me:
                        self._ptr[i][q] = self._own[i]
                        continue
                    ptr = ctypes.c_void_p()
                    raw = (ctypes.c_ubyte * 64)(*hq[i].tolist())
                    check_call(LIB.DGLB200PeerOpen(raw, dev_index, ctypes.byref(ptr)))
                    self._ptr[i][q] = ptr.value
                    self._mapped.append(ptr.value)
        except Exception as e:
            ok, err = 0.0, e
        if not self._agree(ok):
            self.close()
            if required:
                raise RuntimeError("IpcBlocks: peer mapping failed on some rank: %r" % (err,))
            return None
        self._views = [torch.as_tensor(_RawCuda(p, m), device=self.device)
                       for p, m in zip(self._own, sizes)]
        self._flag = torch.zeros(1, device=self.device)
        return self

    def recv_view(self, b):
        return self._views[4 + b]

    def rbases(self, b):
        return self._ptr[4 + b]

    def _agree(self, ok):
        t = torch.tensor([ok], device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return t.item() > 0

    def own_view(self, b):
        return self._views[b]

    def bases(self, b):
        return self._ptr[b]

    def barrier(self):
        dist.all_reduce(self._flag)      # stream-ordered: the current stream waits, not the host

    def close(self):
        if self._mapped or self._own:
            torch.cuda.synchronize(self.device)
        for p in self._mapped:
            self._lib.DGLB200PeerClose(self._ct.c_void_p(p))
        for p in self._own:
            self._lib.DGLB200PeerFree(self._ct.c_void_p(p))
        self._mapped, self._own = [], []


class _RawCuda(object):
    """Exposes a cudaMalloc'ed buffer to torch (torch.as_tensor reads __cuda_array_interface__)."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4",
                                         "data": (ptr, False), "version": 3, "strides": None}


class PeerExchange(object):
    """Halo exchange through peer memory (csrc/peer.cu): every rank owns two forward buffers
    (halo rows it receives) and two backward buffers (halo gradients it receives), mapped into
    all other ranks of the node with CUDA IPC. An exchange is ONE kernel on the sender (gather
    the rows a peer needs and store them straight into its buffer over NVLink) followed by one
    stream-ordered NCCL barrier; buffers alternate, so a buffer is rewritten only after every
    rank has passed the barrier that follows its last readers."""

    def __init__(self, part, device, max_feat):
        import ctypes
        from ._lib import LIB, check_call
        self.part, self.device, self.max_feat = part, torch.device(device), int(max_feat)
        self._lib, self._check, self._ct = LIB, check_call, ctypes
        k, me = part.world, part.rank
        dev_index = self.device.index or 0
        self.total_send = int(part.send_cat.numel())
        sizes = [max(part.n_halo, 1) * max_feat] * 2 + [max(self.total_send, 1) * max_feat] * 2
        self._own, handles = [], torch.zeros(4, 64, dtype=torch.uint8)
        for i, n in enumerate(sizes):
            ptr = ctypes.c_void_p()
            h = (ctypes.c_ubyte * 64)()
            check_call(LIB.DGLB200PeerAlloc(int(n) * 4, dev_index, ctypes.byref(ptr), h))
            self._own.append(ptr.value)
            handles[i] = torch.tensor(list(h), dtype=torch.uint8)
        self._views = [torch.as_tensor(_RawCuda(p, n), device=self.device)
                       for p, n in zip(self._own, sizes)]
        hd = handles.to(self.device)
        allh = [torch.empty_like(hd) for _ in range(k)]
        dist.all_gather(allh, hd)
        counts = torch.tensor(part.halo_counts, dtype=torch.int64, device=self.device)
        allc = [torch.empty_like(counts) for _ in range(k)]
        dist.all_gather(allc, counts)
        need = torch.stack(allc).cpu()          # need[q][r] = rows rank q receives from rank r
        self._mapped = []
        self._peer = [[0] * k for _ in range(4)]  # [buffer][rank] -> device pointer
        for q in range(k):
            hq = allh[q].cpu()
            for i in range(4):
                if q == me:
                    self._peer[i][q] = self._own[i]
                    continue
                ptr = ctypes.c_void_p()
                raw = (ctypes.c_ubyte * 64)(*hq[i].tolist())
                check_call(LIB.DGLB200PeerOpen(raw, dev_index, ctypes.byref(ptr)))
                self._peer[i][q] = ptr.value
                self._mapped.append(ptr.value)
        i64 = ctypes.c_int64
        # forward: my rows for q land after the blocks of the owners below me in q's halo order
        self._f_beg = (i64 * (k + 1))(*([0] + torch.cumsum(torch.tensor(part.send_counts), 0).tolist()))
        self._f_dst = (i64 * k)(*[int(need[q, :me].sum()) for q in range(k)])
        # backward: my halo block of owner q lands after the blocks of the senders below me in
        # q's send order (q sends need[r][q] rows to rank r)
        self._b_beg = (i64 * (k + 1))(*([0] + torch.cumsum(torch.tensor(part.halo_counts), 0).tolist()))
        self._b_dst = (i64 * k)(*[int(need[:me, q].sum()) for q in range(k)])
        self._rows = part.send_cat.to(self.device, torch.int32).contiguous()
        self._flag = torch.zeros(1, device=self.device)
        self._side = torch.cuda.Stream(device=self.device)
        self._turn = [0, 0]
        self._k = k

    def usable(self, x):
        """Decided from rank-invariant facts only (dtype, rank, width): ranks that disagreed
        would issue mismatched collectives. Layout / alignment are fixed locally by _aligned()."""
        return x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and \
            x.shape[1] <= self.max_feat and x.shape[1] % 4 == 0   # 128-bit peer stores only

    @staticmethod
    def _aligned(x):
        x = x.contiguous()
        return x if x.data_ptr() % 16 == 0 else x.clone()

    def _push(self, x, which, beg, dst, rows):
        ct = self._ct
        b = 2 * which + self._turn[which]
        self._turn[which] ^= 1
        cur = torch.cuda.current_stream(self.device)
        self._side.wait_stream(cur)
        with torch.cuda.stream(self._side):
            peers = (ct.c_void_p * self._k)(*self._peer[b])
            self._check(self._lib.DGLB200HaloPush(
                ct.c_void_p(x.data_ptr()), int(x.shape[1]), self._k, peers, beg, dst,
                ct.c_void_p(rows.data_ptr()) if rows is not None else None,
                (self.part.rank + 1) % self._k, self.device.index or 0, ct.c_void_p(self._side.cuda_stream)))
            x.record_stream(self._side)
            work = dist.all_reduce(self._flag, async_op=True)  # barrier: all pushes have landed
        return work, self._views[b]

    def forward(self, x_inner):
        x_inner = self._aligned(x_inner)
        work, buf = self._push(x_inner, 0, self._f_beg, self._f_dst, self._rows)
        d = x_inner.shape[1]
        return work, buf[: self.part.n_halo * d].view(self.part.n_halo, d)

    def backward(self, g_halo):
        g_halo = self._aligned(g_halo)
        work, buf = self._push(g_halo, 1, self._b_beg, self._b_dst, None)
        d = g_halo.shape[1]
        return work, buf[: self.total_send * d].view(self.total_send, d)

    def close(self):
        torch.cuda.synchronize(self.device)
        for p in self._mapped:
            self._lib.DGLB200PeerClose(self._ct.c_void_p(p))
        for p in self._own:
            self._lib.DGLB200PeerFree(self._ct.c_void_p(p))
        self._mapped, self._own = [], []
