"""Direct `update_all / apply_edges` over the B200 kernels -- the caller side of the hot path.

The reference turns `g.update_all(fn.copy_src('h','m'), fn.sum('m','h'))` into a scheduler
`Program` (python/dgl/runtime/scheduler.py:196-252 schedule_update_all, :334-375
schedule_apply_edges, runtime/spmv.py:15-90 gen_v2v_spmv_schedule / gen_e2v_spmv_schedule),
executes it node by node (runtime/ir/executor.py:1000-1248) and writes the result through
FrameRef. SURVEY §8(f) rank 1 asks for the bypass: here a (builtin message, builtin reduce) pair
is resolved to ONE `binary_reduce` / `copy_reduce` call with no IR, no frame temporaries and no
id maps, and the result column is stored by reference.

`Graph` carries only what those calls need of `DGLGraph` / the bipartite `DGLHeteroGraph` block
(python/dgl/graph.py, heterograph.py): a device `GraphIndex`, node/edge feature dicts, degrees,
`local_var / local_scope`. Mutation, queries, batching and UDF reducers (degree bucketing) are out
of scope. Same semantics as the reference where it matters for results:

* every destination row is written; rows without in-edges get the reducer's identity
  (0 for sum/mean, -/+FLT_MAX for max/min, 1 for prod) -- binary_reduce_common.h:447-470;
* a message UDF (callable taking an `EdgeBatch`) is evaluated on all edges in edge-id order and
  its output reduced with `copy_reduce(EDGE)`, as the reference does (scheduler.py:801-960);
* `apply_node_func` receives a `NodeBatch` whose `data` holds the stored node features updated
  with the reduced fields (scheduler.py _apply_with_accum).
"""
from contextlib import contextmanager

import torch as th

from . import backend as F
from . import function as fn
from ._lib import DGLError
from .graph_index import GraphIndex

__all__ = ["Graph", "DGLGraph", "bipartite", "EdgeBatch", "NodeBatch"]


class EdgeBatch(object):
    """What a message / edge UDF sees: `src`, `dst`, `data` feature dicts gathered per edge (in
    edge-id order) -- python/dgl/udf.py EdgeBatch."""

    def __init__(self, graph, src_data, dst_data, edge_data):
        self._g = graph
        self.src, self.dst, self.data = src_data, dst_data, edge_data

    def edges(self):
        return self._g._src, self._g._dst, th.arange(self._g.number_of_edges(),
                                                     device=self._g.device)

    def batch_size(self):
        return self._g.number_of_edges()

    def __len__(self):
        return self.batch_size()


class NodeBatch(object):
    """What an apply-node UDF sees -- python/dgl/udf.py NodeBatch (no mailbox: reducers are
    builtins here)."""

    def __init__(self, graph, data):
        self._g = graph
        self.data = data

    def nodes(self):
        return th.arange(self._g.number_of_dst_nodes(), device=self._g.device)

    def batch_size(self):
        return self._g.number_of_dst_nodes()

    def __len__(self):
        return self.batch_size()


class _LazyGather(dict):
    """Feature dict that gathers `frame[key][index]` on first access, so a UDF that reads one
    field does not pay for the others."""

    def __init__(self, frame, index):
        super(_LazyGather, self).__init__()
        self._frame, self._index = frame, index

    def __missing__(self, key):
        val = self._frame[key].index_select(0, self._index)
        self[key] = val
        return val

    def __contains__(self, key):
        return key in self._frame

    def keys(self):
        return self._frame.keys()


class Graph(object):
    """Immutable graph with node / edge feature storage on one CUDA device.

    Graph(src, dst, num_nodes)                     homogeneous (srcdata is dstdata is ndata)
    Graph(src, dst, num_src=..., num_dst=...)      one bipartite block (unit graph)
    Edge id = position in (src, dst). `gidx` may pass an existing GraphIndex to share.
    """

    def __init__(self, src, dst, num_nodes=None, num_src=None, num_dst=None, device="cuda:0",
                 gidx=None):
        self.device = th.device(device)
        src = th.as_tensor(src, dtype=th.int64)
        dst = th.as_tensor(dst, dtype=th.int64)
        if num_nodes is None and num_src is None:
            num_nodes = int(max(src.max().item(), dst.max().item())) + 1 if src.numel() else 0
        self._bipartite = num_nodes is None
        if self._bipartite:
            if num_src is None or num_dst is None:
                raise DGLError("a bipartite block needs both num_src and num_dst")
            self._ns, self._nd = int(num_src), int(num_dst)
        else:
            self._ns = self._nd = int(num_nodes)
        self._src, self._dst = src.to(self.device), dst.to(self.device)
        self._gidx = gidx if gidx is not None else GraphIndex.from_coo(
            self._ns, self._nd, self._src, self._dst, self.device)
        self._node_frame = {}
        self._dst_frame = {} if self._bipartite else self._node_frame
        self._edge_frame = {}

    # ---- structure -----------------------------------------------------------------------
    @property
    def gidx(self):
        """The device graph index (get_immutable_gidx(ctx) / get_unitgraph(0, ctx))."""
        return self._gidx

    def is_homograph(self):
        return not self._bipartite

    def number_of_nodes(self):
        if self._bipartite:
            raise DGLError("number_of_nodes() is ambiguous on a bipartite block")
        return self._nd

    def number_of_src_nodes(self):
        return self._ns

    def number_of_dst_nodes(self):
        return self._nd

    def number_of_edges(self):
        return self._gidx.number_of_edges()

    def edges(self):
        return self._src, self._dst

    def all_edges(self):
        return self.edges()

    def in_degrees(self):
        return self._gidx.in_degrees()

    def out_degrees(self):
        return self._gidx.out_degrees()

    def to(self, device):
        if th.device(device) != self.device:
            raise DGLError("Graph lives on %s; build a new Graph for another device" % self.device)
        return self

    # ---- feature storage -----------------------------------------------------------------
    @property
    def ndata(self):
        if self._bipartite:
            raise DGLError("use srcdata / dstdata on a bipartite block")
        return self._node_frame

    @property
    def srcdata(self):
        return self._node_frame

    @property
    def dstdata(self):
        return self._dst_frame

    @property
    def edata(self):
        return self._edge_frame

    def local_var(self):
        """A graph sharing the structure whose feature writes do not reach the caller's frames
        (python/dgl/graph.py local_var)."""
        g = object.__new__(Graph)
        g.__dict__.update(self.__dict__)
        g._node_frame = dict(self._node_frame)
        g._dst_frame = dict(self._dst_frame) if self._bipartite else g._node_frame
        g._edge_frame = dict(self._edge_frame)
        return g

    @contextmanager
    def local_scope(self):
        saved = (self._node_frame, self._dst_frame, self._edge_frame)
        self._node_frame = dict(saved[0])
        self._dst_frame = dict(saved[1]) if self._bipartite else self._node_frame
        self._edge_frame = dict(saved[2])
        try:
            yield
        finally:
            self._node_frame, self._dst_frame, self._edge_frame = saved

    # ---- message passing -----------------------------------------------------------------
    def _frame_of(self, target):
        return (self._node_frame, self._dst_frame, self._edge_frame)[target]

    def _col(self, target, field):
        frame = self._frame_of(target)
        if field not in frame:
            raise DGLError('no %s feature named "%s"' % (("source node", "destination node",
                                                           "edge")[target], field))
        return frame[field]

    def _edge_batch(self):
        return EdgeBatch(self, _LazyGather(self._node_frame, self._src),
                         _LazyGather(self._dst_frame, self._dst), self._edge_frame)

    def _message_only(self, mfn):
        """Edge output of one builtin message function: reducer `none`
        (scheduler.py:334-375 -> _gen_send -> BINARY_REDUCE / COPY_REDUCE with 'none')."""
        ne = self.number_of_edges()
        if isinstance(mfn, fn.CopyMessageFunction):
            if mfn.target == fn.TargetCode.EDGE:
                return self._col(2, mfn.in_field)
            return F.copy_reduce("none", self._gidx, mfn.target, self._col(mfn.target, mfn.in_field),
                                 ne)
        return F.binary_reduce("none", mfn.binary_op, self._gidx, mfn.lhs, mfn.rhs,
                               self._col(mfn.lhs, mfn.lhs_field), self._col(mfn.rhs, mfn.rhs_field),
                               ne)

    def apply_edges(self, func):
        """Writes `func`'s output to the edge features of ALL edges."""
        funcs = func if isinstance(func, (list, tuple)) else [func]
        if all(isinstance(f, fn.MessageFunction) for f in funcs):
            for f in funcs:
                self._edge_frame[f.out_field] = self._message_only(f)
            return
        if len(funcs) != 1 or not callable(funcs[0]):
            raise DGLError("apply_edges takes builtin message functions or one callable")
        self._edge_frame.update(funcs[0](self._edge_batch()))

    def update_all(self, message_func, reduce_func, apply_node_func=None):
        """Send along all edges, reduce at every destination, optionally apply."""
        mfuncs = message_func if isinstance(message_func, (list, tuple)) else [message_func]
        rfuncs = reduce_func if isinstance(reduce_func, (list, tuple)) else [reduce_func]
        if not all(isinstance(r, fn.ReduceFunction) for r in rfuncs):
            raise DGLError("only builtin reduce functions are supported (degree-bucketed UDF "
                           "reducers are outside this path)")
        nd = self._nd
        reduced = {}
        if all(isinstance(m, fn.MessageFunction) for m in mfuncs):
            by_field = {m.out_field: m for m in mfuncs}
            for r in rfuncs:
                if r.msg_field not in by_field:
                    raise DGLError('Reduce function requires message field "%s", but no message '
                                   'function generates it.' % r.msg_field)
                m = by_field[r.msg_field]
                if isinstance(m, fn.CopyMessageFunction):
                    reduced[r.out_field] = F.copy_reduce(
                        r.name, self._gidx, m.target, self._col(m.target, m.in_field), nd)
                else:
                    reduced[r.out_field] = F.binary_reduce(
                        r.name, m.binary_op, self._gidx, m.lhs, m.rhs,
                        self._col(m.lhs, m.lhs_field), self._col(m.rhs, m.rhs_field), nd)
        else:
            if len(mfuncs) != 1 or not callable(mfuncs[0]):
                raise DGLError("message_func must be builtin functions or one callable")
            msgs = mfuncs[0](self._edge_batch())
            for r in rfuncs:
                if r.msg_field not in msgs:
                    raise DGLError('Reduce function requires message field "%s", but the message '
                                   'function does not return it.' % r.msg_field)
                reduced[r.out_field] = F.copy_reduce(r.name, self._gidx, fn.TargetCode.EDGE,
                                                     msgs[r.msg_field], nd)
        if apply_node_func is not None:
            data = dict(self._dst_frame)
            data.update(reduced)
            reduced.update(apply_node_func(NodeBatch(self, data)))
        self._dst_frame.update(reduced)

    # ---- partial message passing (id-mapped kernels) --------------------------------------
    def edge_subindex(self, eids):
        """Unit graph over the selected edges with the parent's node ids, plus the (forward,
        backward) edge-id maps the kernels take: what runtime/spmv.py:146-180
        build_gidx_and_mapping_uv builds for send_and_recv / pull (forward map in out-CSR order,
        backward map in in-CSR order), constructed on the device."""
        eids = th.as_tensor(eids, dtype=th.int64, device=self.device)
        gidx = GraphIndex.from_coo(self._ns, self._nd, self._src[eids], self._dst[eids],
                                   self.device)
        e32 = eids.to(th.int32)
        fwd = e32[gidx.get_adj_csr(True)[2].long()]
        bwd = e32[gidx.get_adj_csr(False)[2].long()]
        return gidx, (fwd, bwd)

    def _reduce_on(self, eids, recv, message_func, reduce_func, apply_node_func):
        mfuncs = message_func if isinstance(message_func, (list, tuple)) else [message_func]
        rfuncs = reduce_func if isinstance(reduce_func, (list, tuple)) else [reduce_func]
        if not (all(isinstance(m, fn.MessageFunction) for m in mfuncs)
                and all(isinstance(r, fn.ReduceFunction) for r in rfuncs)):
            raise DGLError("partial message passing takes builtin message and reduce functions")
        gidx, emap = self.edge_subindex(eids)
        o2n = th.full((self._nd,), -1, dtype=th.int32, device=self.device)
        o2n[recv] = th.arange(recv.numel(), dtype=th.int32, device=self.device)
        omap, none, n_out = (o2n, o2n), (None, None), int(recv.numel())
        by_field = {m.out_field: m for m in mfuncs}
        reduced = {}
        for r in rfuncs:
            if r.msg_field not in by_field:
                raise DGLError('Reduce function requires message field "%s", but no message '
                               'function generates it.' % r.msg_field)
            m = by_field[r.msg_field]
            if isinstance(m, fn.CopyMessageFunction):
                reduced[r.out_field] = F.copy_reduce(
                    r.name, gidx, m.target, self._col(m.target, m.in_field), n_out,
                    emap if m.target == fn.TargetCode.EDGE else none, omap)
            else:
                reduced[r.out_field] = F.binary_reduce(
                    r.name, m.binary_op, gidx, m.lhs, m.rhs, self._col(m.lhs, m.lhs_field),
                    self._col(m.rhs, m.rhs_field), n_out,
                    emap if m.lhs == fn.TargetCode.EDGE else none,
                    emap if m.rhs == fn.TargetCode.EDGE else none, omap)
        if apply_node_func is not None:
            data = {k: v.index_select(0, recv) for k, v in self._dst_frame.items()}
            data.update(reduced)
            reduced.update(apply_node_func(NodeBatch(self, data)))
        for k, v in reduced.items():  # WRITE_ROW_: only the receiving rows change
            if k in self._dst_frame and self._dst_frame[k].shape[1:] == v.shape[1:]:
                base = self._dst_frame[k]
            else:
                base = v.new_zeros((self._nd,) + tuple(v.shape[1:]))
            self._dst_frame[k] = base.index_copy(0, recv, v)

    def send_and_recv(self, edges, message_func, reduce_func, apply_node_func=None):
        """Messages along the edges with ids `edges`, reduced at their destinations; only those
        destination rows are written (scheduler.py schedule_snr)."""
        eids = th.as_tensor(edges, dtype=th.int64, device=self.device)
        if eids.numel() == 0:
            return
        recv = th.unique(self._dst[eids])
        self._reduce_on(eids, recv, message_func, reduce_func, apply_node_func)

    def pull(self, v, message_func, reduce_func, apply_node_func=None):
        """All in-edges of the nodes `v` send, `v` receive (scheduler.py schedule_pull): nodes of
        `v` without in-edges get the reducer's identity."""
        from . import sampling
        v = th.unique(th.as_tensor(v, dtype=th.int64, device=self.device))
        if v.numel() == 0:
            return
        eids = sampling.in_subgraph(self, v).eid
        if eids.numel() == 0:
            if apply_node_func is not None:
                data = {k: x.index_select(0, v) for k, x in self._dst_frame.items()}
                for k, x in apply_node_func(NodeBatch(self, data)).items():
                    self._dst_frame[k] = self._dst_frame[k].index_copy(0, v, x)
            return
        self._reduce_on(eids, v, message_func, reduce_func, apply_node_func)

    def push(self, u, message_func, reduce_func, apply_node_func=None):
        """All out-edges of the nodes `u` send, their destinations receive (scheduler.py
        schedule_push = send_and_recv over the out-edges)."""
        u = th.as_tensor(u, dtype=th.int64, device=self.device)
        mark = th.zeros(self._ns, dtype=th.bool, device=self.device)
        mark[u] = True
        self.send_and_recv(th.nonzero(mark[self._src]).flatten(), message_func, reduce_func,
                           apply_node_func)

    def apply_nodes(self, func):
        self._dst_frame.update(func(NodeBatch(self, dict(self._dst_frame))))

    def __repr__(self):
        if self._bipartite:
            return "Graph(num_src=%d, num_dst=%d, num_edges=%d, device=%s)" % (
                self._ns, self._nd, self.number_of_edges(), self.device)
        return "Graph(num_nodes=%d, num_edges=%d, device=%s)" % (
            self._nd, self.number_of_edges(), self.device)


DGLGraph = Graph


def bipartite(src, dst, num_src, num_dst, device="cuda:0"):
    """One bipartite block (dgl.bipartite / a NodeFlow block as a unit graph)."""
    return Graph(src, dst, num_src=num_src, num_dst=num_dst, device=device)
