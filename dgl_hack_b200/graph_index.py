"""Device-resident graph index consumed by the kernels.

Host-side mirror of what the kernels see of python/dgl/graph_index.py::GraphIndex /
heterograph_index.py (get_immutable_gidx(ctx), get_unitgraph(ctx): in-CSR + out-CSR, int32, on
the device) -- nothing else of the graph API is rebuilt. The CSRs are produced ON THE GPU by
DGLB200GraphCreateFromCOO (bit-identical to aten::COOToCSR + CSRTranspose), replacing the
reference's single-threaded CPU conversion + H2D copy (graph_index.py:671-686,
src/graph/immutable_graph.cc:548-590).
"""
import ctypes

import torch

from . import _lib
from ._lib import LIB, check_call


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class GraphIndex(object):
    """in-CSR + out-CSR of a (num_src x num_dst) graph on one CUDA device."""

    def __init__(self, handle, device):
        self._handle = handle
        self.device = torch.device(device)
        info = (ctypes.c_int64 * 4)()
        check_call(LIB.DGLB200GraphInfo(handle, info))
        self.num_src, self.num_dst, self._num_edges = int(info[0]), int(info[1]), int(info[2])

    # ---- constructors
    @classmethod
    def from_coo(cls, num_src, num_dst, src, dst, device="cuda:0"):
        """src/dst: int64 torch tensors (CPU or CUDA) or anything torch.as_tensor accepts;
        edge id = position. Equivalent of graph_index.from_coo(...).get_immutable_gidx(ctx)
        (or create_unitgraph_from_coo(...).get_unitgraph(ctx) when num_src != num_dst)."""
        device = torch.device(device)
        if device.type != "cuda":
            raise _lib.DGLError("GraphIndex lives on a CUDA device; there is no CPU path")
        src = torch.as_tensor(src, dtype=torch.int64).contiguous()
        dst = torch.as_tensor(dst, dtype=torch.int64).contiguous()
        if src.shape != dst.shape or src.dim() != 1:
            raise _lib.DGLError("src and dst must be 1-D arrays of equal length")
        # device tensors are read in stream order (torch's allocator keeps a freed block out of
        # reach of later work on other streams only); host buffers must outlive the copy
        on_device = src.is_cuda and dst.is_cuda
        handle = _lib.GraphHandle()
        dev_index = device.index or 0
        with torch.cuda.device(dev_index):
            check_call(LIB.DGLB200GraphCreateFromCOO(
                int(num_src), int(num_dst), int(src.shape[0]), ctypes.c_void_p(src.data_ptr()),
                ctypes.c_void_p(dst.data_ptr()), dev_index, _stream(device),
                ctypes.byref(handle)))
            if not on_device:
                torch.cuda.current_stream(device).synchronize()  # src/dst may be freed by the caller
        return cls(handle, device)

    @classmethod
    def from_csr(cls, num_src, num_dst, in_csr, out_csr, device="cuda:0"):
        """in_csr / out_csr: (indptr, indices, eids) triples (numpy / torch, int32 or int64), e.g.
        the outputs of the reference's _CAPI_DGLGraphGetAdj(g, transpose, "csr")."""
        device = torch.device(device)
        arrs = [torch.as_tensor(a).contiguous() for a in tuple(in_csr) + tuple(out_csr)]
        bits = 64 if arrs[0].dtype == torch.int64 else 32
        want = torch.int64 if bits == 64 else torch.int32
        arrs = [a.to(want) for a in arrs]
        handle = _lib.GraphHandle()
        dev_index = device.index or 0
        with torch.cuda.device(dev_index):
            check_call(LIB.DGLB200GraphCreateFromCSR(
                int(num_src), int(num_dst), int(arrs[1].shape[0]), bits,
                *[ctypes.c_void_p(a.data_ptr()) for a in arrs], dev_index, _stream(device),
                ctypes.byref(handle)))
        return cls(handle, device)

    @classmethod
    def from_coo_with_type(cls, num_nodes, src, dst, etypes, num_types=None, device="cuda:0"):
        """DGLGraph.add_edges_with_type(src, dst, etypes) (the fork's typed multigraph,
        Graph::AddEdgesWithType src/graph/graph.cc:140-165) as an immutable device index."""
        g = cls.from_coo(num_nodes, num_nodes, src, dst, device)
        g.set_edge_types(etypes, num_types)
        return g

    def set_edge_types(self, etypes, num_types=None):
        """Attaches a relation id to every edge id and builds the typed adjacencies
        (Graph::GetCsrSortedByEdgeType, graph.cc:690-741) on the device."""
        et = torch.as_tensor(etypes)
        if et.dtype not in (torch.int32, torch.int64):
            et = et.to(torch.int64)
        et = et.contiguous()
        if et.dim() != 1 or et.shape[0] != self._num_edges:
            raise _lib.DGLError("etypes must hold one type id per edge")
        if num_types is None:
            num_types = int(et.max().item()) + 1 if et.numel() else 1
        with torch.cuda.device(self.device):
            check_call(LIB.DGLB200GraphSetEdgeTypes(
                self._handle, ctypes.c_void_p(et.data_ptr()), 64 if et.dtype == torch.int64 else 32,
                int(num_types), _stream(self.device)))
        self.num_types = int(num_types)
        self.__dict__.pop("_rel_segments", None)     # derived from the previous typing

    def get_csr_sorted_by_edge_type(self, transpose):
        """[ranges, ids, eids, types] int32 CUDA tensors, as Graph::GetCsrSortedByEdgeType
        (rows ordered by (type, edge id))."""
        n_rows = self.num_src if transpose else self.num_dst
        e = self._num_edges
        out = [torch.empty((n,), dtype=torch.int32, device=self.device)
               for n in (n_rows + 1, e, e, e)]
        with torch.cuda.device(self.device):
            check_call(LIB.DGLB200GraphCopyCsrSortedByEdgeType(
                self._handle, int(transpose), *[ctypes.c_void_p(t.data_ptr()) for t in out],
                _stream(self.device)))
        return out

    def relation_segments(self):
        """(segment graph, seg_dst, seg_type): the relation-segment graph of a typed graph
        (DGLB200GraphCreateRelationSegments; cached). One destination row per (relation, node)
        pair that receives edges; ONE copy_reduce / binary_reduce on it = the reference's loop
        over the relations' unit graphs (python/dgl/nn/pytorch/hetero.py:120-171). seg_dst /
        seg_type are int64 CUDA tensors."""
        cached = self.__dict__.get("_rel_segments")
        if cached is None:
            handle = _lib.GraphHandle()
            with torch.cuda.device(self.device):
                check_call(LIB.DGLB200GraphCreateRelationSegments(
                    self._handle, _stream(self.device), ctypes.byref(handle)))
                sg = GraphIndex(handle, self.device)
                tabs = [torch.empty((sg.num_dst,), dtype=torch.int32, device=self.device)
                        for _ in range(2)]
                check_call(LIB.DGLB200GraphCopyRelationSegments(
                    handle, *[ctypes.c_void_p(t.data_ptr()) for t in tabs],
                    _stream(self.device)))
            cached = self.__dict__["_rel_segments"] = (sg, tabs[0].long(), tabs[1].long())
        return cached

    def __del__(self):
        h, self._handle = getattr(self, "_handle", None), None
        if h:
            try:
                LIB.DGLB200GraphFree(h)
            except Exception:  # interpreter shutdown
                pass

    # ---- queries (graph_index.py number_of_nodes / number_of_edges / in_degrees / get_adj)
    @property
    def handle(self):
        return self._handle

    def number_of_edges(self):
        return self._num_edges

    def number_of_nodes(self):
        return self.num_dst

    def ctx(self):
        return self.device

    def _csr_view(self, transpose):
        n_rows = self.num_src if transpose else self.num_dst
        out = [torch.empty((n,), dtype=torch.int32, device=self.device)
               for n in (n_rows + 1, self._num_edges, self._num_edges)]
        with torch.cuda.device(self.device):
            check_call(LIB.DGLB200GraphCopyCSR(
                self._handle, int(transpose), *[ctypes.c_void_p(t.data_ptr()) for t in out],
                _stream(self.device)))
        return tuple(out)

    def get_adj_csr(self, transpose):
        """(indptr, indices, edge_ids) int32 CUDA tensors; transpose=False -> in-CSR, exactly
        _CAPI_DGLGraphGetAdj(g, transpose, "csr") (src/graph/graph_apis.cc:285-292)."""
        return self._csr_view(transpose)

    @property
    def in_csr(self):
        return tuple(t.cpu().numpy() for t in self._csr_view(False))

    @property
    def out_csr(self):
        return tuple(t.cpu().numpy() for t in self._csr_view(True))

    def _degrees(self, which):
        cache = self.__dict__.setdefault("_deg_cache", {})
        if which not in cache:
            cache[which] = self._degrees_uncached(which)
        return cache[which]

    def mean_divisor(self, dtype):
        """clamp(in_degree, min=1) as a float vector, cached (used by the 'mean' reducer)."""
        cache = self.__dict__.setdefault("_deg_cache", {})
        key = ("mean", dtype)
        if key not in cache:
            cache[key] = self._degrees(0).to(dtype).clamp(min=1)
        return cache[key]

    def _degrees_uncached(self, which):
        n = self.num_src if which else self.num_dst
        out = torch.empty((n,), dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            check_call(LIB.DGLB200GraphDegrees(self._handle, which, ctypes.c_void_p(out.data_ptr()),
                                               _stream(self.device)))
        return out

    def in_degrees(self):
        """A fresh tensor per call (as the reference): in-place edits by the caller must not reach
        the cached degrees behind `mean_divisor` and GraphConv's norms."""
        return self._degrees(0).clone()

    def out_degrees(self):
        return self._degrees(1).clone()

    def index_bytes(self):
        """Bytes of index arrays resident on the device right now (DGLB200GraphIndexBytes)."""
        n = ctypes.c_int64(0)
        check_call(LIB.DGLB200GraphIndexBytes(self._handle, ctypes.byref(n)))
        return int(n.value)
