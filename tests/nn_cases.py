"""Module-level parity cases shared by tests/golden/make_golden_nn.py (which runs them through the
UNMODIFIED reference's dgl.nn.pytorch modules on CPU) and tests/test_gpu_nn.py (which runs them
through dgl_hack_b200.nn on the GPU). Pure numpy + torch; no reference / product imports.

A case = (module class name, constructor kwargs, graph spec, input spec). Parameters are NOT taken
from the modules' random initialisers: every parameter (sorted by name) is overwritten with
uniform(-0.5, 0.5) values from RandomState(case seed), so both sides hold identical weights.
"""
import numpy as np
import torch

import cases

ACT = {"relu": torch.relu, None: None}

# name -> (n_src, n_dst, n_edges, alpha, seed, bipartite)
GRAPHS = {
    "square": (300, 300, 6000, 0.6, 11, False),
    "sparse": (400, 400, 500, 0.0, 13, False),      # many zero-in-degree rows
    "bip": (200, 350, 5000, 0.4, 12, True),
    "cora": (2708, 2708, 10556, 0.3, 1, False),     # BASELINE configs[0] shape
}


def graph_arrays(name):
    ns, nd, ne, alpha, seed, bip = GRAPHS[name]
    src, dst = cases.powerlaw_graph(ns, nd, ne, alpha, seed)
    if name == "cora":  # self loops on every node, as the GCN example adds them
        src = np.concatenate([src[:ne - ns], np.arange(ns)])
        dst = np.concatenate([dst[:ne - ns], np.arange(ns)])
    return ns, nd, src, dst, bip


# (case name, module, kwargs, graph, input feature widths (src[, dst]), extra)
CASES = [
    ("gc_both_wide", "GraphConv", dict(in_feats=40, out_feats=12, norm="both", activation="relu"), "square", (40,), {}),
    ("gc_both_narrow", "GraphConv", dict(in_feats=10, out_feats=24, norm="both"), "square", (10,), {}),
    ("gc_right", "GraphConv", dict(in_feats=10, out_feats=24, norm="right", bias=False), "sparse", (10,), {}),
    ("gc_none", "GraphConv", dict(in_feats=16, out_feats=16, norm="none"), "sparse", (16,), {}),
    ("gc_bip", "GraphConv", dict(in_feats=12, out_feats=20, norm="both"), "bip", (12,), {}),
    ("sage_mean", "SAGEConv", dict(in_feats=20, out_feats=9, aggregator_type="mean"), "square", (20,), {}),
    ("sage_gcn", "SAGEConv", dict(in_feats=20, out_feats=9, aggregator_type="gcn", activation="relu"), "sparse", (20,), {}),
    ("sage_pool", "SAGEConv", dict(in_feats=20, out_feats=9, aggregator_type="pool"), "square", (20,), {}),
    ("sage_mean_bip", "SAGEConv", dict(in_feats=(14, 6), out_feats=9, aggregator_type="mean"), "bip", (14, 6), {}),
    ("gat", "GATConv", dict(in_feats=24, out_feats=8, num_heads=4), "square", (24,), {}),
    ("gat_res", "GATConv", dict(in_feats=24, out_feats=8, num_heads=3, residual=True, activation="relu", negative_slope=0.1), "sparse", (24,), {}),
    ("gat_bip", "GATConv", dict(in_feats=(12, 7), out_feats=5, num_heads=2), "bip", (12, 7), {}),
    ("rgc_basis", "RelGraphConv", dict(in_feat=16, out_feat=16, num_rels=7, regularizer="basis", num_bases=3), "square", (16,), {"norm": True}),
    ("rgc_basis_full", "RelGraphConv", dict(in_feat=12, out_feat=20, num_rels=5, regularizer="basis", self_loop=True, activation="relu"), "sparse", (12,), {"norm": False}),
    ("rgc_basis_big", "RelGraphConv", dict(in_feat=48, out_feat=40, num_rels=4, regularizer="basis", num_bases=2), "square", (48,), {"norm": True}),
    ("rgc_bdd", "RelGraphConv", dict(in_feat=16, out_feat=24, num_rels=6, regularizer="bdd", num_bases=4), "square", (16,), {"norm": True}),
]


def make_module(namespace, cls, kwargs):
    kw = dict(kwargs)
    if "activation" in kw:
        kw["activation"] = ACT[kw["activation"]]
    return getattr(namespace, cls)(**kw)


def set_params(module, seed):
    rs = np.random.RandomState(seed)
    with torch.no_grad():
        for name, p in sorted(module.named_parameters(), key=lambda kv: kv[0]):
            p.copy_(torch.from_numpy(rs.uniform(-0.5, 0.5, tuple(p.shape)).astype(np.float32)))


def case_inputs(idx, graph, widths, extra, num_rels=None):
    """numpy inputs of case #idx: feature matrices, edge types / norm, grad_out seed."""
    ns, nd, src, dst, bip = graph_arrays(graph)
    rs = np.random.RandomState(1000 + idx)
    feats = [rs.uniform(-1, 1, (ns, widths[0])).astype(np.float32)]
    if len(widths) == 2:
        feats.append(rs.uniform(-1, 1, (nd, widths[1])).astype(np.float32))
    out = {"feats": feats}
    if num_rels is not None:
        out["etypes"] = rs.randint(0, num_rels, size=src.shape[0]).astype(np.int64)
        if extra.get("norm"):
            out["norm"] = rs.uniform(0.1, 1, (src.shape[0], 1)).astype(np.float32)
    return out


def run_case(idx, namespace, make_graph, device):
    """Runs case #idx forward + backward. `namespace` provides the module classes,
    `make_graph(ns, nd, src, dst, bip)` the graph object. Returns dict of numpy arrays:
    out, grads of the input features and of every parameter."""
    name, cls, kwargs, graph, widths, extra = CASES[idx]
    ns, nd, src, dst, bip = graph_arrays(graph)
    g = make_graph(ns, nd, src, dst, bip)
    mod = make_module(namespace, cls, kwargs).to(device)
    set_params(mod, 500 + idx)
    inp = case_inputs(idx, graph, widths, extra, kwargs.get("num_rels"))
    feats = [torch.from_numpy(f).to(device).requires_grad_(True) for f in inp["feats"]]
    x = feats[0] if len(feats) == 1 else tuple(feats)
    if cls == "RelGraphConv":
        et = torch.from_numpy(inp["etypes"]).to(device)
        norm = torch.from_numpy(inp["norm"]).to(device) if "norm" in inp else None
        out = mod(g, x, et, norm)
    else:
        out = mod(g, x)
    gout = torch.from_numpy(cases.grad_out_for(tuple(out.shape), 77 + idx)).to(device)
    out.backward(gout)
    res = {"out": out.detach().cpu().numpy()}
    for i, f in enumerate(feats):
        res["gfeat%d" % i] = f.grad.cpu().numpy()
    for pname, p in mod.named_parameters():
        res["gparam/" + pname] = (p.grad if p.grad is not None else torch.zeros_like(p)).cpu().numpy()
    return res


# ---- edge_softmax: logits [E, 4, 1] / [E, 1]
SOFTMAX_CASES = [("square", (4, 1)), ("sparse", (1,)), ("bip", (3, 1)), ("square", (40, 1))]


def softmax_inputs(idx):
    graph, tail = SOFTMAX_CASES[idx]
    ns, nd, src, dst, bip = graph_arrays(graph)
    rs = np.random.RandomState(2000 + idx)
    z = rs.uniform(-3, 3, (src.shape[0],) + tail).astype(np.float32)
    go = rs.uniform(-1, 1, z.shape).astype(np.float32)
    return graph, z, go


# ---- BASELINE configs[0]: 2-layer GCN on the Cora-shaped graph, forward + loss + backward
def run_gcn_cora(namespace, make_graph, device, widths=(1433, 16, 7)):
    ns, nd, src, dst, bip = graph_arrays("cora")
    g = make_graph(ns, nd, src, dst, bip)
    rs = np.random.RandomState(4242)
    x = torch.from_numpy(rs.uniform(-1, 1, (ns, widths[0])).astype(np.float32)).to(device)
    labels = torch.from_numpy(rs.randint(0, widths[2], size=ns)).to(device)
    l1 = namespace.GraphConv(widths[0], widths[1], norm="both", activation=torch.relu).to(device)
    l2 = namespace.GraphConv(widths[1], widths[2], norm="both").to(device)
    set_params(l1, 1)
    set_params(l2, 2)
    with torch.no_grad():  # keep logits in a sane range: 1433 inputs at |w| <= 0.5 explode
        l1.weight.mul_(0.05)
    x.requires_grad_(True)
    logits = l2(g, l1(g, x))
    loss = torch.nn.functional.cross_entropy(logits, labels)
    loss.backward()
    return {"logits": logits.detach().cpu().numpy(), "loss": np.float32(loss.item()),
            "gx": x.grad.cpu().numpy()[:64].copy(),  # a slice: the full gradient is 15 MB
            "gw1": l1.weight.grad.cpu().numpy(), "gb1": l1.bias.grad.cpu().numpy(),
            "gw2": l2.weight.grad.cpu().numpy(), "gb2": l2.bias.grad.cpu().numpy()}
