"""Seeded inputs for the partial message-passing parity tests (send_and_recv / pull /
edge_softmax over an edge subset), shared by tests/golden/make_golden_partial.py (reference) and
tests/test_gpu_partial.py (this library). Each case returns the resulting node / edge column and
the gradients of sum(column * gout) w.r.t. the node feature h and the edge feature w."""
import numpy as np
import torch

import cases

N, E = 300, 6000


def inputs():
    src, dst = cases.powerlaw_graph(N, N, E, 0.6, 11)
    rs = np.random.RandomState(77)
    h = rs.uniform(-1, 1, (N, 5)).astype(np.float32)
    w = rs.uniform(0.1, 1, (E, 5)).astype(np.float32)
    w1 = rs.uniform(0.1, 1, (E, 1)).astype(np.float32)
    eids = rs.permutation(E)[:700].astype(np.int64)
    nodes = rs.permutation(N)[:60].astype(np.int64)
    logits = rs.uniform(-2, 2, (700, 3, 1)).astype(np.float32)
    return src, dst, h, w, w1, eids, nodes, logits


# (name, kind, message builtin name, message args, reducer, out field)
CASES = [
    ("snr_copy_sum", "snr", "copy_src", ("h", "m"), "sum"),
    ("snr_umule_max", "snr", "u_mul_e", ("h", "w", "m"), "max"),
    ("snr_copye_mean", "snr", "copy_e", ("w", "m"), "mean"),
    ("snr_uadde_bcast_sum", "snr", "u_add_e", ("h", "w1", "m"), "sum"),
    ("pull_umule_sum", "pull", "u_mul_e", ("h", "w", "m"), "sum"),
    ("pull_copy_min", "pull", "copy_src", ("h", "m"), "min"),
]


def run(fn, make_graph, edge_softmax, device):
    """`fn`: builtin-function module; make_graph(n, src, dst) -> graph with ndata/edata;
    edge_softmax(graph, logits, eids)."""
    src, dst, h, w, w1, eids, nodes, logits = inputs()
    t = lambda a: torch.from_numpy(a).to(device)  # noqa: E731
    out = {}
    for name, kind, mname, margs, red in CASES:
        g = make_graph(N, src, dst)
        ht, wt, w1t = t(h).requires_grad_(True), t(w).requires_grad_(True), t(w1).requires_grad_(True)
        g.ndata["h"] = ht
        g.edata["w"] = wt
        g.edata["w1"] = w1t
        mfunc = getattr(fn, mname)(*margs)
        rfunc = getattr(fn, red)("m", "out")
        if kind == "snr":
            g.send_and_recv(t(eids), mfunc, rfunc)
        else:
            g.pull(t(nodes), mfunc, rfunc)
        col = g.ndata["out"]
        gout = t(cases.grad_out_for(tuple(col.shape), 5))
        (col * gout).sum().backward()
        out[name + "/out"] = col.detach().cpu().numpy()
        for k, x in (("gh", ht), ("gw", wt), ("gw1", w1t)):
            out[name + "/" + k] = (x.grad if x.grad is not None else torch.zeros_like(x)).cpu().numpy()
    g = make_graph(N, src, dst)
    z = t(logits).requires_grad_(True)
    a = edge_softmax(g, z, t(eids))
    go = t(cases.grad_out_for(tuple(a.shape), 6))
    a.backward(go)
    out["softmax_eids/out"] = a.detach().cpu().numpy()
    out["softmax_eids/gz"] = z.grad.cpu().numpy()
    return out
