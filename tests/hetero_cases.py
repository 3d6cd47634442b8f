"""Shared definitions of the HeteroGraphConv cases: used by tests/golden/make_golden_hetero.py
(the UNMODIFIED reference on CPU) and tests/test_gpu_hetero.py (the B200 path)."""
import numpy as np

NTYPES = {"user": 40, "item": 55, "tag": 13}
# (stype, etype, dtype, number of edges); 'likes-nothing' is empty on purpose (skipped relation)
RELATIONS = [("user", "follows", "user", 300), ("user", "buys", "item", 500),
             ("item", "bought-by", "user", 420), ("item", "has", "tag", 150),
             ("tag", "of", "item", 160), ("user", "likes-nothing", "tag", 0)]

# name, in_feats, out_feats, GraphConv norm, cross-relation aggregate, which ntypes get inputs
CASES = [("fused_both_sum", 6, 6, "both", "sum", ("user", "item", "tag")),
         ("fused_right_sum", 4, 9, "right", "sum", ("user", "item", "tag")),
         ("multfirst_max", 8, 4, "both", "max", ("user", "item", "tag")),
         ("mean_none", 5, 7, "none", "mean", ("user", "item", "tag")),
         ("missing_input_sum", 6, 6, "both", "sum", ("user", "item"))]


def edges():
    rs = np.random.RandomState(11)
    out = {}
    for s, e, d, n in RELATIONS:
        src = rs.randint(0, NTYPES[s], n)
        dst = rs.randint(0, NTYPES[d], n)
        if n:
            dst[: n // 6] = 3            # a hub per relation; many nodes get no edge of it
        out[(s, e, d)] = (src.astype(np.int64), dst.astype(np.int64))
    return out


def case_arrays(idx):
    """inputs {ntype: [n, in]}, parameters {etype: (weight [in, out], bias [out])}, output
    gradients {ntype: [n, out]} of case idx."""
    name, fin, fout, norm, agg, have = CASES[idx]
    rs = np.random.RandomState(100 + idx)
    x = {t: rs.uniform(-1, 1, (NTYPES[t], fin)).astype(np.float32) for t in sorted(NTYPES)}
    params = {e: (rs.uniform(-1, 1, (fin, fout)).astype(np.float32),
                  rs.uniform(-1, 1, (fout,)).astype(np.float32)) for _, e, _, _ in RELATIONS}
    gout = {t: rs.uniform(-1, 1, (NTYPES[t], fout)).astype(np.float32) for t in sorted(NTYPES)}
    return {t: x[t] for t in have}, params, gout


def run_case(idx, make_graph, graphconv, heteroconv, torch, device):
    """Builds the modules with the case's parameters, runs forward + backward; returns a dict of
    numpy arrays: out/<ntype>, gx/<ntype>, gw/<etype>, gb/<etype>."""
    name, fin, fout, norm, agg, have = CASES[idx]
    x_np, params, gout = case_arrays(idx)
    g = make_graph(edges(), NTYPES)
    mods = {}
    for _, e, _, _ in RELATIONS:
        m = graphconv(fin, fout, norm=norm)
        with torch.no_grad():
            m.weight.copy_(torch.from_numpy(params[e][0]))
            m.bias.copy_(torch.from_numpy(params[e][1]))
        mods[e] = m
    conv = heteroconv(mods, aggregate=agg).to(device)
    x = {t: torch.from_numpy(a).to(device).requires_grad_(True) for t, a in x_np.items()}
    out = conv(g, x)
    loss = sum((out[t] * torch.from_numpy(gout[t]).to(device)).sum() for t in out)
    loss.backward()
    res = {}
    for t, o in out.items():
        res["out/" + t] = o.detach().cpu().numpy()
    for t, a in x.items():
        res["gx/" + t] = a.grad.detach().cpu().numpy() if a.grad is not None else \
            np.zeros_like(x_np[t])
    for e, m in conv.mods.items():
        if m.weight.grad is not None:
            res["gw/" + e] = m.weight.grad.detach().cpu().numpy()
            res["gb/" + e] = m.bias.grad.detach().cpu().numpy()
    return res, conv
