"""A/B of the row schedule (B200_SPMM_SCHED=fine: degree classes for every row; default: rows
below 64 edges in natural order) on the products-shaped SBM graph: scaled_copy_sum fwd + bwd."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "profiles"))
    import bench_partitioned as BP
    import bench_configs as BC
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F
    dev = torch.device("cuda:0")
    n = (BP.PRODUCTS_N // 64) * 64
    src, dst = BP.sbm_graph(n, BP.PRODUCTS_E, 64, 0.85, 0.5, 5, dev)
    g = B.GraphIndex.from_coo(n, n, src, dst, dev)
    cs = torch.rand(n, device=dev) + 0.5
    rs = torch.rand(n, device=dev) + 0.5
    res = {}
    for d in (48, 100, 256):
        x = torch.rand(n, d, device=dev)
        res["scaled fwd D=%d" % d] = round(BC.timeit(lambda: F.scaled_copy_sum(g, x, cs, rs)), 4)
        res["plain fwd D=%d" % d] = round(BC.timeit(lambda: F.copy_reduce("sum", g, 0, x, n)), 4)
        from dgl_hack_b200 import kernel as K
        o = torch.empty_like(x)
        res["scaled bwd D=%d" % d] = round(BC.timeit(lambda: K.scaled_copy_sum(g, True, x, rs, cs, o)), 4)
        del x
    print(json.dumps(res))
else:
    out = {}
    for mode in ("fine", "auto"):
        env = dict(os.environ)
        if mode == "fine":
            env["B200_SPMM_SCHED"] = "fine"
        r = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
        out[mode] = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else r.stderr[-500:]
        print(mode, out[mode], flush=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "r2_sched_ab.json"), "w"), indent=1)
