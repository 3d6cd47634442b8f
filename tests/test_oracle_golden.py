"""Pins the CPU oracle (oracle/dgl_oracle.c) to outputs of the unmodified reference
(tests/golden/*.npz) and to the reference's own known-answer index fixtures."""
import numpy as np

import golden_util
from oracle import oracle as O
from oracle_adapter import OracleImpl


def test_oracle_kernel_graph_all_builtins():
    # 6 target pairs x 5 ops x 5 reducers x 3 broadcast modes + copy_u/copy_e x 5 reducers
    n = golden_util.replay_kernel_graph(OracleImpl)
    assert n == 450 + 10


def test_oracle_mid_graphs():
    golden_util.replay_mid(OracleImpl)


def test_oracle_id_mappings():
    golden_util.replay_mapped(OracleImpl)


# ---- known-answer fixtures transcribed from the reference's tests/cpp/test_spmat.cc ----
# CSR1 (test_spmat.cc:14-27):  [[0,1,1,0,0],[1,0,0,0,0],[0,0,1,1,0],[0,0,0,0,0]]
CSR1 = dict(n=4, m=5, indptr=[0, 2, 3, 5, 5], indices=[1, 2, 0, 2, 3], data=[0, 2, 3, 1, 4])
# CSR2 has a duplicate entry (test_spmat.cc:30-43)
CSR2 = dict(n=4, m=5, indptr=[0, 3, 4, 6, 6], indices=[1, 2, 2, 0, 2, 3],
            data=[0, 2, 5, 3, 1, 4])


def test_csr_transpose_known_answer():
    # TestCSRTranspose, test_spmat.cc:229-252 (on CSR2)
    bp, bi, bx = O.csr_transpose(CSR2["n"], CSR2["m"], CSR2["indptr"], CSR2["indices"],
                                 CSR2["data"])
    assert bp.tolist() == [0, 1, 2, 5, 6, 6]
    assert bi.tolist() == [1, 0, 0, 0, 2, 2]
    assert bx.tolist() == [3, 0, 2, 5, 1, 4]


def test_csr_to_coo_known_answer():
    # TestCSRToCOO, test_spmat.cc:254-281: rows of CSR2 = {0,0,0,1,2,2}
    assert O.csr_to_coo_rows(CSR1["indptr"]).tolist() == [0, 0, 1, 2, 2]
    assert O.csr_to_coo_rows(CSR2["indptr"]).tolist() == [0, 0, 0, 1, 2, 2]


def test_coo_to_csr_known_answer():
    # TestCOOToCSR, test_spmat.cc:357-421: COO1 (row {0,2,0,1,2}, col {1,2,2,0,3}) -> CSR1,
    # COO2 (row {0,2,0,1,2,0}, col {1,2,2,0,3,2}) -> CSR2 with data = COO position.
    ip, ix, dt = O.coo_to_csr(4, [0, 2, 0, 1, 2], [1, 2, 2, 0, 3])
    assert ip.tolist() == CSR1["indptr"] and ix.tolist() == CSR1["indices"]
    assert dt.tolist() == CSR1["data"]
    ip, ix, dt = O.coo_to_csr(4, [0, 2, 0, 1, 2, 0], [1, 2, 2, 0, 3, 2])
    assert ip.tolist() == CSR2["indptr"] and ix.tolist() == CSR2["indices"]
    assert dt.tolist() == CSR2["data"]


def test_empty_and_isolated():
    g = O.Graph(3, 3, [], [])
    x = np.ones((3, 4), np.float32)
    assert np.array_equal(O.copy_reduce("sum", g, 0, x, 3), np.zeros((3, 4), np.float32))
    mx = O.copy_reduce("max", g, 0, x, 3)
    assert (mx == np.finfo(np.float32).min).all()  # zero in-degree keeps the identity
    assert (O.copy_reduce("min", g, 0, x, 3) == np.finfo(np.float32).max).all()
    assert (O.copy_reduce("prod", g, 0, x, 3) == 1).all()
