"""GPU: connected components through the C ABI against the oracle's min-id union-find --
bit-exact canonical labels (smallest vertex id of the component) and component counts."""
import numpy as np
import pytest

from conftest import load_golden, rmat_graph

pytestmark = pytest.mark.gpu


def test_shared_lib_fixture_through_reference_api(gr, shared_lib):
    """shared_lib_tests/test_cc.c; golden CMakeLists.txt:223-225: node 1 -> component 0"""
    labels = gr.gunrock_cc_func(shared_lib["row_offsets"], shared_lib["col_indices"])
    assert labels.tolist() == shared_lib["cc_labels"]


@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_pr", "test_bc_dir", "test_cc_dir",
                                  "bips98_606_dir"])
def test_fixtures(gr, orc, name):
    n, ro, co, _ = load_golden(name)
    want, ncomp = orc.cc(ro, co)
    got = gr.gunrock_cc_func(ro, co)
    assert np.array_equal(got, want)
    for sym in (False, name.endswith("_dir") is False):
        g = gr.Graph.from_csr(ro, co, symmetric=sym)
        labels, k, st = gr.cc(g)
        assert np.array_equal(labels, want) and k == ncomp
        assert st["gpu_launches"] == 3
        g.close()


def test_test_cc_fixture_has_two_components(gr):
    """dataset/small/test_cc.mtx: {1..7}, {8..11} (1-based)"""
    n, ro, co, _ = load_golden("test_cc")
    labels, k, _ = gr.cc(gr.Graph.from_csr(ro, co, symmetric=True))
    assert k == 2 and labels.tolist() == [0] * 7 + [7] * 4


@pytest.mark.parametrize("scale,ef", [(10, 16), (16, 16), (18, 2), (20, 16)])
def test_rmat(gr, orc, scale, ef):
    ro, co = rmat_graph(orc, scale, ef)
    want, ncomp = orc.cc(ro, co)
    for fob in (4, 8):
        g = gr.Graph.from_csr(ro, co, symmetric=True, force_offset_bytes=fob)
        labels, k, st = gr.cc(g)
        assert np.array_equal(labels, want) and k == ncomp
        assert st["edges_inspected"] == co.size
        g.close()


def test_directed_input_gives_weak_components(gr, orc):
    """every CSR entry is one hook edge whatever its direction (cc_problem.cuh:262-272)"""
    ro, co = rmat_graph(orc, 14, 4, undirected=False)
    want, ncomp = orc.cc(ro, co)
    labels, k, _ = gr.cc(gr.Graph.from_csr(ro, co))
    assert np.array_equal(labels, want) and k == ncomp


def test_edge_cases(gr, orc):
    for ro, co in [([0, 0], []), ([0, 0, 0, 0], []), ([0, 1, 2, 2], [1, 0])]:
        got = gr.gunrock_cc_func(ro, co)
        assert np.array_equal(got, orc.cc(ro, co)[0])
    # a long path: worst case for pointer chasing; and a star
    n = 200000
    rows = np.arange(n - 1)
    ro, co, _ = orc.coo_to_csr(n, np.concatenate([rows, rows + 1]), np.concatenate([rows + 1, rows]))
    labels, k, _ = gr.cc(gr.Graph.from_csr(ro, co, symmetric=True))
    assert k == 1 and not labels.any()
    # reversed ids: the smallest id sits at the far end of the path
    perm = np.arange(n)[::-1]
    ro2, co2, _ = orc.coo_to_csr(n, np.concatenate([perm[rows], perm[rows + 1]]), np.concatenate([perm[rows + 1], perm[rows]]))
    labels, k, _ = gr.cc(gr.Graph.from_csr(ro2, co2, symmetric=True))
    assert k == 1 and not labels.any()


def test_idempotent_across_runs(gr, orc):
    ro, co = rmat_graph(orc, 16)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    a = gr.cc(g)[0]
    for _ in range(3):
        assert np.array_equal(gr.cc(g)[0], a)
