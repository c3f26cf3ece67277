"""Generates tests/golden/*.npz FROM THE REFERENCE (run in the build container, from the repo root:
``python tests/golden/make_golden.py``).

Inputs are seeded numpy data; every output is what the UNMODIFIED reference modules
(/root/reference/dask_ml/cluster/k_means.py, metrics/pairwise.py) return when run through
``oracle/reference_loader.py`` on the eager dask stand-in (``oracle/dask_stub``).  The reference tree does not
exist on the GPU box, so these files are how its answers travel there.  The oracle is NOT used to produce the
fixtures; it is only asserted equal to them afterwards (and again by tests/test_oracle_pin.py::test_golden_vectors).

k-means|| fixtures: the reference's ``init_scalable`` takes any object with dask's RandomState methods, so it is
driven here by the counter-based Philox stream the CUDA path uses (``oracle.philox_uniform_f64``; the published
Philox4x32-10 algorithm, pinned by Random123's known-answer vectors) -- the same draws on both sides make the
picked rows comparable bit for bit.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")

from conftest import blobs  # noqa: E402
from oracle import kmeans_oracle as orc  # noqa: E402
from oracle import reference_loader  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
ref = reference_loader.load()


def lloyd_case(name, n, d, k, dtype, seed, max_iter, tol, chunks, init_mode="rows"):
    X, true = blobs(n, d, k, seed, dtype)
    if init_mode == "rows":
        init = X[:k].copy()
    elif init_mode == "near":
        init = (true + 0.5).astype(dtype)
    else:   # "far": the last centre is far from every row, its cluster is empty after the first pass
        init = X[:k].copy()
        init[-1] = 1e3
    Xd = ref.da.from_array(X, chunks=(tuple(chunks), (d,)))
    lab, inertia, C, n_iter = ref.k_means._kmeans_single_lloyd(Xd, k, init=init, max_iter=max_iter, tol=tol)
    lab, inertia = lab.compute(), float(inertia.compute())
    o = orc.kmeans_single_lloyd(X, k, init, max_iter=max_iter, tol=tol, chunks=chunks)
    assert (o[0] == lab).all() and o[3] == n_iter and np.allclose(o[2], C, rtol=1e-12, atol=0) \
        and abs(o[1] - inertia) <= 1e-12 * abs(inertia), name
    np.savez_compressed(os.path.join(OUT, name), kind="lloyd", source="reference", X=X, init=init, k=k,
                        max_iter=max_iter, tol=tol, chunks=np.array(chunks), labels=lab, centers=C, inertia=inertia,
                        n_iter=n_iter)
    print(name, "n_iter", n_iter, "inertia", inertia)


class PhiloxForReference:
    """dask-RandomState-shaped provider of the CUDA path's draws, for the reference's init_scalable."""

    def __init__(self, seed, supplement_seed):
        self.seed, self.round = int(seed), 0
        self._supp_rs = np.random.RandomState(supplement_seed)
        self.supplement = np.empty(0, np.int64)

    def uniform(self, size, chunks):
        u = orc.philox_uniform_f64(self.seed, self.round, 0, size)
        self.round += 1
        return ref.da.from_array(u, chunks=chunks)

    def choice(self, a, size, replace, chunks):
        assert replace is False
        self.supplement = np.sort(self._supp_rs.choice(len(a), size, replace=False)).astype(np.int64)
        return ref.da.from_array(self.supplement, chunks=len(self.supplement))

    def randint(self, low, high, chunks=()):
        return ref.da.from_array(np.asarray(orc.PhiloxDraws(self.seed).randint(low, high)), chunks=())


def _row_ids(X, rows):
    """Indices of ``rows`` (exact copies of rows of X) in X."""
    key = {X[i].tobytes(): i for i in range(len(X) - 1, -1, -1)}
    return np.array([key[r.tobytes()] for r in np.ascontiguousarray(rows)], dtype=np.int64)


def init_case(name, n, d, k, dtype, seed, l, chunks):
    X, _ = blobs(n, d, max(k // 4, 3), seed, dtype)
    Xd = ref.da.from_array(X, chunks=(tuple(chunks), (d,)))
    rs = PhiloxForReference(seed, seed)
    C = ref.k_means.init_scalable(Xd, k, rs, None, l)
    supp = rs.supplement
    trace = []
    Co = orc.init_scalable(X, k, orc.PhiloxDraws(seed), None, l, "auto", chunks,
                           supplement=lambda nn, need: supp, trace=trace)
    picked = np.array(sorted(set(v for t in trace if "picked" in t for v in t["picked"])), dtype=np.int64)
    assert np.allclose(C, Co, rtol=1e-12, atol=0) and trace[0]["n_iter"] == rs.round, name
    if len(supp):   # candidates + supplement are exact rows of X: recover the reference's own picks from them
        ids = _row_ids(X, C[: len(C) - len(supp)])
        assert sorted(set(ids) - {0}) == sorted(set(picked) - {0}), name
    np.savez_compressed(os.path.join(OUT, name), kind="init", source="reference", X=X, k=k, seed=seed, l=l,
                        chunks=np.array(chunks), picked=picked, supplement=supp, centers=C,
                        cost0=float(ref.dask.compute(ref.k_means.evaluate_cost(Xd, X[0, np.newaxis]))[0]),
                        n_rounds=rs.round)
    print(name, "rounds", rs.round, "picked", len(picked), "supplement", len(supp), "centers", C.shape)


if __name__ == "__main__":
    for old in os.listdir(OUT):
        if old.endswith(".npz"):
            os.remove(os.path.join(OUT, old))
    # Lloyd, shapes after BASELINE.json's configs (C1 d=16 k=8 f64 | C2 d=32 k=256 f32 | C3 d=64 k=1024 f32 | C4 d=128)
    lloyd_case("lloyd_f64_n2000_d16_k8.npz", 2000, 16, 8, np.float64, 0, 6, 0.0, (500, 500, 500, 500))
    lloyd_case("lloyd_f32_n3000_d32_k40.npz", 3000, 32, 40, np.float32, 1, 5, 0.0, (1000, 1000, 1000))
    lloyd_case("lloyd_f32_n2500_d20_k150_conv.npz", 2500, 20, 150, np.float32, 2, 100, 1e-4, (1250, 1250), "near")
    lloyd_case("lloyd_f64_n1500_d5_k3_conv.npz", 1500, 5, 3, np.float64, 3, 300, 1e-9, (1500,))
    lloyd_case("lloyd_f32_n4000_d32_k256_c2.npz", 4000, 32, 256, np.float32, 4, 3, 0.0, (1000,) * 4)
    lloyd_case("lloyd_f32_n3000_d64_k1024_c3.npz", 3000, 64, 1024, np.float32, 5, 2, 0.0, (800, 800, 800, 600))
    lloyd_case("lloyd_f32_n2000_d128_k1500_c4.npz", 2000, 128, 1500, np.float32, 6, 2, 0.0, (1000, 1000))
    lloyd_case("lloyd_f64_n2500_d64_k300.npz", 2500, 64, 300, np.float64, 7, 3, 0.0, (900, 900, 700))
    lloyd_case("lloyd_f32_n2000_d8_k6_empty.npz", 2000, 8, 6, np.float32, 8, 1, 0.0, (500,) * 4, "far")
    lloyd_case("lloyd_f32_n2500_d20_k40_tolbreak.npz", 2500, 20, 40, np.float32, 9, 100, 5.0, (900, 900, 700))   # exit 4 with 1e-7 < shift < tol
    # k-means||: the supplement branch (< k candidates) and the re-cluster branch (>= k)
    init_case("init_f32_n4000_d8_k64.npz", 4000, 8, 64, np.float32, 5, 2, (2000, 2000))
    init_case("init_f64_n3000_d4_k6.npz", 3000, 4, 6, np.float64, 6, 2, (3000,))
    init_case("init_f64_n3000_d16_k40.npz", 3000, 16, 40, np.float64, 7, 2, (1000, 1000, 1000))
    init_case("init_f32_n2500_d6_k5.npz", 2500, 6, 5, np.float32, 8, 4, (500,) * 5)
