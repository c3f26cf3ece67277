"""Generates tests/golden/*.npz from the CPU oracle (run from the repo root:
``python tests/golden/make_golden.py``).  Inputs are seeded numpy data; outputs are what
oracle/kmeans_oracle.py -- the restatement of the reference -- returns.  The reference itself
cannot be imported in the build container (dask is absent), see DESIGN.md."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from conftest import blobs  # noqa: E402
from oracle import kmeans_oracle as orc  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def lloyd_case(name, n, d, k, dtype, seed, max_iter, tol, chunks, init_mode="rows"):
    X, true = blobs(n, d, k, seed, dtype)
    if init_mode == "rows":
        init = X[:k].copy()
    else:
        init = (true + 0.5).astype(dtype)
    lab, inertia, C, n_iter = orc.kmeans_single_lloyd(X, k, init, max_iter=max_iter, tol=tol, chunks=chunks)
    np.savez_compressed(os.path.join(OUT, name), kind="lloyd", X=X, init=init, k=k, max_iter=max_iter, tol=tol,
                        chunks=np.array(chunks), labels=lab, centers=C, inertia=inertia, n_iter=n_iter)
    print(name, "n_iter", n_iter, "inertia", inertia)


def init_case(name, n, d, k, dtype, seed, l, chunks):
    X, _ = blobs(n, d, max(k // 4, 3), seed, dtype)
    trace = []
    need_rs = np.random.RandomState(seed)
    supp = {}

    def supplement(nn, need):
        supp["v"] = np.sort(need_rs.choice(nn, need, replace=False)).astype(np.int64)
        return supp["v"]

    C = orc.init_scalable(X, k, orc.PhiloxDraws(seed), None, l, "auto", chunks, supplement=supplement, trace=trace)
    picked = sorted(set(v for t in trace if "picked" in t for v in t["picked"]))
    np.savez_compressed(os.path.join(OUT, name), kind="init", X=X, k=k, seed=seed, l=l, chunks=np.array(chunks),
                        picked=np.array(picked, dtype=np.int64), supplement=supp.get("v", np.empty(0, np.int64)),
                        centers=C, cost0=trace[0]["cost0"], n_rounds=trace[0]["n_iter"])
    print(name, "rounds", trace[0]["n_iter"], "picked", len(picked), "centers", C.shape)


if __name__ == "__main__":
    lloyd_case("lloyd_f64_n2000_d16_k8.npz", 2000, 16, 8, np.float64, 0, 6, 0.0, (500, 500, 500, 500))
    lloyd_case("lloyd_f32_n3000_d32_k40.npz", 3000, 32, 40, np.float32, 1, 5, 0.0, (1000, 1000, 1000))
    lloyd_case("lloyd_f32_n2500_d20_k150_conv.npz", 2500, 20, 150, np.float32, 2, 100, 1e-4, (1250, 1250), "near")
    lloyd_case("lloyd_f64_n1500_d5_k3_conv.npz", 1500, 5, 3, np.float64, 3, 300, 1e-9, (1500,))
    init_case("init_f32_n4000_d8_k64.npz", 4000, 8, 64, np.float32, 5, 2, (2000, 2000))
    init_case("init_f64_n3000_d4_k6.npz", 3000, 4, 6, np.float64, 6, 2, (3000,))
