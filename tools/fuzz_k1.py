"""One-off stress run: many random (n, d, k, variant, offset, scale, stride) cases through K1 + re-decision against
float64 numpy.  Labels must match outside the float64 expanded form's own rounding; winner distances within the float32
accuracy model written at the check.  Where the Lloyd iteration would take the labels-only form (k <= 256, d <= 32:
b200km_assign_lloyd_f32) that form must give the same labels.  argv: trials seed [small]
("small": only shapes of the small-k tensor-core kernel)"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dask_ml_b200._engine import get_engine
eng = get_engine(); lib = eng.lib
trials = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rs = np.random.RandomState(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
small = len(sys.argv) > 3 and sys.argv[3] == "small"
for t in range(trials):
    d = int(rs.choice([1, 2, 3, 4, 7, 8, 13, 16, 24, 31, 32] if small else
                      [1, 2, 3, 4, 7, 8, 13, 16, 24, 31, 32, 33, 48, 64, 65, 100, 128, 129, 160]))
    k = int(rs.choice([17, 31, 32, 33, 64, 100, 129, 200, 255, 256] if small else
                      [1, 2, 5, 16, 17, 31, 64, 100, 255, 256, 257, 300, 512, 513, 1024, 1500]))
    n = int(rs.choice([1, 2, 31, 127, 128, 129, 1000, 4095, 4096, 4097, 10007, 40000, 150001]))
    variant = int(rs.choice([0, 0, 0, 1, 2, 3]))
    offset = float(rs.choice([0.0, 0.0, 5.0, 300.0, -2000.0]))
    scale = float(rs.choice([1e-3, 1.0, 1.0, 50.0]))
    pad = int(rs.choice([0, 0, 1, 5]))
    dup = rs.rand() < 0.2
    kk = max(1, k // 4)
    cen = rs.uniform(-10, 10, size=(kk, d))
    Xw = ((cen[rs.randint(0, kk, n)] + rs.standard_normal((n, d))) * scale + offset).astype(np.float32)
    if pad:
        wide = np.zeros((n, d + pad), np.float32); wide[:, :d] = Xw
        Xd = eng.to_device(wide)[:, :d]
    else:
        Xd = eng.to_device(Xw)
    C = (Xw[rs.randint(0, n, k)].astype(np.float64) + 0.2 * scale * rs.standard_normal((k, d)))
    if dup and k >= 4:
        C[k // 2:k // 2 + min(9, k // 2)] = C[0]
    X64 = Xw.astype(np.float64)
    # exact float64 reference in DIRECT form (the expanded form loses ~1e-16 * (|x|^2 + |c|^2), which matters for
    # data far from the origin), in row blocks
    D = np.empty((n, k))
    for r0 in range(0, n, 2048):
        blk = X64[r0:r0 + 2048]
        D[r0:r0 + 2048] = ((blk[:, None, :] - C[None, :, :]) ** 2).sum(-1) if blk.shape[0] * k * d <= 4e7 else \
            np.stack([((blk - C[j]) ** 2).sum(1) for j in range(k)], axis=1)
    lab_ref = D.argmin(1)
    lib.b200km_set_option(1, variant)
    try:
        pack = eng.pack(eng.to_device(C, torch.float64))
        lab, m2 = eng.assign(Xd, pack)
        lab_l = None
        if eng.lloyd_split(Xd, k, d):
            scratch = torch.full((n,), float("nan"), dtype=torch.float32, device=eng.device)
            lab_l, _ = eng.assign(Xd, pack, mind2=scratch, lloyd=True)
        torch.cuda.synchronize()
    finally:
        lib.b200km_set_option(1, 0)
    lab = lab.cpu().numpy(); m2 = m2.double().cpu().numpy()
    diff = lab != lab_ref
    rows_i = np.arange(n)
    xn = (X64 ** 2).sum(1); cn = (C ** 2).sum(1)
    # a different label is legitimate only inside the float64 EXPANDED form's own rounding (what sklearn computes)
    dd = np.abs(D[rows_i, lab] - D[rows_i, lab_ref])
    hard = diff & (dd > 1e-9 * D[rows_i, lab_ref] + 64 * 2.2e-16 * (xn + cn[lab_ref]))
    m_at = D[rows_i, lab]                       # exact distance to the centre WE chose
    # float32 accuracy model of the winner distance: the row and the centre are rounded once each after centring on
    # mu = mean of the centres, then sum (x' - c')^2 in fp32:  |err| <= 3e-6 m + 8 * 2^-24 * sqrt(m) * (|x'| + |c'|)
    mu = C.mean(0)
    xc = np.sqrt(((X64 - mu) ** 2).sum(1)); cc = np.sqrt(((C - mu) ** 2).sum(1))[lab]
    allowed = 3e-6 * m_at + 8 * 2.0 ** -24 * np.sqrt(m_at) * (xc + cc) + 1e-37
    err = np.abs(m2 - m_at) / allowed
    same = lab_l is None or bool((lab_l == torch.as_tensor(lab, device=lab_l.device)).all().item())
    ok = (not hard.any()) and err.max() < 1.0 and same
    if not ok:
        bad += 1
        print(f"FAIL trial {t}: n={n} d={d} k={k} variant={variant} offset={offset} scale={scale} pad={pad} dup={dup}: "
              f"{int(hard.sum())} hard label mismatches, mind2 error / allowed {err.max():.2e}, "
              f"labels-only form equal: {same}", flush=True)
print(f"{trials} trials, {bad} failures")
