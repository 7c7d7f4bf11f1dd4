"""Generates tests/golden/linalg_cv2.npz with OpenCV (cv2) as the authority for the LA primitives the
reference takes from OpenCV (cv::solve DECOMP_QR, cv::eigen, Mat::inv, gemm): SURVEY.md App. B.3.
Run in the build container (cv2 4.13.0): python tests/golden/make_linalg_golden.py"""
import os
import numpy as np
import cv2

rng = np.random.default_rng(20261104)
out = {"cv2_version": np.array(cv2.__version__)}
# 6x6 normal equations J^T J x = J^T b (laserOdometry.cpp:765-768)
A6, b6, x6 = [], [], []
for t in range(200):
    J = (rng.normal(size=(int(rng.integers(12, 400)), 6)) * rng.uniform(0.05, 20, size=6)).astype(np.float32)
    A = (J.T.astype(np.float64) @ J.astype(np.float64)).astype(np.float32)
    b = (J.T @ rng.normal(size=(J.shape[0], 1)).astype(np.float32)).astype(np.float32)
    ok, x = cv2.solve(A, b, flags=cv2.DECOMP_QR)
    A6.append(A); b6.append(b[:, 0]); x6.append(x[:, 0])
out["qr6_A"], out["qr6_b"], out["qr6_x"] = np.array(A6), np.array(b6), np.array(x6)
# 5x3 plane fit (laserMapping.cpp:828-832)
A5, x5 = [], []
for t in range(200):
    nrm = rng.normal(size=3); nrm /= np.linalg.norm(nrm)
    base = rng.normal(size=(1, 3)) * 30
    P = base + rng.normal(size=(5, 3)) * 0.4
    P -= ((P - base) @ nrm)[:, None] * nrm[None, :] * 0.97
    A = P.astype(np.float32)
    ok, x = cv2.solve(A, -np.ones((5, 1), np.float32), flags=cv2.DECOMP_QR)
    A5.append(A); x5.append(x[:, 0])
out["qr53_A"], out["qr53_x"] = np.array(A5), np.array(x5)
for n in (3, 6):
    As, Ws, Vs = [], [], []
    for t in range(200):
        J = (rng.normal(size=(int(rng.integers(n, 60)), n)) * rng.uniform(0.01, 10, size=n)).astype(np.float32)
        A = (J.T @ J).astype(np.float32)
        A = ((A + A.T) * np.float32(0.5)).astype(np.float32)
        ok, w, v = cv2.eigen(A)
        As.append(A); Ws.append(w[:, 0]); Vs.append(v)
    out[f"eig{n}_A"], out[f"eig{n}_w"], out[f"eig{n}_v"] = np.array(As), np.array(Ws), np.array(Vs)
Ai, Ii = [], []
for t in range(200):
    Q, _ = np.linalg.qr(rng.normal(size=(6, 6)))
    A = (Q + rng.normal(size=(6, 6)) * 0.05).astype(np.float32)
    ok, inv = cv2.invert(A, flags=cv2.DECOMP_LU)
    Ai.append(A); Ii.append(inv)
out["inv6_A"], out["inv6_inv"] = np.array(Ai), np.array(Ii)
Ag, Cg = [], []
for t in range(50):
    A = rng.normal(size=(6, 257)).astype(np.float32)
    Ag.append(A); Cg.append(cv2.gemm(A, A.T.copy(), 1, None, 0))
out["gemm_A"], out["gemm_C"] = np.array(Ag), np.array(Cg)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "linalg_cv2.npz"), **out)
print("written", {k: v.shape for k, v in out.items()})
