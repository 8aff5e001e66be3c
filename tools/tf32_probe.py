"""Layout probe for the MN-major tf32 operand tiles: C = A . I with A[m][r] = m, then = r, shows which
(row, reduction index) the tensor core read at every logical position (diagnostic, GPU box only)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200")]
import alegpu

M, N, R = 128, 64, 32
eye = np.zeros((N, R), np.float32)
eye[np.arange(R), np.arange(R)] = 1.0          # B[n][r] = delta(n, r): C[m][n] = A(m, n) for n < R
Am = np.repeat(np.arange(M, dtype=np.float32)[:, None], R, 1)   # A(m, r) = m
Ar = np.repeat(np.arange(R, dtype=np.float32)[None, :], M, 0)   # A(m, r) = r
for name, a_mn, b_mn in (("A MN-major", True, False), ("B MN-major", False, True)):
    outs = []
    for A in (Am, Ar):
        if a_mn:
            got = alegpu.selftest_tc_gemm(A.T.copy(), eye, M, N, R, True, False, 64, tf32=True)[0]
        else:  # swap roles: C[m][n] = sum_r I(m,r) B(n,r) -> put the pattern in B (N=64 rows)
            Bp = A[:N]
            I2 = np.zeros((M, R), np.float32); I2[np.arange(R), np.arange(R)] = 1.0
            got = alegpu.selftest_tc_gemm(I2, Bp.T.copy(), M, N, R, False, True, 64, tf32=True)[0].T
        outs.append(got)
    gm, gr = outs
    rows = gm.shape[0]
    ok = np.array_equal(gm[:, :R], (Am if a_mn else Am[:N])[:, :R]) and np.array_equal(gr[:, :R], (Ar if a_mn else Ar[:N])[:, :R])
    print(name, "ok" if ok else "WRONG")
    if not ok:
        np.set_printoptions(linewidth=250, threshold=100000)
        print("read row index at logical (m, r), m = 0..15 and 32..39:")
        print(gm[list(range(16)) + list(range(32, 40)), :R].astype(int))
        print("read reduction index at logical (m, r):")
        print(gr[list(range(16)) + list(range(32, 40)), :R].astype(int))
