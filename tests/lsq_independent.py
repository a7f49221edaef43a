"""An INDEPENDENT restatement of the reference's non-minimal least-squares fit for the tests (TEST INFRASTRUCTURE): rows as
estimateNonMinimalModel builds them with unit weights (solver_rectifying_homography_two_sift.hpp:228-259,423-523: one row
[x, y, s^(1/3)] = 1 per scale feature, one row [vp0, vp1, 0] = vp2 per PAIR of orientation features, vp = l_i x l_j scaled by its largest
magnitude when that exceeds 1; three_sift.hpp:195-254 for the one-class problem), solved by LAPACK through scipy's column-pivoted QR with
Eigen's ColPivHouseholderQR conventions (rank = pivots above eps * 3 * largest pivot; a rank-deficient system gets the BASIC solution: free
components zero).  The golden vectors under tests/golden/ come from the reference running on the builder-written QR of oracle/shim/Eigen;
this module pins them, the host fit and the device fit against a solver nobody here wrote."""
import numpy as np
import scipy.linalg


def line_from_point_and_angle(x, y, t):  # math_utils.hpp:104-109
    c, s = np.cos(t), np.sin(t)
    return np.array([s, -c, y * c - x * s])


def rows(scale, orient):
    """Coefficient matrix [n_rows,3] and right-hand side of the hybrid non-minimal fit (unit weights), in the reference's row order."""
    A, b = [], []
    for x, y, s in scale:
        A.append([x, y, s ** (1.0 / 3.0)])
        b.append(1.0)
    lines = [line_from_point_and_angle(*r) for r in orient]
    for i in range(len(lines) - 1):
        for j in range(i + 1, len(lines)):
            vp = np.cross(lines[i], lines[j])
            m = np.abs(vp).max()
            if m > 1.0:
                vp = vp / m
            A.append([vp[0], vp[1], 0.0])
            b.append(vp[2])
    return np.array(A, float).reshape(-1, 3), np.array(b, float)


def solve_colpiv(A, b):
    """x = A.colPivHouseholderQr().solve(b) by LAPACK's pivoted QR: (solution, rank)."""
    Q, R, P = scipy.linalg.qr(A, mode="economic", pivoting=True)
    d = np.abs(np.diag(R))
    rank = int((d > np.finfo(float).eps * min(A.shape) * d.max()).sum()) if d.max() > 0 else 0
    c = Q.T @ b
    y = np.zeros(A.shape[1])
    if rank:
        y[:rank] = scipy.linalg.solve_triangular(R[:rank, :rank], c[:rank])
    x = np.zeros(A.shape[1])
    x[P] = y
    return x, rank


def fit(scale, orient):
    """(h7, h8, alpha) or None, with the reference's acceptance rules (>= 2 scale rows, >= 1 pair row, finite, alpha >= 1e-9)."""
    if len(scale) < 2 or len(orient) < 2:
        return None
    A, b = rows(scale, orient)
    x, _ = solve_colpiv(A, b)
    if not np.all(np.isfinite(x)) or x[2] < 1e-9:
        return None
    return x
