"""CPU restatement (numpy, plain loops) of the linear-algebra half of the path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module; the product never does.
Each function follows the reference routine it cites, loop for loop, so that results agree with the reference to
round-off; `tests/test_oracle_cpu.py` pins it against arrays dumped from the unmodified reference
(tests/golden/*.npz: Jacobian, ILU factors, inverse diagonals, right-hand side, solution, iteration count).

  block_spmv        CSysMatrix<T>::MatrixVectorProduct      Common/src/linear_algebra/CSysMatrix.cpp:617-645
  matrix_inverse    CSysMatrix<T>::MatrixInverse            CSysMatrix.cpp:548-601 (RegularizePivot :500-508)
  ilu0_build        CSysMatrix<T>::BuildILUPreconditioner   CSysMatrix.cpp:673-767 (one partition = serial order)
  ilu0_apply        CSysMatrix<T>::ComputeILUPreconditioner CSysMatrix.cpp:770-823
  fgmres            CSysSolve<T>::FGMRES_LinSolver          Common/src/linear_algebra/CSysSolve.cpp:418-613
                    ModGramSchmidt :179-240, Generate/ApplyGivens :132-162, SolveReduced :165-176
"""
import math
import numpy as np


def dia_ptr_of(row_ptr, col_ind):
    n = len(row_ptr) - 1
    dp = np.zeros(n, dtype=np.int64)
    for i in range(n):
        for k in range(row_ptr[i], row_ptr[i + 1]):
            if col_ind[k] == i:
                dp[i] = k
                break
        else:
            raise ValueError("row %d has no diagonal entry" % i)
    return dp


def block_spmv(row_ptr, col_ind, A, x, nPointDomain=None):
    """y_i = sum_k A_ik x_k over the owned rows; A [nnz][nVar][nVar], x [P][nVar]."""
    nv = A.shape[1]
    x = np.asarray(x, dtype=np.float64).reshape(-1, nv)
    n = len(row_ptr) - 1 if nPointDomain is None else nPointDomain
    y = np.zeros_like(x)
    for i in range(n):
        acc = np.zeros(nv)
        for k in range(row_ptr[i], row_ptr[i + 1]):
            acc += A[k] @ x[col_ind[k]]
        y[i] = acc
    return y


def _regularize(p):
    eps = float(np.float32(1e-12))
    if abs(p) < eps:
        return math.copysign(eps, p)
    return p


def matrix_inverse(a):
    """Gauss elimination without pivoting on a copy of `a`, applied to the identity, then back substitution."""
    a = np.array(a, dtype=np.float64)
    n = a.shape[0]
    m = np.eye(n)
    for i in range(1, n):
        for j in range(i):
            a[j, j] = _regularize(a[j, j])
            w = a[i, j] / a[j, j]
            for k in range(j, n):
                a[i, k] -= w * a[j, k]
            for k in range(j + 1):
                m[i, k] -= w * m[j, k]
    for i in range(n - 1, -1, -1):
        for j in range(i + 1, n):
            for k in range(n):
                m[i, k] -= a[i, j] * m[j, k]
        a[i, i] = _regularize(a[i, i])
        for k in range(n):
            m[i, k] /= a[i, i]
    return m


def ilu0_build(row_ptr, col_ind, dia_ptr, A, nPointDomain=None):
    """Returns (ILU [nnz][nVar][nVar], invM [Pd][nVar][nVar]); rows/columns >= Pd take no part."""
    n = len(row_ptr) - 1 if nPointDomain is None else nPointDomain
    ilu = np.array(A, dtype=np.float64)
    nv = ilu.shape[1]
    invM = np.zeros((n, nv, nv))
    pos = [dict((int(col_ind[k]), k) for k in range(row_ptr[i], row_ptr[i + 1])) for i in range(n)]
    for i in range(1, n):
        invM[i - 1] = matrix_inverse(ilu[dia_ptr[i - 1]])
        for idx in range(row_ptr[i], dia_ptr[i]):
            j = int(col_ind[idx])
            weight = ilu[idx] @ invM[j]
            for idx2 in range(dia_ptr[j] + 1, row_ptr[j + 1]):
                k = int(col_ind[idx2])
                if k >= n:
                    break
                p = pos[i].get(k)
                if p is not None:
                    ilu[p] = ilu[p] - weight @ ilu[idx2]
            ilu[idx] = weight
    if n > 0:
        invM[n - 1] = matrix_inverse(ilu[dia_ptr[n - 1]])
    return ilu, invM


def ilu0_apply(row_ptr, col_ind, dia_ptr, ilu, invM, x, nPointDomain=None):
    """y = U^-1 L^-1 x (forward with unit lower factor, backward with the stored inverse diagonal blocks)."""
    nv = ilu.shape[1]
    n = len(row_ptr) - 1 if nPointDomain is None else nPointDomain
    y = np.array(x, dtype=np.float64).reshape(-1, nv).copy()
    for i in range(1, n):
        for idx in range(row_ptr[i], dia_ptr[i]):
            y[i] -= ilu[idx] @ y[col_ind[idx]]
    for i in range(n - 1, -1, -1):
        aux = y[i].copy()
        for idx in range(dia_ptr[i] + 1, row_ptr[i + 1]):
            j = col_ind[idx]
            if j >= n:
                break
            aux -= ilu[idx] @ y[j]
        y[i] = invM[i] @ aux
    return y


def gauss_elimination(a, v):
    """Solve a x = v without pivoting (CSysMatrix::Gauss_Elimination, CSysMatrix.cpp:510-545): forward elimination of
    matrix and right-hand side together, regularised pivots, back substitution."""
    a = np.array(a, dtype=np.float64)
    v = np.array(v, dtype=np.float64)
    n = a.shape[0]
    for i in range(1, n):
        for j in range(i):
            a[j, j] = _regularize(a[j, j])
            w = a[i, j] / a[j, j]
            for k in range(j, n):
                a[i, k] -= w * a[j, k]
            v[i] -= w * v[j]
    for i in range(n - 1, -1, -1):
        for j in range(i + 1, n):
            v[i] -= a[i, j] * v[j]
        a[i, i] = _regularize(a[i, i])
        v[i] /= a[i, i]
    return v


def lusgs_apply(row_ptr, col_ind, dia_ptr, A, x, nPointDomain=None):
    """One symmetric Gauss-Seidel sweep, single partition (ComputeLU_SGSPreconditioner, CSysMatrix.cpp:826-888):
    (D+L) x* = b, then (D+U) x = D x*; every row solved with the diagonal block by Gauss elimination."""
    nv = A.shape[1]
    n = len(row_ptr) - 1 if nPointDomain is None else nPointDomain
    b = np.array(x, dtype=np.float64).reshape(-1, nv)
    y = np.zeros_like(b)
    for i in range(n):
        low = np.zeros(nv)
        for idx in range(row_ptr[i], dia_ptr[i]):
            low += A[idx] @ y[col_ind[idx]]
        y[i] = gauss_elimination(A[dia_ptr[i]], b[i] - low)
    for i in range(n - 1, -1, -1):
        dia = A[dia_ptr[i]] @ y[i]
        up = np.zeros(nv)
        for idx in range(dia_ptr[i] + 1, row_ptr[i + 1]):
            up += A[idx] @ y[col_ind[idx]]  # halo columns included (they hold the exchanged x*)
        y[i] = gauss_elimination(A[dia_ptr[i]], dia - up)
    return y


def _sign(x, y):
    return -abs(x) if y < 0.0 else abs(x)


def _generate_givens(dx, dy):
    if dx == 0.0 and dy == 0.0:
        return 0.0, 0.0, 0.0, 1.0
    if abs(dy) > abs(dx):
        tmp = dx / dy
        dx = abs(dy) * math.sqrt(1.0 + tmp * tmp)
        s = _sign(1.0 / math.sqrt(1.0 + tmp * tmp), dy)
        c = tmp * s
    elif abs(dy) <= abs(dx):
        tmp = dy / dx
        dx = abs(dx) * math.sqrt(1.0 + tmp * tmp)
        c = _sign(1.0 / math.sqrt(1.0 + tmp * tmp), dx)
        s = tmp * c
    else:
        dx, s, c = 0.0, 0.0, 1.0
    return dx, 0.0, s, c


def fgmres(matvec, precond, b, tol, m, nOwned=None):
    """Flexible GMRES(m) from x0 = 0 with the reference's stopping rule (|r| < tol*|b|). Returns (x, iterations,
    residual). `matvec`/`precond` map flat vectors to flat vectors; dots run over the first nOwned entries."""
    b = np.asarray(b, dtype=np.float64).ravel()
    no = b.size if nOwned is None else nOwned
    dot = lambda u, v: float(np.dot(u[:no], v[:no]))
    eps = 1e-14
    norm0 = math.sqrt(dot(b, b))
    x = np.zeros_like(b)
    V = [-1.0 * b]
    Z = []
    beta = norm0
    if beta < tol * norm0 or beta < eps:
        return x, 0, beta
    V[0] = V[0] / (-beta)
    g = np.zeros(m + 1)
    sn, cs = np.zeros(m + 1), np.zeros(m + 1)
    H = np.zeros((m + 1, m))
    g[0] = beta
    i = 0
    while i < m:
        if beta < tol * norm0:
            break
        Z.append(precond(V[i]))
        w = matvec(Z[i]).copy()
        nrm = dot(w, w)
        thr = nrm * 0.98
        if nrm <= 0.0 or nrm != nrm:
            raise RuntimeError("FGMRES orthogonalization failed, linear solver diverged.")
        for k in range(i + 1):
            prod = dot(w, V[k])
            h = prod
            w = w - prod * V[k]
            if prod * prod > thr:
                prod = dot(w, V[k])
                h += prod
                w = w - prod * V[k]
            H[k, i] = h
            nrm = max(nrm - h * h, 0.0)
            thr = nrm * 0.98
        nrm = math.sqrt(dot(w, w))
        H[i + 1, i] = nrm
        V.append(w / nrm)
        for k in range(i):
            h1, h2 = H[k, i], H[k + 1, i]
            H[k, i], H[k + 1, i] = cs[k] * h1 + sn[k] * h2, cs[k] * h2 - sn[k] * h1
        H[i, i], H[i + 1, i], sn[i], cs[i] = _generate_givens(H[i, i], H[i + 1, i])
        g1, g2 = g[i], g[i + 1]
        g[i], g[i + 1] = cs[i] * g1 + sn[i] * g2, cs[i] * g2 - sn[i] * g1
        beta = abs(g[i + 1])
        i += 1
    y = g[:i].copy()
    for k in range(i - 1, -1, -1):
        y[k] /= H[k, k]
        for j in range(k - 1, -1, -1):
            y[j] -= H[j, k] * y[k]
    for k in range(i):
        x = x + y[k] * Z[k]
    return x, i, beta / norm0
