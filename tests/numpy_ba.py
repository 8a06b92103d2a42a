"""Independent numpy fp64 restatement of the BA path (second check): written from
the formulas, not from oracle/ba_oracle.c, so that the C oracle and the CUDA code cannot
share one misreading.  Dense (no Schur): solves the full (6Nc+3Np) damped normal equations.

Reference conventions: projection include/keyframe.h:82-89; 2x6 Jacobian src/viso.cpp:640-658;
pose update exp(d)*T include/bundle_adjuster.h:51-55; point update :34-38; LM = g2o policy.
"""
import numpy as np


def quat_to_R(q):
    x, y, z, w = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def hat(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0.0]])


def se3_exp(xi):
    """(R, t) of exp((v, w)) by the closed form (Rodrigues + V matrix)."""
    v, w = np.asarray(xi[:3], float), np.asarray(xi[3:], float)
    th = np.linalg.norm(w)
    W = hat(w)
    if th < 1e-6:
        R = np.eye(3) + W + 0.5 * W @ W
        V = np.eye(3) + 0.5 * W + W @ W / 6.0
    else:
        R = np.eye(3) + np.sin(th) / th * W + (1 - np.cos(th)) / th**2 * W @ W
        V = np.eye(3) + (1 - np.cos(th)) / th**2 * W + (th - np.sin(th)) / th**3 * W @ W
    return R, V @ v


def R_to_quat(R):
    """unit quaternion (x, y, z, w), w >= 0 branch-free enough for tests"""
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        q = np.array([(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s])
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k]) * 2
        q = np.zeros(4)
        q[i] = 0.25 * s
        q[j] = (R[j, i] + R[i, j]) / s
        q[k] = (R[k, i] + R[i, k]) / s
        q[3] = (R[k, j] - R[j, k]) / s
    return q / np.linalg.norm(q)


def project(K, R, t, P):
    fx, fy, cx, cy = K
    Pc = R @ P + t
    return np.array([fx * Pc[0] / Pc[2] + cx, fy * Pc[1] / Pc[2] + cy]), Pc


def jac_pose(K, Pc):
    fx, fy = K[0], K[1]
    x, y, z = Pc
    return np.array([
        [fx / z, 0, -fx * x / z**2, -fx * x * y / z**2, fx + fx * x * x / z**2, -fx * y / z],
        [0, fy / z, -fy * y / z**2, -fy - fy * y * y / z**2, fy * x * y / z**2, fy * x / z]])


def jac_point(K, Pc, R):
    fx, fy = K[0], K[1]
    x, y, z = Pc
    Jp = np.array([[fx / z, 0, -fx * x / z**2], [0, fy / z, -fy * y / z**2]])
    return Jp @ R


def huber(s, delta):
    if delta <= 0 or s <= delta * delta:
        return s, 1.0
    r = np.sqrt(s)
    return 2 * r * delta - delta * delta, delta / r


class Problem:
    def __init__(self, K, poses7, points3, edge_pose, edge_point, edge_uv, huber_delta=0.0, fixed=(0,)):
        self.K = K
        self.R = [quat_to_R(p[:4] / np.linalg.norm(p[:4])) for p in poses7]
        self.t = [np.array(p[4:7], float) for p in poses7]
        self.P = [np.array(p, float) for p in points3]
        self.ep, self.el = list(map(int, edge_pose)), list(map(int, edge_point))
        self.uv = np.asarray(edge_uv, np.float32).astype(np.float64).reshape(-1, 2)
        self.delta = huber_delta
        self.free = [c for c in range(len(self.R)) if c not in set(fixed)]
        self.slot = {c: i for i, c in enumerate(self.free)}

    def cost(self, R=None, t=None, P=None):
        R, t, P = R or self.R, t or self.t, P or self.P
        tot = 0.0
        for e in range(len(self.ep)):
            uv, _ = project(self.K, R[self.ep[e]], t[self.ep[e]], P[self.el[e]])
            r = uv - self.uv[e]
            tot += huber(r @ r, self.delta)[0]
        return tot

    def normal_equations(self):
        """dense H (undamped) and b = -J^T W r over [free poses (6 each), points (3 each)]"""
        nc, npt = len(self.free), len(self.P)
        n = 6 * nc + 3 * npt
        H, b = np.zeros((n, n)), np.zeros(n)
        for e in range(len(self.ep)):
            c, l = self.ep[e], self.el[e]
            uv, Pc = project(self.K, self.R[c], self.t[c], self.P[l])
            r = uv - self.uv[e]
            _, w = huber(r @ r, self.delta)
            Jl = jac_point(self.K, Pc, self.R[c])
            il = 6 * nc + 3 * l
            H[il:il + 3, il:il + 3] += w * Jl.T @ Jl
            b[il:il + 3] -= w * Jl.T @ r
            if c in self.slot:
                Jp = jac_pose(self.K, Pc)
                ic = 6 * self.slot[c]
                H[ic:ic + 6, ic:ic + 6] += w * Jp.T @ Jp
                H[ic:ic + 6, il:il + 3] += w * Jp.T @ Jl
                H[il:il + 3, ic:ic + 6] += w * Jl.T @ Jp
                b[ic:ic + 6] -= w * Jp.T @ r
        return H, b

    def reduced_system(self, lam):
        nc = len(self.free)
        H, b = self.normal_equations()
        H = H + lam * np.eye(len(b))
        m = 6 * nc
        Hpp, Hpl, Hll = H[:m, :m], H[:m, m:], H[m:, m:]
        Hll_inv = np.linalg.inv(Hll)  # block diagonal, tiny problems only
        return Hpp - Hpl @ Hll_inv @ Hpl.T, b[:m] - Hpl @ Hll_inv @ b[m:]

    def step(self, lam):
        H, b = self.normal_equations()
        x = np.linalg.solve(H + lam * np.eye(len(b)), b)
        return x, H, b

    def apply(self, x):
        nc = len(self.free)
        R, t, P = list(self.R), list(self.t), list(self.P)
        for i, c in enumerate(self.free):
            dR, dt = se3_exp(x[6 * i:6 * i + 6])
            R[c], t[c] = dR @ self.R[c], dR @ self.t[c] + dt
        for l in range(len(P)):
            P[l] = self.P[l] + x[6 * nc + 3 * l:6 * nc + 3 * l + 3]
        return R, t, P

    def lm(self, iterations, tau=1e-5, max_trials=10):
        """g2o OptimizationAlgorithmLevenberg policy; returns per-iteration (cost, lambda, trials)."""
        chi = self.cost()
        lam, nu, hist = 0.0, 2.0, []
        for it in range(iterations):
            H, b = self.normal_equations()
            if it == 0:
                lam = tau * np.max(np.abs(np.diag(H)))
            q, rho = 0, 0.0
            while True:
                x = np.linalg.solve(H + lam * np.eye(len(b)), b)
                R, t, P = self.apply(x)
                chi_new = self.cost(R, t, P)
                scale = x @ (lam * x + b) + 1e-3
                rho = (chi - chi_new) / scale
                if rho > 0 and np.isfinite(chi_new):
                    lam *= max(1 / 3, min(2 / 3, 1 - (2 * rho - 1) ** 3))
                    nu = 2.0
                    chi, self.R, self.t, self.P = chi_new, R, t, P
                else:
                    lam *= nu
                    nu *= 2
                q += 1
                if not (rho < 0 and q < max_trials):
                    break
            hist.append((chi, lam, q))
            if q == max_trials or rho == 0:
                break
        return hist

    def poses7(self):
        return np.array([np.concatenate([R_to_quat(R), t]) for R, t in zip(self.R, self.t)])

    def points3(self):
        return np.array(self.P)
