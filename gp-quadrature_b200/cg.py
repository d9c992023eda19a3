"""Conjugate gradients with the reference's interface (cg.py:5-261 of danbider/gp-quadrature).

`ConjugateGradients(A_apply_function, b, x0, tol, max_iter, early_stopping, M_inv_apply).solve()`
and the `iters_completed` attribute behave as in the reference, for single (n,) and batched (B, n)
right-hand sides with per-system early stopping.

When the operator is one of this package's fused Toeplitz operators (efgpnd.create_A_mean /
create_A_var / create_Gv) and the preconditioner is absent or the Jacobi object from
efgpnd.create_jacobi_precond, the whole solve runs inside libefgp_b200 (lockstep batched CG,
csrc/cg.cu).  Any other callable or dense tensor goes through the generic loop below, which uses
whatever device the tensors live on and only torch elementwise ops.
"""
import torch


class ConjugateGradients:
    def __init__(self, A_apply_function, b, x0, tol=1e-6, max_iter=None, early_stopping=True, M_inv_apply=None):
        self.device = b.device
        if isinstance(A_apply_function, torch.Tensor):
            mat = A_apply_function
            self.A_apply_function = lambda v: mat @ v
            self.dtype = mat.dtype
        elif callable(A_apply_function):
            self.A_apply_function = A_apply_function
            self.dtype = x0.dtype
        else:
            raise ValueError("A_apply_function must be a torch.Tensor or a callable")
        self.b = b.to(dtype=self.dtype, device=self.device)
        self.x0 = x0.to(dtype=self.dtype, device=self.device)
        # efgpnd tags the cold-start guesses it creates with torch.zeros_like: the library then skips A x0
        self._x0_zero = bool(getattr(x0, "_efgp_zero", False))
        self.is_batched = b.dim() > 1
        self.tol = tol
        self.div_eps = 1e-16
        if max_iter is None:
            max_iter = 2 * (self.b.shape[1] if self.is_batched else len(self.b))
        self.max_iter = max_iter
        self.early_stopping = early_stopping
        self.M_inv_apply = M_inv_apply
        self._iters = 0

    @property
    def iters_completed(self):
        """Iterations of the last solve (cg.py:152, 243).  A solve that the library only enqueued (one cooperative
        kernel, no host decisions) reports it lazily: reading the attribute waits for that solve."""
        if not isinstance(self._iters, int):
            self._iters = int(self._iters)
        return self._iters

    @iters_completed.setter
    def iters_completed(self, value):
        self._iters = value

    # ------------------------------------------------------------------ dispatch
    def _fused(self):
        op = self.A_apply_function
        if not getattr(op, "is_efgp_fused_operator", False) or not self.b.is_cuda:
            return None
        pre = self.M_inv_apply
        if pre is not None and not getattr(pre, "is_efgp_jacobi", False):
            return None
        if self.b.dim() > 2 or not self.b.is_complex():
            return None
        if not getattr(op, "_real_ws", False):
            return None      # complex weights: the library solve would drop Im(ws); use the operator's own __call__
        return op

    def solve(self):
        op = self._fused()
        if op is not None:
            minv = self.M_inv_apply.inv_diag if self.M_inv_apply is not None else None
            x, it = op.toeplitz.plan.cg(self.b, self.x0, op.mode, op.ws_real, op.sigmasq, self.tol,
                                        max_iter=self.max_iter, early_stopping=self.early_stopping,
                                        minv=minv, batched=self.is_batched,
                                        x0_zero=self._x0_zero,
                                        hermitian=bool(getattr(self.b, "_efgp_hermitian", False)), lazy_iters=True)
            self._iters = it                    # a LazyIters: resolved when iters_completed is read
            return x
        return self._solve_batched() if self.is_batched else self._solve_single()

    # ------------------------------------------------------------------ generic paths
    def _precond(self, r):
        return self.M_inv_apply(r) if self.M_inv_apply is not None else r

    def _solve_single(self):
        """One right-hand side; the stop test runs right after the residual update (cg.py:116-153)."""
        with torch.no_grad():
            A = self.A_apply_function
            x = self.x0.clone()
            r = self.b - A(x)
            z = self._precond(r)
            p = z.clone()
            rz = self._inner_product(r, z).real
            bnorm = torch.linalg.norm(self.b).real
            scale = bnorm if bnorm > 0 else torch.ones((), device=self.device, dtype=bnorm.dtype)
            k = -1
            for k in range(self.max_iter):
                Ap = A(p)
                step = rz / (self._inner_product(p, Ap).real + self.div_eps)
                x = x + step * p
                r = r - step * Ap
                if self.early_stopping and torch.linalg.norm(r).real / (scale + self.div_eps) < self.tol:
                    break
                z = self._precond(r)
                rz_next = self._inner_product(r, z).real
                p = z + (rz_next / (rz + self.div_eps)) * p
                rz = rz_next
            self.iters_completed = k + 1
            return x

    def _solve_batched(self):
        """(B, n) systems; rows leave the active set once rel. residual < tol or ||r|| < 1e-12 (cg.py:193-244)."""
        with torch.no_grad():
            A = self.A_apply_function
            x = self.x0.clone()
            r = self.b - A(x)
            z = self._precond(r)
            p = z.clone()
            rz = torch.sum(r.conj() * z, dim=1).real
            bnorm = torch.linalg.norm(self.b, dim=1).real
            scale = torch.where(bnorm > 0, bnorm, torch.ones_like(bnorm))
            live = torch.ones(x.shape[0], dtype=torch.bool, device=self.device)
            k = -1
            for k in range(self.max_iter):
                rows = torch.nonzero(live, as_tuple=False).flatten()
                if rows.numel() == 0:
                    break
                pk = p[rows]
                Ap = A(pk)
                step = rz[rows] / (torch.sum(pk.conj() * Ap, dim=1).real + self.div_eps)
                x[rows] += step.unsqueeze(1) * pk
                rk = r[rows] - step.unsqueeze(1) * Ap
                r[rows] = rk
                zk = self._precond(rk)
                rz_next = torch.sum(rk.conj() * zk, dim=1).real
                p[rows] = zk + (rz_next / (rz[rows] + self.div_eps)).unsqueeze(1) * pk
                rz[rows] = rz_next
                if self.early_stopping:
                    rn = torch.linalg.norm(rk, dim=1).real
                    done = (rn / (scale[rows] + self.div_eps) < self.tol) | (rn < 1e-12)
                    live[rows[done]] = False
            self.iters_completed = k + 1
            return x

    def _inner_product(self, a, b):
        if a.dim() == 1:
            return torch.dot(torch.conj(a), b)
        if a.dim() == 2:
            return torch.sum(a.conj() * b, dim=1)
        raise ValueError(f"Unsupported tensor dimension: {a.dim()}")
