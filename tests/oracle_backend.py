"""TEST INFRASTRUCTURE: the CPU oracle behind the kernel interface of
cedar_b200.mpi2.DistSolver2, so that the host logic of the block-decomposed solver
(partition, ghost-exchange schedule, agglomeration, reduction) can be tested on CPU
with the gloo backend.  Never used by the product: DistSolver2 builds CudaKernels
unless a test passes this object explicitly."""
import numpy as np
import torch

import oracle
from oracle import P


def _n(t):
    a = t.numpy()
    assert a.flags["C_CONTIGUOUS"]
    return a


class OracleKernels:
    def __init__(self):
        self.L = oracle.lib()
        self.launches = 0

    def recip(self, A, sor):
        nst, jj, ii = A.shape
        self.L.orc2_setup_recip(P(_n(A)), P(_n(sor)), ii, jj, nst, 2)

    def relax_pass(self, A, b, x, sor, updown, p):
        nst, jj, ii = A.shape
        assert self.L.orc2_relax_gs_pass(1, P(_n(A)), P(_n(b)), P(_n(x)), P(_n(sor)), ii, jj, 1, 1 if nst == 3 else 0,
                                         nst, 2, 1, updown, 0, p) == 0

    def relax_oop(self, A, b, x, xout, sor, updown):
        nst, jj, ii = A.shape
        xout.copy_(x)
        assert self.L.orc2_relax_gs(1, P(_n(A)), P(_n(b)), P(_n(xout)), P(_n(sor)), ii, jj, 1, 1 if nst == 3 else 0, nst,
                                    2, 1, updown, 0) == 0

    def residual(self, A, b, x, r):
        nst, jj, ii = A.shape
        self.L.orc2_residual(0, P(_n(A)), P(_n(b)), P(_n(x)), P(_n(r)), ii, jj, 0, 1 if nst == 3 else 0, nst)

    def restrict(self, r, bc, Pw):
        jj, ii = r.shape
        jjc, iic = bc.shape
        assert self.L.orc2_restrict(P(_n(r)), P(_n(bc)), P(_n(Pw)), ii, jj, iic, jjc, 0) == 0

    def interp_add(self, x, xc, r, A, Pw):
        nst, jj, ii = A.shape
        jjc, iic = xc.shape
        assert self.L.orc2_interp_add(P(_n(x)), P(_n(xc)), P(_n(r)), P(_n(A)), P(_n(Pw)), iic, jjc, ii, jj, nst, 0) == 0

    def setup_interp_stage(self, A, Pw, stage):
        nst, jj, ii = A.shape
        _, jjc, iic = Pw.shape
        assert self.L.orc2_setup_interp_oi_stage(P(_n(A)), None, P(_n(Pw)), ii, jj, iic, jjc, 1 if nst == 3 else 0,
                                                 nst, 0, 0, stage) == 0

    def galerkin(self, A, Ac, Pw):
        nst, jj, ii = A.shape
        _, jjc, iic = Pw.shape
        assert self.L.orc2_setup_itli_ex(P(_n(A)), P(_n(Ac)), P(_n(Pw)), ii, jj, iic, jjc, 1 if nst == 3 else 0, nst,
                                         0) == 0

    def line_copy(self, a, mask, direction, index, buf, to_buf):
        a3 = a if a.dim() == 3 else a.unsqueeze(0)
        sel = [p for p in range(a3.shape[0]) if mask & (1 << p)]
        line = a3[sel, :, index - 1] if direction == 0 else a3[sel, index - 1, :]
        if to_buf:
            buf.copy_(line.reshape(-1))
        else:
            line_view = buf.reshape(len(sel), -1)
            if direction == 0:
                a3[sel, :, index - 1] = line_view
            else:
                a3[sel, index - 1, :] = line_view

    def sumsq(self, r, out):
        out[0] = float((r[1:-1, 1:-1].double() ** 2).sum())


class OracleCoarse:
    """agglomerated levels on the CPU oracle (mirror of mpi2._SerialCoarse)"""

    def __init__(self, g, npre, npost, nlevels):
        self.so = np.ascontiguousarray(g.numpy())
        self.sol = oracle.Solver2(self.so, "point", npre, npost, num_levels=nlevels)
        assert self.sol.nlevels == nlevels

    def vcycle(self, x, b):
        self.sol.vcycle(_n(x), _n(b))
