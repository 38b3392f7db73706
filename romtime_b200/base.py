"""Solution containers with the reference's attribute names (reference base.py:22-79).

``RomSolutionsStorage`` is what ``RomConstructor.solve`` leaves in ``.solutions``:
``ts`` (nt,), ``mu`` dict, ``domain`` (Nh x nt), ``fom`` (Nh x nt), ``rom`` (N x nt) -- NumPy arrays on the
host, deep-copied on construction like the reference.  ``RomBatchSolutions`` is the batched counterpart
returned by ``solve_batch``: the reduced trajectories stay on the GPU until asked for."""
import pickle
from copy import deepcopy

import numpy as np

from .conventions import PistonParameters


class SolutionsStorage:
    def __init__(self, ts, mu, domain, fom, snapshots=None):
        self.ts = np.array(ts, copy=True)
        self.mu = deepcopy(mu)
        self.snapshots = deepcopy(snapshots)
        self.fom = deepcopy(fom)
        self.domain = deepcopy(domain)

    def to_pickle(self, name):
        with open(name + ".pkl", mode="wb") as fp:
            pickle.dump(self, fp)

    def compute_at(self, x):
        """Probe the stored FOM fields at physical position(s) ``x`` for every time step (linear
        interpolation on the moving mesh, scaled to physical units by a0; reference base.py:45-67).
        The mesh runs from L down to 0, hence the flips."""
        cols = range(len(self.ts))
        pts = np.array([np.interp(x, self.domain[::-1, k], self.fom[::-1, k]) for k in cols])
        return pts * self.mu[PistonParameters.A0]


class RomSolutionsStorage(SolutionsStorage):
    def __init__(self, ts, mu, domain, fom, rom):
        super().__init__(ts=ts, mu=mu, domain=domain, fom=fom)
        self.rom = deepcopy(rom)


class RomBatchSolutions:
    """Result of a batched sweep: ``rom_device`` is a CUDA tensor (B, nt, N) -- [b, n, :] the reduced
    solution of parameter b after step n (a pinned host tensor after ``solve_batch(..., to_host=True)``); ``info`` (B,) int32, 0 = ok, n+1 = zero pivot at step n."""

    def __init__(self, ts, mus, rom_device, info, basis=None):
        self.ts = np.asarray(ts, dtype=float)
        self.mus = mus
        self.rom_device = rom_device
        self.info = info
        self._basis = basis

    def __len__(self):
        return self.rom_device.shape[0]

    def rom(self, b):
        """Reduced coefficients of parameter b as the reference stores them: (N, nt) NumPy array."""
        return self.rom_device[b].t().cpu().numpy()

    def fom(self, b):
        """Lazy FOM reconstruction V uN for parameter b (Nh x nt), reference rom/rom.py:496 -- one lift
        GEMM on the device (homogeneous part only; the lifting g_h is host/FEniCS data)."""
        from . import kernels

        if self._basis is None:
            raise ValueError("no reduced basis attached to this batch")
        return kernels.pod_lift(self._basis, kernels.to_device(self.rom_device[b]).t().contiguous(), order="C").cpu().numpy()

    @property
    def has_basis(self):
        return self._basis is not None

    def fom_batch(self, b0=0, b1=None, lifting=None):
        """Lazy FOM reconstruction of parameters b0 .. b1 - 1 as ONE lift GEMM on the device (reference rom/rom.py:496-522
        per step and parameter):  uc_h[b, n, :] = V uN[b, n, :] + g_h(mu_b, t_n).  ``lifting``: None (homogeneous part
        only) or an evaluator from romtime_b200/evaluators.py (``PistonP1`` / ``Mfp1P1``) whose boundary lifting is a
        combination of k fixed nodal shapes, g_h = coef(mu, t) @ shapes -- the coefficients ride along as k extra
        columns of the GEMM's left operand.  Returns a CUDA tensor (b1 - b0, nt, Nh)."""
        import torch

        from . import kernels

        if self._basis is None:
            raise ValueError("no reduced basis attached to this batch")
        b1 = len(self) if b1 is None else b1
        u = kernels.to_device(self.rom_device[b0:b1])
        nb, nt, N = u.shape
        W = kernels.to_device(np.ascontiguousarray(np.asarray(self._basis).T))            # N x Nh
        left = u.reshape(nb * nt, N)
        if lifting is not None:
            coef, shapes = lifting.lifting(self.mus[b0:b1], self.ts)                      # (nb, nt, k), (k, Nh)
            left = torch.cat([left, kernels.to_device(coef).reshape(nb * nt, -1)], dim=1).contiguous()
            W = torch.cat([W, kernels.to_device(shapes)], dim=0).contiguous()
        return kernels.pod_lift(left, W, order="C").reshape(nb, nt, -1)

    def storage(self, b, domain=None, lifting=None):
        """A reference-style RomSolutionsStorage for parameter b."""
        fom = self.fom(b) if self._basis is not None else None
        if fom is not None and lifting is not None:
            fom = fom + lifting
        mu = self.mus[b]
        return RomSolutionsStorage(ts=self.ts, mu=mu, domain=domain, fom=fom, rom=self.rom(b))
