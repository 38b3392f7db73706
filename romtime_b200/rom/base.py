"""``Reductor``: bookkeeping shared by the ROM constructor and the (M)DEIM reductors
(reference rom/base.py:9-163): parameter spaces per stage, the offline report, RMS errors."""
from collections import defaultdict

import numpy as np

from ..conventions import ProblemType, Stage, Treewalk, TreewalkNonlinear


class Reductor:
    FOM = ProblemType.FOM
    ROM = ProblemType.ROM

    BASIS_AFTER_WALK = Treewalk.BASIS_AFTER_WALK
    BASIS_FINAL = Treewalk.BASIS_FINAL
    BASIS_TIME = Treewalk.BASIS_TIME
    ENERGY_MU = Treewalk.ENERGY_MU
    ENERGY_TIME = Treewalk.ENERGY_TIME
    SPECTRUM_MU = Treewalk.SPECTRUM_MU
    SPECTRUM_TIME = Treewalk.SPECTRUM_TIME

    def __init__(self, grid=None):
        self.grid = grid
        self.mu_space = {stage: [] for stage in (Stage.OFFLINE, Stage.ONLINE, Stage.VALIDATION)}
        self.report = defaultdict(dict)
        self.errors_rom = defaultdict(list)
        self.summary_errors = None
        self.mu = None
        self.random_state = None

    @staticmethod
    def _compute_error(u, ue):
        """RMS difference ||u - ue||_2 / sqrt(len(u))  (reference rom/base.py:52-73)."""
        u = np.asarray(u)
        return np.linalg.norm(u - np.asarray(ue), ord=2) / np.sqrt(len(u))

    def add_mu(self, step, mu):
        """Record ``mu`` under stage ``step``; returns (index of its first occurrence, mu)
        (reference rom/base.py:75-97: list.index, so a repeated mu maps to the earlier index)."""
        space = self.mu_space[step]
        space.append(mu)
        self.mu = mu
        return space.index(mu), mu

    def build_sampling_space(self, num, rnd=None):
        """Random parameter sampler over ``self.grid`` (reference rom/base.py:99-120)."""
        from sklearn.model_selection import ParameterSampler

        return ParameterSampler(param_distributions=self.grid, n_iter=num, random_state=rnd)

    def setup(self, rnd=None):
        """Seed + empty offline report (reference rom/base.py:122-152)."""
        self.random_state = rnd
        off = self.report[Stage.OFFLINE]
        for keys in (Treewalk, TreewalkNonlinear):
            for scalar in (keys.BASIS_AFTER_WALK, keys.BASIS_FINAL, keys.SPECTRUM_MU, keys.ENERGY_MU):
                off[scalar] = None
            for per_mu in (keys.BASIS_TIME, keys.SPECTRUM_TIME, keys.ENERGY_TIME):
                off[per_mu] = dict()

    def create_errors_summary(self):
        """mean / median / max / min of the online errors per parameter (reference rom/base.py:154-163)."""
        import pandas as pd

        rows = {idx: {"mean": np.mean(err), "median": np.median(err), "max": np.max(err), "min": np.min(err)}
                for idx, err in self.errors_rom.items()}
        self.summary_errors = pd.DataFrame(rows).T
