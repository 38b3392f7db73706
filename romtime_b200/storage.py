"""On-disk result tree of a (hyper-)reduced model run, with the reference's file names and formats
(conventions.py:4-12 ``StorageNames``; the ``dump_*`` / ``load_*`` methods of rom/hrom.py:137-227; the IO helpers of
utils.py:248-279): JSON parameter spaces, pickled bases and validation solutions, CSV error tables -- so that trees
written by the reference load here and vice versa.  Batched sweeps (``RomBatchSolutions``) are written as the
reference's per-parameter ``RomSolutionsStorage`` pickles (base.py:70-79), the FOM fields reconstructed lazily, block by
block, by ONE lift GEMM per block (``RomBatchSolutions.fom_batch``: V uN + g_h, rom/rom.py:496-522)."""
from pathlib import Path

import numpy as np

from .conventions import OperatorType, StorageNames
from .utils import dump_json, dump_pickle, read_json, read_pickle


def _plain(obj):
    """NumPy scalars / arrays -> JSON-able Python values (ujson did this implicitly for the reference)."""
    if isinstance(obj, dict):
        return {str(k): _plain(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [_plain(v) for v in obj]
    if isinstance(obj, np.ndarray):
        return obj.tolist()
    if isinstance(obj, np.generic):
        return obj.item()
    return obj


class ResultTree:
    def __init__(self, path="."):
        self.path = Path(path)
        self.path.mkdir(parents=True, exist_ok=True)

    # ---- parameter spaces (rom/hrom.py:137-151)
    def dump_mu_space(self, rom, name=None):
        dump_json(self.path / (name or StorageNames.MU_SPACE), _plain(rom.mu_space))

    def dump_mu_space_deim(self, mu_space_deim, name=None):
        dump_json(self.path / (name or StorageNames.MU_SPACE_DEIM), _plain(mu_space_deim))

    def load_mu_space(self, name=None):
        return read_json(self.path / (name or StorageNames.MU_SPACE))

    def load_mu_space_deim(self, name=None):
        return read_json(self.path / (name or StorageNames.MU_SPACE_DEIM))

    # ---- bases (rom/hrom.py:153-170; deim/deim.py:77-81, :166-173 for the collateral ones)
    def dump_reduced_basis(self, rom, srom=None):
        dump_pickle(self.path / StorageNames.ROM, np.asarray(rom.basis))
        if srom is not None:
            dump_pickle(self.path / StorageNames.SROM, np.asarray(srom.basis))

    def dump_nonlinear_basis(self, srom):
        dump_pickle(self.path / f"basis_fom_n-mdeim_{OperatorType.TRILINEAR}.pkl", np.asarray(srom.basis_nonlinear))

    def dump_collateral_bases(self, rom):
        """``basis_fom_<type>_<name>.pkl`` of every attached reductor (the file ``load_fom_basis`` reads back)."""
        for red in rom._reductors():
            if red is not None and red.basis_fom is not None:
                dump_pickle(self.path / red.basis_pickle_name, np.asarray(red.basis_fom))

    def load_collateral_basis(self, reductor, keep=None):
        """``reductor.load_fom_basis`` from this tree (deim/deim.py:133-164): basis, interpolation mesh, PT_U."""
        reductor.load_fom_basis(keep=keep, basis=read_pickle(self.path / reductor.basis_pickle_name))
        return reductor

    def load_reduced_basis(self):
        rom = read_pickle(self.path / StorageNames.ROM)
        srom_file = self.path / StorageNames.SROM
        return rom, (read_pickle(srom_file) if srom_file.exists() else None)

    def restore(self, rom, stage=None):
        """``rom.load_from_basis`` from this tree (rom/rom.py:170-182): the basis and the stored parameter space."""
        basis, _ = self.load_reduced_basis()
        rom.load_from_basis(basis=basis, mu_space=self.load_mu_space())
        return rom

    # ---- validation FOM solutions (rom/hrom.py:172-179, :218-226)
    def dump_validation_fom(self, validation_solutions):
        dump_pickle(self.path / StorageNames.VALIDATION_SOLUTIONS, validation_solutions)

    def load_validation_fom(self):
        try:
            return read_pickle(self.path / StorageNames.VALIDATION_SOLUTIONS)
        except FileNotFoundError:
            print("Validation solutions are not defined.")
            return None

    # ---- errors (rom/hrom.py:181-205): one CSV per comparison / operator, columns = parameter index, rows = time steps
    def dump_errors(self, errors, which):
        import pandas as pd

        if which not in errors:
            raise Warning(f"These errors ({which}) have not been computed yet.")
        pd.DataFrame(errors[which]).to_csv(self.path / f"errors_{which}.csv")

    def dump_errors_deim(self, summary_errors_deim):
        import pandas as pd

        for operator, errors in summary_errors_deim.items():
            df = pd.DataFrame(errors)
            if not df.empty:
                df.to_csv(self.path / f"errors_deim_{operator.lower()}.csv")

    def load_errors(self, which):
        import pandas as pd

        return pd.read_csv(self.path / f"errors_{which}.csv", index_col=0)

    # ---- run set-up (rom/hrom.py:207-216)
    def dump_setup(self, fom_params=None, grid_params=None, rom_params=None, deim_params=None, mdeim_params=None, online_params=None):
        out = {"fom_params": fom_params, "mu_space": grid_params, "rom_params": rom_params, "deim_params": deim_params,
               "mdeim_params": mdeim_params, "online_params": online_params}
        dump_json(self.path / StorageNames.SETUP, _plain(out))

    def load_setup(self):
        return read_json(self.path / StorageNames.SETUP)

    # ---- batched sweeps
    def dump_batch(self, batch, prefix="solutions_online", domain=None, lifting=None, block=256):
        """One ``<prefix>_<b>.pkl`` RomSolutionsStorage per parameter of a ``RomBatchSolutions`` (base.py:70-79 layout:
        ts, mu, domain (Nh x nt), fom (Nh x nt), rom (N x nt)); FOM fields are reconstructed ``block`` parameters at a time."""
        from .base import RomSolutionsStorage

        names = []
        for b0 in range(0, len(batch), block):
            b1 = min(len(batch), b0 + block)
            fom = batch.fom_batch(b0, b1, lifting=lifting).cpu().numpy() if batch.has_basis else None
            for b in range(b0, b1):
                st = RomSolutionsStorage(ts=batch.ts, mu=batch.mus[b], domain=None if domain is None else domain(batch.mus[b], batch.ts),
                                         fom=None if fom is None else fom[b - b0].T, rom=batch.rom(b))
                name = str(self.path / f"{prefix}_{b}")
                st.to_pickle(name)
                names.append(name + ".pkl")
        return names
