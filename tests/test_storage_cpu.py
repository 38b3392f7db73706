"""CPU: the on-disk result tree (romtime_b200/storage.py) keeps the reference's file names and formats
(conventions.py:4-12, rom/hrom.py:137-227, utils.py:248-279): JSON parameter spaces, pickled bases, CSV error tables."""
import json
import pickle

import numpy as np
import pandas as pd

from romtime_b200.conventions import StorageNames
from romtime_b200.storage import ResultTree


class _Rom:
    def __init__(self):
        self.basis = np.arange(12.0).reshape(4, 3)
        self.mu_space = {"offline": [{"a0": np.float64(9.0), "omega": 17.5}], "online": [], "validation": []}
        self.loaded = None

    def load_from_basis(self, basis, mu_space):
        self.loaded = (basis, mu_space)

    def _reductors(self):
        return []


def test_tree_round_trip(tmp_path):
    tree = ResultTree(tmp_path / "run")
    rom, srom = _Rom(), _Rom()
    srom.basis = rom.basis[:, :2]
    srom.basis_nonlinear = np.ones((5, 2))
    tree.dump_mu_space(rom)
    tree.dump_mu_space_deim({"offline": [{"delta": 0.1}]})
    tree.dump_reduced_basis(rom, srom)
    tree.dump_nonlinear_basis(srom)
    tree.dump_setup(fom_params={"nx": 100, "nt": 50}, rom_params={"num_snapshots": 5}, online_params=None)
    errors = {"rom_srom": {0: np.array([1e-3, 2e-3]), 1: np.array([3e-3, 4e-3])}}
    tree.dump_errors(errors, "rom_srom")
    tree.dump_errors_deim({"MASS": {0: [1e-9, 2e-9]}, "STIFFNESS": {}})
    names = sorted(p.name for p in (tmp_path / "run").iterdir())
    assert names == sorted([StorageNames.MU_SPACE, StorageNames.MU_SPACE_DEIM, StorageNames.ROM, StorageNames.SROM, StorageNames.SETUP,
                            "basis_fom_n-mdeim_trilinear.pkl", "errors_rom_srom.csv", "errors_deim_mass.csv"])
    # formats: plain json / pickle / pandas csv, readable without this package (what the reference's own loaders expect)
    assert json.load(open(tmp_path / "run" / StorageNames.MU_SPACE))["offline"][0] == {"a0": 9.0, "omega": 17.5}
    assert np.array_equal(pickle.load(open(tmp_path / "run" / StorageNames.ROM, "rb")), rom.basis)
    df = pd.read_csv(tmp_path / "run" / "errors_rom_srom.csv", index_col=0)
    assert df.shape == (2, 2) and np.allclose(df["1"].values, [3e-3, 4e-3])
    # and back
    basis, basis_srom = tree.load_reduced_basis()
    assert np.array_equal(basis, rom.basis) and np.array_equal(basis_srom, srom.basis)
    assert tree.load_mu_space()["offline"][0]["omega"] == 17.5
    assert tree.load_setup()["fom_params"] == {"nx": 100, "nt": 50}
    assert np.allclose(tree.load_errors("rom_srom").values, [[1e-3, 3e-3], [2e-3, 4e-3]])
    fresh = _Rom()
    tree.restore(fresh)
    assert np.array_equal(fresh.loaded[0], rom.basis) and fresh.loaded[1]["offline"][0]["a0"] == 9.0
    assert tree.load_validation_fom() is None
    tree.dump_validation_fom({0: np.zeros((3, 2))})
    assert np.array_equal(tree.load_validation_fom()[0], np.zeros((3, 2)))
