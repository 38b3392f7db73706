"""String conventions shared with the reference (reference conventions.py) -- only the groups the
offline-compression / online-solve path reads.  The VALUES are API surface (dictionary keys, ``which``
selectors, file names) and therefore identical to the reference's; everything else is ours."""


class _Names:
    """Namespace of string constants; ``values()`` lists them (used for validation)."""

    @classmethod
    def values(cls):
        return [v for k, v in vars(cls).items() if not k.startswith("_") and isinstance(v, str)]


class ProblemType(_Names):          # conventions.py:15-20 -- the ``which`` selector of interpolate()
    FOM, ROM, SROM, HROM = "fom", "rom", "srom", "hrom"


class Stage(_Names):                # conventions.py:99-103 -- keys of Reductor.mu_space / report
    OFFLINE, VALIDATION, ONLINE = "offline", "validation", "online"


class BDF(_Names):                  # conventions.py:42-45 -- fom.BDF_SCHEME values
    ONE, TWO = "1", "2"


class EmpiricalInterpolation(_Names):   # conventions.py:92-96 -- reductor TYPE tags
    DEIM, MDEIM, NONLINEAR = "DEIM", "MDEIM", "N-MDEIM"


class OperatorType(ProblemType):    # conventions.py:48-59 -- add_hyper_reductor(which=...)
    CONVECTION = "convection"
    FORCING = "forcing"
    LIFTING = "lifting"
    MASS = "mass"
    TRILINEAR = "trilinear"
    NONLINEAR = "nonlinear"
    NONLINEAR_LIFTING = "nonlinear-lifting"
    REDUCED_BASIS = "reduced-basis"
    RHS = "rhs"
    STIFFNESS = "stiffness"


def _treewalk(prefix):
    keys = dict(BASIS_AFTER_WALK="basis-shape-after-tree-walk", BASIS_FINAL="basis-shape-final",
                BASIS_TIME="basis-shape-time", ENERGY_MU="energy-mu", ENERGY_TIME="energy-time",
                SPECTRUM_MU="spectrum-mu", SPECTRUM_TIME="spectrum-time")
    return type("Treewalk" + ("Nonlinear" if prefix else ""), (_Names,), {k: prefix + v for k, v in keys.items()})


Treewalk = _treewalk("")                # conventions.py:62-69 -- report keys of the tree walk
TreewalkNonlinear = _treewalk("N-")     # conventions.py:72-79


class RomParameters(_Names):        # conventions.py:121-139 -- tree_walk_params / tolerances keys
    NUM_ONLINE = "num_online"
    SROM_TRUNCATE = "srom_truncate"
    SROM_KEEP = "srom_num"
    NMDEIM_SIZE = "mdeim_truncate"
    NUM_BASIS = "num_phi"
    NUM_MU = "num_mu"
    NUM_SNAPSHOTS = "num_snapshots"
    NUM_TIME = "num_time"
    TOL_BASIS = "tol_phi"
    TOL_MU = "tol_mu"
    TOL_TIME = "tol_time"
    TS = "ts"


class PistonParameters(_Names):     # conventions.py:148-155 -- keys of the mu dictionaries
    ALPHA, DELTA, GAMMA, OMEGA, A0 = "alpha", "delta", "gamma", "omega", "a0"
    MACH_PISTON, NONLINEARITY = "piston_mach", "eta"


class StorageNames(_Names):         # conventions.py:4-12 -- on-disk artefacts
    ROM, SROM = "basis_rom.pkl", "basis_srom.pkl"
    VALIDATION_SOLUTIONS = "validation_solutions.pkl"
    SETUP, MU_SPACE, MU_SPACE_DEIM = "setup.json", "mu_space.json", "mu_space_deim.json"


class Domain(_Names):               # conventions.py:114-118 -- fom.domain keys
    NT, NX, T, L0 = "nt", "nx", "T", "L0"
