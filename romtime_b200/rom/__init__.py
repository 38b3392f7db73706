"""POD + reduced-order model constructors (reference rom/__init__.py)."""
from .base import Reductor
from .pod import orth, orth_batch
from .rom import Mfp1Evaluator, PistonEvaluator, RomConstructor, RomConstructorMoving, RomConstructorNonlinear

__all__ = ["Reductor", "RomConstructor", "RomConstructorMoving", "RomConstructorNonlinear", "PistonEvaluator", "Mfp1Evaluator", "orth", "orth_batch"]
