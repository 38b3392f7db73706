"""CPU: the build's guard against heavily spilling GEMM-core builds (romtime_b200/build.py) reads ptxas -v output correctly."""
from romtime_b200 import build


LOG = """ptxas info    : Compiling entry function '_ZN2dg7k_dgemmILi1ELi1ELi0EEEv14CUtensorMap_stS1_NS_6ParamsE' for 'sm_100a'
ptxas info    : Function properties for _ZN2dg7k_dgemmILi1ELi1ELi0EEEv14CUtensorMap_stS1_NS_6ParamsE
    {frame} bytes stack frame, {st} bytes spill stores, {ld} bytes spill loads
ptxas info    : Used 168 registers, used 1 barriers
ptxas info    : Compiling entry function '_ZN4test9k_gj_smemEv' for 'sm_100a'
ptxas info    : Function properties for _ZN4test9k_gj_smemEv
    400 bytes stack frame, 396 bytes spill stores, 400 bytes spill loads
"""


def test_shipped_level_of_spills_is_accepted():
    assert build._spilling_gemm_kernels(LOG.format(frame=8, st=4, ld=12)) == []
    assert build._spilling_gemm_kernels(LOG.format(frame=0, st=0, ld=0)) == []


def test_heavily_spilling_gemm_is_refused_and_other_kernels_are_not_its_business():
    bad = build._spilling_gemm_kernels(LOG.format(frame=200, st=196, ld=224))
    assert len(bad) == 1 and "k_dgemm" in bad[0] and "196" in bad[0]
