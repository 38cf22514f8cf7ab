"""-m gpu: tcgen05 / TMEM / TMA descriptor self test (every operand layout the prefill kernels use)."""
import pytest

pytestmark = pytest.mark.gpu


def test_umma_selftest():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import ops
    err = ops.selftest_umma()
    assert err <= 1e-3
