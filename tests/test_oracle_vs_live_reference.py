"""Short run of oracle/validate_against_reference.py wherever the reference can be imported (the live tree in the
build container, else the oracle/_ref bytecode build).  CPU only."""
import pytest

from oracle import ref_shim, validate_against_reference as v

HAVE = ref_shim.reference_available() or ref_shim.prebuilt_available()
pytestmark = pytest.mark.skipif(not HAVE, reason="reference not importable here")


@pytest.mark.parametrize("S", [4, 5, 6, 8, 9, 10, 12, 13, 16])
def test_rules_fuzz(S):
    assert v.fuzz_rules(S, 4 if S <= 10 else 2, 50 + S) > 20


@pytest.mark.parametrize("S,depths", [(4, (1, 2, 3)), (6, (1, 2, 3)), (8, (1, 2)), (10, (1,))])
def test_policy_fuzz(S, depths):
    assert v.fuzz_policies(S, 3, depths, 60 + S) >= 3


def test_prebuilt_reference_is_the_same_code():
    """oracle/_ref (what bench.py times on the GPU box) behaves as the live tree: same fuzz, same seeds."""
    if not (ref_shim.reference_available() and ref_shim.prebuilt_available()):
        pytest.skip("needs both the live tree and the oracle/_ref build")
    import subprocess, sys, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "oracle", "validate_against_reference.py"), "--prebuilt",
                        "--games", "2", "--sizes", "6,8"], capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0 and "OK:" in r.stdout, r.stdout[-500:] + r.stderr[-1500:]
