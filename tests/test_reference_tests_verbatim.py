"""The reference's own tests, run VERBATIM against the drop-in (VERDICT r1 missing #7).

`tests/golden/ref_tests/*.py` are byte-for-byte copies of the reference's two non-empty test files; here
they are loaded with `wedm` aliased to `sparc_b200` and every `Test*` method is called.  test_state.py
needs no device; test_env_integration.py constructs and steps `WireEDMEnv` (one 1-µs launch)."""
import importlib.util
import os
import sys

import pytest

from conftest import GOLDEN_DIR

REF_TESTS = os.path.join(GOLDEN_DIR, "ref_tests")


def _alias_wedm():
    import sparc_b200
    import sparc_b200.core
    import sparc_b200.core.state

    sys.modules.setdefault("wedm", sparc_b200)
    sys.modules.setdefault("wedm.core", sparc_b200.core)
    sys.modules.setdefault("wedm.core.state", sparc_b200.core.state)


def _run_file(name):
    _alias_wedm()
    spec = importlib.util.spec_from_file_location("ref_" + name[:-3], os.path.join(REF_TESTS, name))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    ran = 0
    for cname in dir(mod):
        cls = getattr(mod, cname)
        if isinstance(cls, type) and cname.startswith("Test"):
            for mname in sorted(dir(cls)):
                if mname.startswith("test_"):
                    getattr(cls(), mname)()
                    ran += 1
    return ran


def test_fixture_is_the_reference_file_when_the_reference_is_present():
    ref_dir = "/root/reference/tests"
    if not os.path.isdir(ref_dir):
        pytest.skip("reference tree not present on this machine")
    for name in ("test_state.py", "test_env_integration.py"):
        assert open(os.path.join(REF_TESTS, name), "rb").read() == open(os.path.join(ref_dir, name), "rb").read()


def test_reference_test_state_verbatim():
    assert _run_file("test_state.py") == 2


@pytest.mark.gpu
def test_reference_test_env_integration_verbatim():
    assert _run_file("test_env_integration.py") == 5
