"""pytest configuration: `gpu` marker, import paths, shared fixtures.

CPU suite (`-m "not gpu"`): oracle vs golden vectors / the compiled reference TU, the product's
device headers compiled for the host, C-ABI symbol check, multi-rank tile logic over gloo.
GPU suite (`-m gpu`): parity of the CUDA path, always through the C-ABI of liborb200.so.
Nothing here reads /root/reference at run time (it does not exist on the GPU box).
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    lib = os.path.join(ROOT, "optix_renderer_b200", "liborb200.so")
    ora = os.path.join(ROOT, "oracle", "liboracle.so")
    if not (os.path.exists(lib) and os.path.exists(ora)):
        import __graft_entry__ as g

        if not os.path.exists(lib):
            g.build_library()
        if not os.path.exists(ora):
            g.build_oracle()


_ensure_built()

import _oracle  # noqa: E402


# oracle/_ref (the reference's own device source compiled here) is git-ignored: on a fresh clone without
# /root/reference the strongest checker silently drops out.  Which kind each test got is printed in the session
# summary, and ORB_REQUIRE_REF=1 turns every such skip / fall-back into a failure.
_ORACLE_KINDS_USED = {}


def _note(kind):
    _ORACLE_KINDS_USED[kind] = _ORACLE_KINDS_USED.get(kind, 0) + 1


def _ref_or_skip(kind):
    if not _oracle.available(kind):
        if os.environ.get("ORB_REQUIRE_REF") == "1":
            pytest.fail("ORB_REQUIRE_REF=1 and oracle/_ref (%s) is not built" % kind)
        _note("skipped: " + kind)
        pytest.skip("oracle/_ref not built (reference absent)")
    _note(kind)
    return _oracle.Oracle(kind)


def pytest_terminal_summary(terminalreporter):
    if _ORACLE_KINDS_USED:
        terminalreporter.write_line("oracle kinds used: " + ", ".join("%s x%d" % kv for kv in sorted(_ORACLE_KINDS_USED.items())))


@pytest.fixture(scope="session")
def restatement():
    return _oracle.Oracle("restatement")


@pytest.fixture(scope="session")
def ref_literal():
    return _ref_or_skip("reference-literal")


@pytest.fixture(scope="session")
def ref_intended():
    return _ref_or_skip("reference-intended")


def best_oracle(mode):
    """The strongest checker available: the compiled reference TU, else the restatement."""
    kind = "reference-intended" if mode == 1 else "reference-literal"
    if not _oracle.available(kind):
        if os.environ.get("ORB_REQUIRE_REF") == "1":
            pytest.fail("ORB_REQUIRE_REF=1 and oracle/_ref (%s) is not built" % kind)
        kind = "restatement"
    _note(kind)
    return _oracle.Oracle(kind)


@pytest.fixture(scope="session")
def hostcheck():
    import _hostcheck

    return _hostcheck.HostCheck()


def has_gpu():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=30)
        return out.returncode == 0 and "GPU" in out.stdout
    except Exception:
        return False


@pytest.fixture(scope="session")
def renderer_factory():
    """GPU tests only: makes Renderer objects on cuda:0 and closes them at the end."""
    from optix_renderer_b200 import api

    made = []

    def make(scene=None, **kw):
        r = api.Renderer(scene, **kw)
        made.append(r)
        return r

    yield make
    for r in made:
        r.close()
