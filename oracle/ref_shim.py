"""TEST INFRASTRUCTURE -- import shim for the live Python reference.

In the build container it imports the reference from ``/root/reference`` (golden generators under
``tests/golden/``, ``oracle/validate_against_reference.py``).  On the GPU box, where that tree does not exist,
``load_reference()`` imports the bytecode build of the same three modules that ``oracle/make_ref.py`` left in
``oracle/_ref/`` -- that is how ``bench.py``'s CPU legs time the reference's own code.  Nothing in the product
package imports this file.

The reference's ``othello.py:3-7`` imports ``gym``/``pyglet`` (absent here) but only
uses ``gym.Env`` as a base class and ``spaces.Discrete/Box`` as shape metadata
(``othello.py:245-254``); ``util.py:5-6`` imports ``ppo`` (pulls in tensorboard) and a
``Rainbow`` package that is not in the tree.  We inject inert stand-ins for those
modules so that ``othello``, ``simple_policies`` and ``util`` import UNMODIFIED.
"""
import contextlib
import io
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("OTHELLO_REFERENCE_ROOT", "/root/reference")


def reference_available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "othello.py"))


def _install_stubs():
    if "gym" not in sys.modules:
        gym = types.ModuleType("gym")

        class Env(object):
            metadata = {}

            def close(self):
                pass

        class _Discrete(object):
            def __init__(self, n):
                self.n = n
                self.shape = ()

        class _Box(object):
            def __init__(self, low, high, *a, **k):
                self.low, self.high = low, high
                self.shape = getattr(low, "shape", None)

        spaces = types.ModuleType("gym.spaces")
        spaces.Discrete = _Discrete
        spaces.Box = _Box
        gym.Env = Env
        gym.spaces = spaces
        gym.wrappers = types.ModuleType("gym.wrappers")
        sys.modules["gym"] = gym
        sys.modules["gym.spaces"] = spaces
        sys.modules["gym.wrappers"] = gym.wrappers
    if "pyglet" not in sys.modules:
        pyglet = types.ModuleType("pyglet")
        pyglet.gl = types.ModuleType("pyglet.gl")
        sys.modules["pyglet"] = pyglet
        sys.modules["pyglet.gl"] = pyglet.gl
    if "Rainbow" not in sys.modules:
        rb = types.ModuleType("Rainbow")
        rba = types.ModuleType("Rainbow.agent")
        rba.Agent = type("Agent", (), {})
        rb.agent = rba
        sys.modules["Rainbow"] = rb
        sys.modules["Rainbow.agent"] = rba
    if "ppo" not in sys.modules:
        # util.py:5 only needs the name PPO for create_policy('ppo'); the real module
        # drags in tensorboard.  A stand-in keeps make_state/undo_state importable.
        ppo = types.ModuleType("ppo")
        ppo.PPO = type("PPO", (), {})
        sys.modules["ppo"] = ppo


def prebuilt_available():
    from oracle import make_ref
    return make_ref.available()


def load_reference(prefer_prebuilt=False):
    """Returns (othello, simple_policies, util) modules of the reference: the live tree when it is present, else
    (or when prefer_prebuilt) the bytecode build in oracle/_ref/."""
    if (prefer_prebuilt or not reference_available()) and prebuilt_available():
        from oracle import make_ref
        _install_stubs()
        othello, simple_policies, util = make_ref.load_modules()
        return othello, simple_policies, util
    if not reference_available():
        raise RuntimeError("reference not available: no tree at %s and no build in oracle/_ref/" % REFERENCE_ROOT)
    _install_stubs()
    # The product package has same-named modules (othello.py, simple_policies.py,
    # util.py) *inside* gymothelloenv_b200/, never at top level, so these top-level
    # names resolve to the reference only.
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import othello  # noqa
    import simple_policies  # noqa
    import util  # noqa
    assert os.path.dirname(os.path.abspath(othello.__file__)) == os.path.abspath(REFERENCE_ROOT)
    return othello, simple_policies, util


@contextlib.contextmanager
def muted():
    """The reference wrappers print on every reset/terminal (othello.py:65,156,477-500)."""
    with contextlib.redirect_stdout(io.StringIO()):
        yield
