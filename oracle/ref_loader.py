"""Import the UNMODIFIED reference (Lindeby/toric-RL-decoder) from /root/reference.

TEST INFRASTRUCTURE ONLY, and only usable in the build container: the GPU box
has no /root/reference, so nothing that runs there (gpu tests, smoke, bench)
may import this.  It is used by `oracle/gen_golden.py` (fixture generation) and
by the optional `tests/test_oracle_vs_reference.py` (skipped when the checkout
is absent).

The reference targets python 3.7 / numpy 1.17 / gym 0.15; four shims make it
importable here without touching its sources:
  1. numpy aliases np.int / np.float (ToricCode.py:151, EnvSet.py:9,39, src/util.py:10);
  2. a stub `gym` (Env, spaces.Box, envs.registration.register, make)
     (ToricCode.py:1,8,44-45; gym_ToricCode/__init__.py:1);
  3. a stub `matplotlib.pyplot` (ToricCode.py:3, import only);
  4. sys.path entries for the checkout and its vendored gym_ToricCode submodule.
"""
import os
import sys
import types

REF_ROOT = os.environ.get("TORIC_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "gym_ToricCode"))


def _install_shims():
    import numpy as np
    if not hasattr(np, "int"):
        np.int = int
    if not hasattr(np, "float"):
        np.float = float

    if "gym" not in sys.modules:
        gym = types.ModuleType("gym")

        class Env(object):
            pass

        class Box(object):
            def __init__(self, low, high, shape=None, dtype=None):
                self.low = np.asarray(low)
                self.high = np.asarray(high)
                self.shape = shape

        spaces = types.ModuleType("gym.spaces")
        spaces.Box = Box
        envs = types.ModuleType("gym.envs")
        registration = types.ModuleType("gym.envs.registration")
        _registry = {}

        def register(id, entry_point=None, **kw):
            _registry[id] = entry_point

        def make(id, **kw):
            mod, cls = _registry[id].split(":")
            import importlib
            return getattr(importlib.import_module(mod), cls)(**kw)

        registration.register = register
        envs.registration = registration
        gym.Env, gym.spaces, gym.envs, gym.make = Env, spaces, envs, make
        sys.modules["gym"] = gym
        sys.modules["gym.spaces"] = spaces
        sys.modules["gym.envs"] = envs
        sys.modules["gym.envs.registration"] = registration

    try:
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt


def load():
    """Returns a namespace with the reference's hot-path callables."""
    if not available():
        raise RuntimeError("reference checkout not present at %s" % REF_ROOT)
    _install_shims()
    for p in (REF_ROOT, os.path.join(REF_ROOT, "gym_ToricCode")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import gym
    import gym_ToricCode  # noqa: F401  (registers toric-code-v0)
    from gym_ToricCode.envs.ToricCode import ToricCode
    from src.EnvSet import EnvSet
    import src.util as util
    ns = types.SimpleNamespace(gym=gym, ToricCode=ToricCode, EnvSet=EnvSet, util=util,
                               numba_util=None, numba_util_actor=None, util_actor=None)
    try:
        import src.numba.util as numba_util
        import src.numba.util_actor as numba_util_actor
        ns.numba_util, ns.numba_util_actor = numba_util, numba_util_actor
    except Exception as e:  # numba missing or incompatible: numpy path is still the oracle
        ns.numba_error = repr(e)
    try:
        import src.util_actor as util_actor
        ns.util_actor = util_actor
    except Exception as e:
        ns.util_actor_error = repr(e)
    return ns
