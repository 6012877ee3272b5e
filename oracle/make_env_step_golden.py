"""Fixtures for the discrete env transition, produced by the reference's OWN method objects  --  TEST INFRASTRUCTURE.

The env modules cannot be imported here (gym is absent and meent_env.py:58 does not parse), so the three methods on the
path are cut out of the reference sources with ``ast`` and compiled unchanged:

  * ``DeflectorBase.flip``              deflector-gym/deflector_gym/envs/base.py:37-42
  * ``MeentIndex.reset`` / ``.step``    deflector-gym/deflector_gym/envs/meent_env.py:84-99
  * ``random_bunch_init``               deflector-gym/deflector_gym/envs/utils.py:29-44

They are bound to a stand-in object whose ``get_efficiency`` is a closed-form float64 function of the structure (the RCWA
value itself is pinned elsewhere), and driven through trajectories whose actions include out-of-range positions.  Stored:
every (state, action) -> state, the efficiencies and the rewards, bit for bit.

Run in the build container (needs /root/reference):  python -m oracle.make_env_step_golden
"""
from __future__ import annotations

import ast
import os
import textwrap

import numpy as np

from oracle import ref_bridge

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
ENVS = os.path.join(ref_bridge.REF_ROOT, "deflector-gym", "deflector_gym", "envs")


def _method(path, cls, name):
    """The unmodified source of ``cls.name`` in ``path``, compiled as a plain function."""
    with open(path) as fh:
        src = fh.read()
    try:
        tree = ast.parse(src)
        nodes = [n for c in ast.walk(tree) if isinstance(c, ast.ClassDef) and c.name == cls
                 for n in c.body if isinstance(n, ast.FunctionDef) and n.name == name]
        seg = ast.get_source_segment(src, nodes[0])
    except SyntaxError:            # meent_env.py:58 does not parse as a whole: cut the method out by indentation
        lines = src.splitlines()
        ci = next(i for i, l in enumerate(lines) if l.startswith(f"class {cls}"))
        mi = next(i for i in range(ci, len(lines)) if lines[i].startswith(f"    def {name}("))
        me = next((i for i in range(mi + 1, len(lines)) if lines[i].strip() and not lines[i].startswith("     ")
                   and not lines[i].startswith("\t")), len(lines))
        seg = "\n".join(lines[mi:me])
    ns = {"np": np}
    exec(textwrap.dedent(seg), ns)
    return ns[name]


def stub_efficiency(struct) -> float:
    """Deterministic float64 stand-in for get_efficiency (same function in tests/): position-weighted mean."""
    s = np.asarray(struct, dtype=np.float64)
    w = np.cos(0.37 * np.arange(s.size, dtype=np.float64))
    return float((s * w).sum() / s.size)


def main():
    flip = _method(os.path.join(ENVS, "base.py"), "DeflectorBase", "flip")
    reset = _method(os.path.join(ENVS, "meent_env.py"), "MeentIndex", "reset")
    step = _method(os.path.join(ENVS, "meent_env.py"), "MeentIndex", "step")
    ns = {}
    with open(os.path.join(ENVS, "utils.py")) as fh:
        src = fh.read()
    exec("import numpy as np\n" + src[src.index("def random_bunch_init"):], ns)
    bunch = ns["random_bunch_init"]

    class Stand:                       # carries exactly the attributes the three methods touch
        n_cells = 256
        struct = None
        eff = None
        def initialize_struct(self, *a, **k):      # base.py:30-32
            return bunch(*a, **k)

        def get_efficiency(self, s):
            return stub_efficiency(s)

    Stand.flip, Stand.reset, Stand.step = flip, reset, step       # the reference's functions, as methods

    T, seeds = 48, [3, 17, 2026]
    rng = np.random.default_rng(99)
    out = {"seeds": np.array(seeds), "init": [], "actions": [], "states": [], "effs": [], "rewards": [], "dones": []}
    for seed in seeds:
        np.random.seed(seed)
        env = Stand()
        s0 = env.reset()
        acts = rng.integers(0, 256, T)
        acts[[5, 11, 23, 31, 40]] = [-1, 256, 300, -7, 255]       # out-of-range positions do nothing (base.py:40-42)
        acts[12] = acts[13] = acts[14]                              # the same cell flipped three times
        states, effs, rews, dones = [], [env.eff], [], []
        for a in acts:
            s, r, d, info = env.step(int(a))
            states.append(s); effs.append(env.eff); rews.append(r); dones.append(d)
            assert info == {}
        out["init"].append(s0); out["actions"].append(acts); out["states"].append(np.stack(states))
        out["effs"].append(np.array(effs)); out["rewards"].append(np.array(rews)); out["dones"].append(np.array(dones))
    np.savez_compressed(os.path.join(GOLD, "env_step_golden.npz"),
                        **{k: np.asarray(v) for k, v in out.items()})
    print("env_step_golden.npz:", {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
