"""The oracle is independent of the product for its constants: it carries its own table of the
reference's literals (oracle_config_default, typed from the reference sources) and its own ctypes
twin of mppi_config_t.  These tests check both against the product's twins, field by field, so a
wrong literal on either side is caught."""
import ctypes as C
import os
import subprocess
import sys

import oracle_lib as orc
from motionplanning_b200 import _capi as capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _as_py(v):
    return list(v) if hasattr(v, "__len__") else v


def test_config_twins_have_the_same_layout(built):
    a, b = orc.Config._fields_, capi.Config._fields_
    assert [f[0] for f in a] == [f[0] for f in b]
    for (n, ta), (_, tb) in zip(a, b):
        assert C.sizeof(ta) == C.sizeof(tb), n
        assert getattr(orc.Config, n).offset == getattr(capi.Config, n).offset, n
    assert C.sizeof(orc.Config) == C.sizeof(capi.Config)


def test_oracle_defaults_equal_product_defaults_field_by_field(built):
    o, p = orc.default_config(), capi.default_config()
    for name, _ in orc.Config._fields_:
        assert _as_py(getattr(o, name)) == _as_py(getattr(p, name)), name


def test_oracle_wrapper_never_loads_the_product():
    """oracle_lib + scenarios in a fresh interpreter: no motionplanning_b200 module, no libmppi_b200.so mapped."""
    code = (
        "import sys, os, importlib.util\n"
        f"sys.path.insert(0, {os.path.join(ROOT, 'tests')!r})\n"
        "import oracle_lib as orc\n"
        f"spec = importlib.util.spec_from_file_location('scenarios', {os.path.join(ROOT, 'motionplanning_b200', 'scenarios.py')!r})\n"
        "sc = importlib.util.module_from_spec(spec); spec.loader.exec_module(sc)\n"
        "args, kw, circles, blocks = sc.config(0)\n"
        "cfg = orc.make_config(args, kw)\n"
        "assert cfg.K == 1024 and cfg.obstacle_penalty == 71250.0\n"
        "assert not any(m.startswith('motionplanning_b200') for m in sys.modules), sorted(sys.modules)\n"
        "maps = open('/proc/self/maps').read()\n"
        "assert 'libmppi_oracle.so' in maps and 'libmppi_b200.so' not in maps\n"
        "print('ok')\n"
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr
