"""The C-ABI library loads without a GPU and exports every symbol include/mppi_b200.h declares;
the ctypes twin binds all of them; config defaults reproduce the reference literals."""
import ctypes as C
import math
import os
import re

import pytest

from motionplanning_b200 import _capi as capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "mppi_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mppi_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(built):
    names = header_functions()
    assert len(names) >= 30
    lib = C.CDLL(capi.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
        assert n in capi.SIGNATURES, f"{n} has no ctypes signature"
    assert sorted(capi.SIGNATURES) == names


def test_struct_layout_matches(built):
    cfg = capi.default_config()
    assert cfg.struct_size == C.sizeof(capi.Config)
    assert C.sizeof(capi.Partial) == 8 + 8 + 8 + 4 + 4 + 8 * 2 * capi.MPPI_MAX_N


def test_defaults_are_the_reference_literals(built):
    c = capi.default_config()
    # MPPI/main.jl:7-27, types.jl:24,30, MPPIUtils.jl:54,127, vehicledynamics.jl:2-18,52
    assert (c.n_x, c.m, c.K, c.N, c.T, c.lambda_) == (7, 2, 1500, 20, 3.0, 25.0)
    assert list(c.Sigma) == [0.05, 0.0, 0.0, 0.1]
    assert list(c.X0)[:7] == [0, 0, 0, 0, 0, 5, 0] and list(c.goal) == [110, 0]
    assert list(c.XL)[:7] == [-10, -20, -2, -math.pi / 2, -math.pi / 2, 1, -math.pi / 9]
    assert list(c.CL) == [-0.5, -2.5] and list(c.CU) == [0.5, 2.5]
    assert c.feasibility_count == 1300 and c.slack_penalty == 1e5
    assert c.obstacle_penalty == 71250.0 and c.terminal_scale == 1e4
    assert (c.la, c.lb, c.mass, c.Izz, c.g, c.mu) == (1.56, 1.64, 2020, 4095, 9.81, 0.8)
    assert (c.KFZF, c.KFZR, c.KFZX) == (1018.28 / 2, 963.34 / 2, 186.22)
    assert list(c.tire) == [-10.4, 1.3, 1.0, 0.1556, -0.0023, 41.3705]
    assert list(c.stage_w) == [1, 1, 5, 3, 2, 10]


def test_create_validates_like_the_reference_without_a_gpu(built):
    lib = capi.load()
    h = capi._H()
    for field, value, text in [("m", 0, b"Controls (0) must be > 0"), ("n_x", 0, b"States (0) must be > 0"),
                               ("K", 0, b"SamplingNumber"), ("N", 0, b"N (0)"), ("n_x", 5, b"must match the model")]:
        c = capi.default_config()
        setattr(c, field, value)
        assert lib.mppi_create(C.byref(c), C.byref(h)) == capi.MPPI_EINVAL
        assert text in lib.mppi_last_error(None)
    c = capi.default_config()
    c.Sigma[3] = -1.0
    assert lib.mppi_create(C.byref(c), C.byref(h)) == capi.MPPI_EINVAL


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under motionplanning_b200/ may reference it."""
    pkg = os.path.join(ROOT, "motionplanning_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".h")) or f == "Makefile":
                txt = open(os.path.join(d, f), errors="ignore").read()
                assert "oracle_lib" not in txt and "libmppi_oracle" not in txt and "oracle/" not in txt, os.path.join(d, f)


def test_julia_shim_mirrors_the_config_struct_field_by_field():
    """julia/MPPIB200.jl cannot be executed here (no julia): at least its isbits mirror of mppi_config_t must list the
    same fields, in the same order, with the same widths as the ctypes mirror, which IS checked against the library."""
    import re

    src = open(os.path.join(ROOT, "julia", "MPPIB200.jl")).read()
    body = src[src.index("struct Config"):]
    body = body[:body.index("\nend")]
    fields = []
    for line in body.splitlines()[1:]:
        for part in line.split("#")[0].split(";"):
            mm = re.match(r"\s*(\w+)::(.+?)\s*$", part)
            if mm:
                fields.append((mm.group(1), mm.group(2)))
    size = {"UInt32": 4, "Int32": 4, "UInt64": 8, "Float64": 8}

    def width(t):
        mm = re.match(r"NTuple\{(\d+),(\w+)\}", t)
        return int(mm.group(1)) * size[mm.group(2)] if mm else size[t]

    ct = [(n, C.sizeof(t)) for n, t in capi.Config._fields_]
    jl = [("lambda_" if n == "lambda" else n, width(t)) for n, t in fields]
    assert jl == ct
    assert sum(w for _, w in jl) <= C.sizeof(capi.Config)
