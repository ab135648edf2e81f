"""The C-ABI library loads without a GPU and exports exactly what include/mcov.h declares;
argument validation (no compute) returns the documented status codes; ctypes mirrors of the
structs have the C sizes (checked by compiling the header with gcc as plain C)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mcov.h")


@pytest.fixture(scope="module")
def lib():
    from marl_coverage_b200 import build, _lib
    build.build()
    return _lib.load()


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mcov_[a-z0-9_]+)\s*\(", src)))


def test_exports_match_header(lib):
    from marl_coverage_b200 import _lib
    declared = _declared_functions()
    assert declared == sorted(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r"\bT (mcov_[a-z0-9_]+)$", out, flags=re.M)))
    assert exported == declared                      # nothing undeclared leaks out of the .so


def test_abi_version_and_struct_sizes(lib, tmp_path):
    from marl_coverage_b200 import _lib
    assert lib.mcov_abi_version() == _lib.ABI_VERSION
    prog = tmp_path / "sizes.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "mcov.h"\n'
                    'int main(void){printf("%zu %zu %zu %zu %d\\n", sizeof(mcov_config), sizeof(mcov_layout),'
                    ' offsetof(mcov_config, collision_penalty), offsetof(mcov_layout, off_hdr),'
                    ' MCOV_ABI_VERSION);return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    str(prog), "-o", str(exe)], check=True)
    a, b, c, d, v = (int(t) for t in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split())
    assert a == C.sizeof(_lib.McovConfig) and b == C.sizeof(_lib.McovLayout)
    assert c == _lib.McovConfig.collision_penalty.offset and d == _lib.McovLayout.off_hdr.offset
    assert v == _lib.ABI_VERSION


def test_plane_geometry(lib):
    from marl_coverage_b200 import _lib
    assert _lib.plane_geometry(102, 102) == (4, 104, 416)
    assert _lib.plane_geometry(1075, 1075) == (34, 1076, 34 * 1076)
    with pytest.raises(_lib.McovError) as ei:
        _lib.plane_geometry(0, 5)
    assert ei.value.code == -1


def _cfg(**kw):
    from marl_coverage_b200 import _lib
    c = _lib.McovConfig()
    c.struct_size = C.sizeof(_lib.McovConfig)
    c.env_type, c.num_envs, c.num_robots = _lib.ENV_DEC, 4, 4
    c.sensor_type, c.sensor_range, c.egoradius, c.num_lasers = _lib.SENSOR_SQUARE, 2, 5, 21
    for k, v in kw.items():
        setattr(c, k, v)
    return c


@pytest.mark.parametrize("kw,code", [
    (dict(struct_size=8), -6), (dict(env_type=7), -1), (dict(num_envs=0), -1), (dict(num_robots=0), -1),
    (dict(num_robots=33), -1), (dict(sensor_type=5), -1), (dict(sensor_range=16), -1),
    (dict(egoradius=-1), -1), (dict(mini_map_rad=99), -1), (dict(sensor_type=1, num_lasers=20), -1),
    (dict(comm_radius=-2), -1), (dict(env_type=1, senseradius=40), -1)])
def test_create_rejects_bad_config(lib, kw, code):
    h = C.c_void_p()
    rc = lib.mcov_create(C.byref(_cfg(**kw)), C.byref(h))
    assert rc == code
    assert lib.mcov_last_error()                       # a message is always left behind
    assert not h.value


def test_null_and_state_errors(lib):
    h = C.c_void_p()
    assert lib.mcov_create(None, C.byref(h)) == -2
    assert lib.mcov_step(None, None, 0, None, None, None, None, None, None) == -2
    assert lib.mcov_reset(None, None, None, None, 0, None, None) == -2
    assert lib.mcov_observe(None, None, None) == -2
    assert lib.mcov_percent_covered(None, None, None) == -2
    assert lib.mcov_destroy(None) == 0
    assert lib.mcov_launch_count(None) == 0
    assert lib.mcov_pack_maps(None, 1, 4, 4, 0, None, None, None, None) == -2


def test_no_cpu_fallback():
    """Without a CUDA device the product path raises; it never computes on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from marl_coverage_b200 import DecGridRL, VecDecGridRL, gridload
    train, test = gridload(None)
    cfg = dict(numrobot=2, maxsteps=10, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30,
               dist_reward=0, train_maxsteps=10, test_maxsteps=10, egoradius=2, mini_map_rad=0, comm_radius=0,
               allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0,
               sensor_type="square_sensor", sensor_config={"range": 1})
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        VecDecGridRL(train, cfg, num_envs=2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        DecGridRL(train, cfg, test_set=test)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "marl_coverage_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
