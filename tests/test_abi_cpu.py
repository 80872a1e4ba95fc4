"""The C-ABI library loads on a machine without a GPU and exports every symbol
include/mpci.h declares; argument validation works without touching a device."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as entry
    from mpc_interface_b200 import device

    if not os.path.exists(device.LIB_PATH):
        entry.build()
    return device.lib()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mpci.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mpci_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(lib):
    from mpc_interface_b200 import device

    names = declared_symbols()
    assert names == device.exported_symbols()
    for name in names:
        assert getattr(lib, name) is not None


def test_version_and_error_string(lib):
    assert lib.mpci_version() == 100
    assert isinstance(lib.mpci_last_error(), bytes)


def test_argument_validation_needs_no_gpu(lib):
    # negative sizes are refused before any CUDA call
    rc = lib.mpci_extend_matrices(None, None, None, None, 1, 0, 1, 5, None)
    assert rc < 0 and b"sizes must be positive" in lib.mpci_last_error()
    rc = lib.mpci_extend_matrices(None, None, None, None, 1, 64, 1, 5, None)
    assert rc < 0 and b"exceed" in lib.mpci_last_error()
    rc = lib.mpci_gemm(0, -1, 1, 1, 1.0, None, 1, 0, None, 1, 0, 0.0, None, 1, 0, 1, None)
    assert rc < 0
    rc = lib.mpci_assemble(None, None, None, None, None, None, None, 4, None)
    assert rc < 0 and b"NULL plan" in lib.mpci_last_error()
    assert lib.mpci_plan_create(None, None) < 0
    assert lib.mpci_plan_destroy(None) == 0


def test_product_path_fails_loudly_without_gpu(lib):
    """No CPU fallback: without a device every matrix-producing call raises."""
    import numpy as np
    from mpc_interface_b200 import tools

    if lib.mpci_device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(RuntimeError):
        tools.extend_matrices(5, np.eye(2), np.ones([2, 1]))
    from mpc_interface_b200 import Formulation

    with pytest.raises(RuntimeError):
        Formulation()._ops()


def test_cpp_header_builds_and_links(lib):
    """include/mpci_gecko.hh (gecko::tools::extend_matrices, cpp/include/mpc-interface/
    tools.hh:24-25) compiles with g++ and links against libmpci.so; the values are
    checked on the GPU (test_gpu_parity.py::test_gecko_cpp_extend_matrices)."""
    import _gecko_build

    done = _gecko_build.run("--link-only")
    assert done.returncode == 0, done.stdout + done.stderr
    assert "mpci version 100" in done.stdout
    if lib.mpci_device_count() == 0:  # no CPU fallback behind the C++ face either
        failed = _gecko_build.run()
        assert failed.returncode != 0 and "extend_matrices" in (failed.stdout + failed.stderr)


def test_plan_desc_of_the_documentation_matches_the_header():
    """The PlanDesc printed in INTEGRATION.md, the one in device.py and mpci_plan_desc of
    include/mpci.h list the same fields in the same order with the same C types."""
    from mpc_interface_b200 import device

    header = open(os.path.join(ROOT, "include", "mpci.h")).read()
    body = re.search(r"typedef struct \{(.*?)\} mpci_plan_desc;", header, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        ctype, names = decl.rsplit(" ", 1)[0], decl
        pointer = "*" in decl
        for name in re.sub(r"^(const\s+)?(int|double)\s*\*?\s*", "", names).split(","):
            fields.append((name.strip().lstrip("*"), "p" if pointer else "i"))
    own = [(n, "p" if t is ctypes.c_void_p else "i") for n, t in device.PlanDesc._fields_]
    assert own == fields
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    block = re.search(r"class PlanDesc\(ctypes.Structure\):.*?_fields_ = \[(.*?)\]\n", doc, flags=re.S).group(1)
    assert re.findall(r'\("(\w+)", (\w)\)', block) == fields


def _header_fields(struct_name):
    header = open(os.path.join(ROOT, "include", "mpci.h")).read()
    end = header.index("} %s;" % struct_name)
    body = header[header.rindex("typedef struct {", 0, end) + len("typedef struct {"):end]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        pointer = "*" in decl
        for name in re.sub(r"^(const\s+)?(int|double)\s*\*?\s*", "", decl).split(","):
            fields.append((name.strip().lstrip("*").strip(), "p" if pointer else "i"))
    return fields


@pytest.mark.parametrize("struct_name,mirror", [("mpci_lti_desc", "LtiDesc"), ("mpci_lti_tables", "LtiTables"),
                                                ("mpci_lti_embed", "LtiEmbed"),
                                                ("mpci_lti_sched_desc", "LtiSchedDesc"),
                                                ("mpci_lti_coupled_tables", "LtiCoupledTables")])
def test_lti_structs_of_the_binding_match_the_header(struct_name, mirror):
    """The ctypes mirrors in lti.py list the fields of include/mpci.h in the same order with
    the same kinds (int / pointer): a field added on one side only would shift every later one."""
    from mpc_interface_b200 import lti

    own = [(n, "p" if t is ctypes.c_void_p else "i") for n, t in getattr(lti, mirror)._fields_]
    assert own == _header_fields(struct_name)


def test_embedded_plan_validation_needs_no_gpu(lib):
    """mpci_lti_plan_create_embedded refuses NULL / overlapping layouts before any CUDA call."""
    import numpy as np
    from mpc_interface_b200 import lti

    assert lib.mpci_lti_plan_create_embedded(None, None, None) < 0
    n, m, N = 2, 1, 4
    z = lambda *shape: np.zeros(shape)  # noqa: E731
    t = dict(Qxx=z(n, n), Quu=z(m, m), Tx=z(n, 1), Tu=z(m, 1))
    desc = lti.LtiDesc(n, m, N, 0, 0, t["Qxx"].ctypes.data, t["Quu"].ctypes.data, t["Tx"].ctypes.data,
                       t["Tu"].ctypes.data, None, None, None, None, None, None)
    voff, zcol = np.array([0], dtype=np.int32), np.array([2], dtype=np.int32)  # z inside the input block
    pzz, hptr, kind = z(1, 1), np.zeros(1, dtype=np.int32), np.zeros(0, dtype=np.int32)
    embed = lti.LtiEmbed(N + 1, 0, voff.ctypes.data, None, 1, zcol.ctypes.data, 0, 0, None, None, None, None,
                         None, None, None, pzz.ctypes.data, None, hptr.ctypes.data, None, None, kind.ctypes.data)
    handle = ctypes.c_void_p()
    rc = lib.mpci_lti_plan_create_embedded(ctypes.byref(desc), ctypes.byref(embed), ctypes.byref(handle))
    assert rc < 0 and b"not a free column" in lib.mpci_last_error()
    rc = lib.mpci_lti_assemble_coupled(None, None, 0, None, 0, None, None, None, None, None, None, None, None,
                                       None, None, 1, 1, None)
    assert rc < 0 and b"NULL plan" in lib.mpci_last_error()
    assert lib.mpci_extend_matrices_static(None, None, None, None, 1, 0, 1, 5, None) < 0
    assert lib.mpci_extend_static_capacity() == 524288


def test_scheduled_cost_stage_validation_needs_no_gpu(lib):
    """mpci_lti_sched_create / _add refuse malformed descriptions before any CUDA call."""
    import numpy as np
    from mpc_interface_b200 import lti

    handle = ctypes.c_void_p()
    assert lib.mpci_lti_sched_create(None, ctypes.byref(handle)) < 0
    i32 = lambda *v: np.array(v, dtype=np.int32)  # noqa: E731
    ptr, rows, zcol = i32(0, 2), i32(1, 4), i32(2)

    def create(n=2, m=1, N=4, nv=5, nz=0, zc=None, rp=ptr, rk=rows, terms=1, ep=None):
        desc = lti.LtiSchedDesc(n, m, N, nv, None, nz, zc, 0, terms, rp.ctypes.data, rk.ctypes.data,
                                None, None, None, None, ep, None, None, None, None, None, None)
        return lib.mpci_lti_sched_create(ctypes.byref(desc), ctypes.byref(handle))

    assert create() < 0 and b"outside the horizon" in lib.mpci_last_error()  # sample 4 of 4
    assert create(nv=3, rk=i32(1, 3)) < 0 and b"columns" in lib.mpci_last_error()  # the input does not fit
    assert create(nz=1, zc=zcol.ctypes.data, rk=i32(1, 3)) < 0 and b"inside input" in lib.mpci_last_error()
    assert create(terms=0) < 0 and b"bad sizes" in lib.mpci_last_error()
    assert create(rp=i32(0, 0), rk=i32(1, 3)) < 0 and b"no scheduled rows" in lib.mpci_last_error()
    assert create(rk=i32(1, 3), ep=ptr.ctypes.data) < 0 and b"ent_k / ent_val missing" in lib.mpci_last_error()
    assert not handle.value
    rc = lib.mpci_lti_sched_rows(None, None, 0, None, 0, None, None, 0, None, None, 4, 1, None)
    assert rc < 0 and b"null argument" in lib.mpci_last_error()
    rc = lib.mpci_lti_sched_add(None, None, 0, None, 0, None, None, None, 0, None, None, 1, None)
    assert rc < 0 and b"null argument" in lib.mpci_last_error()
    assert lib.mpci_lti_sched_destroy(None) == 0
