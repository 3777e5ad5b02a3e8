"""The ctypes structures of ops/native.py against the C structs of include/swarm_b200.h, as the C compiler lays them out:
sizes and the offset of every field.  The binding passes these structs by pointer; a field added on one side only would
corrupt kernel arguments silently (SURVEY.md section 4.2 item 4 guards numerics, this guards the ABI under it)."""
import ctypes
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PAIRS = {"IGemmParams": "sw_igemm_params", "AttnParams": "sw_attn_params", "SchedParams": "sw_sched_params",
         "Peers": "sw_peers", "Xfer": "sw_xfer"}


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs a C compiler")
def test_ctypes_structures_match_the_header(tmp_path):
    from chiaswarm_b200.ops import native

    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "swarm_b200.h"', "int main(void) {"]
    for py, c in PAIRS.items():
        lines.append(f'  printf("{py} size %zu\\n", sizeof({c}));')
        for name, *_ in getattr(native, py)._fields_:
            lines.append(f'  printf("{py} {name} %zu\\n", offsetof({c}, {name}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "abi.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    build = subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                           capture_output=True, text=True)
    assert build.returncode == 0, build.stderr[-2000:]   # (a field the header does not have fails here, by name)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split("\n")
    seen = 0
    for line in filter(None, out):
        py, field, value = line.split()
        struct = getattr(native, py)
        want = ctypes.sizeof(struct) if field == "size" else getattr(struct, field).offset
        assert int(value) == want, f"{py}.{field}: C says {value}, ctypes says {want}"
        seen += 1
    assert seen == sum(len(getattr(native, py)._fields_) + 1 for py in PAIRS)


def test_every_header_struct_has_a_binding_and_every_entry_point_exists_in_the_library_source():
    """Names only (no compiler, no GPU): each `sw_*` struct of the header is bound, and each function the binding declares
    argtypes for is declared in the header."""
    import re

    from chiaswarm_b200.ops import native

    header = open(os.path.join(ROOT, "include", "swarm_b200.h")).read()
    structs = set(re.findall(r"^\} (sw_\w+);", header, re.M))
    assert structs == set(PAIRS.values()), structs ^ set(PAIRS.values())
    declared = set(re.findall(r"\b(sw_\w+)\s*\(", header))
    src = open(native.__file__).read()
    used = set(re.findall(r"\bL\.(sw_\w+)\b", src)) | set(re.findall(r"lib\(\)\.(sw_\w+)\b", src))
    assert used, "binding calls not found"
    missing = sorted(n for n in used if n not in declared)
    assert not missing, f"bound but not declared in the header: {missing}"


def test_argtypes_match_the_header_prototypes():
    """Every function the binding declares `argtypes` for: same number of parameters as the header's prototype and the
    same class per parameter (pointer / 32-bit int / 64-bit int / float).  A missing declaration truncates 64-bit
    pointers to int -- the illegal-address failure this guards against."""
    import re

    from chiaswarm_b200.ops import native

    if not native.LIB_PATH.exists():
        pytest.skip("native library not built")
    lib = native.lib()
    header = re.sub(r"/\*.*?\*/", " ", open(os.path.join(ROOT, "include", "swarm_b200.h")).read(), flags=re.S)
    protos = {m.group(1): m.group(2) for m in re.finditer(r"\b(sw_\w+)\s*\(([^()]*)\)\s*;", header)}

    def c_class(param):
        p = param.strip()
        if "*" in p or "[" in p:
            return "ptr"
        if "float" in p:
            return "f32"
        if "int64_t" in p or "long long" in p or "size_t" in p:
            return "i64"
        return "i32"

    def py_class(t):
        if t in (ctypes.c_void_p, ctypes.c_char_p) or hasattr(t, "_type_") and not isinstance(t._type_, str):
            return "ptr"
        return {ctypes.c_float: "f32", ctypes.c_int64: "i64", ctypes.c_longlong: "i64", ctypes.c_int: "i32",
                ctypes.c_int32: "i32"}[t]

    src = open(native.__file__).read()
    bound = sorted(set(re.findall(r"\bl\.(sw_\w+)\.argtypes", src)))
    assert len(bound) > 25
    called = set(re.findall(r"\bL\.(sw_\w+)\(", src)) | set(re.findall(r"lib\(\)\.(sw_\w+)\(", src))
    assert not sorted(n for n in called if n not in bound and n != "sw_last_error"), "called without argtypes"
    for name in bound:
        assert name in protos, f"{name} has argtypes but no prototype in the header"
        params = [p for p in protos[name].split(",") if p.strip() and p.strip() != "void"]
        want = [c_class(p) for p in params]
        got = [py_class(t) for t in getattr(lib, name).argtypes]
        assert got == want, f"{name}: header {want}, binding {got}"
