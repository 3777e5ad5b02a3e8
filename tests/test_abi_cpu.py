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
