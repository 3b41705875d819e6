"""The ctypes mirrors of the C ABI (`include/b200seq.h`) are checked against the header itself: a C program that
includes the header prints sizeof / offsetof of every struct field, and the numbers must equal what ctypes lays out.
A field added to one side only, a reordered field or a different type fails here instead of corrupting a launch."""
import ctypes
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pairs():
    from hetseq_9cme_b200.ops import kernels, native
    from hetseq_9cme_b200.parallel import p2p
    return [("b200_gemm_desc", native.GemmDesc), ("b200_adam_args", kernels.AdamArgs),
            ("b200_adadelta_args", kernels.AdadeltaArgs), ("b200_ln_args", kernels.LnArgs),
            ("b200_attn_args", kernels.AttnArgs), ("b200_comm_buffer", p2p.CommBuffer)]


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs a C compiler")
def test_ctypes_structs_match_the_header(tmp_path):
    pairs = _pairs()
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "b200seq.h"', 'int main(void) {']
    for cname, ct in pairs:
        lines.append('  printf("%s sizeof %%zu\\n", sizeof(%s));' % (cname, cname))
        for fname, _ in ct._fields_:
            lines.append('  printf("%s %s %%zu\\n", offsetof(%s, %s));' % (cname, fname, cname, fname))
    lines += ['  return 0;', '}']
    src = tmp_path / "abi.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    r = subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]          # a ctypes field the header does not have fails to compile
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    c_layout = {}
    for line in out.splitlines():
        s, f, v = line.split()
        c_layout[(s, f)] = int(v)
    for cname, ct in pairs:
        assert c_layout[(cname, "sizeof")] == ctypes.sizeof(ct), cname      # also catches a header field ctypes lacks
        for fname, _ in ct._fields_:
            assert c_layout[(cname, fname)] == getattr(ct, fname).offset, (cname, fname)


def test_every_declared_entry_point_is_exported():
    """Each `int b200_*(` / `long long b200_*(` prototype of the header resolves in the built library."""
    import re
    from hetseq_9cme_b200.ops import native
    if not native.cuda_lib_available():
        pytest.skip("libb200seq.so not built")
    text = open(os.path.join(ROOT, "include", "b200seq.h")).read()
    names = sorted(set(re.findall(r"^\s*(?:const\s+char\s*\*|int|long long|void)\s+(b200_\w+)\s*\(", text, flags=re.M)))
    assert len(names) > 40
    lib = ctypes.CDLL(str(native.lib_path()))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_every_exported_entry_point_is_declared():
    """... and nothing is exported that the header does not document."""
    import re
    from hetseq_9cme_b200.ops import native
    if not native.cuda_lib_available() or shutil.which("nm") is None:
        pytest.skip("libb200seq.so not built / no nm")
    out = subprocess.run(["nm", "-D", "--defined-only", str(native.lib_path())], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-1].startswith("b200_")}
    declared = set(re.findall(r"\b(b200_\w+)\s*\(", open(os.path.join(ROOT, "include", "b200seq.h")).read()))
    assert exported - declared == set()
