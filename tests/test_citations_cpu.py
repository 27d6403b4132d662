"""Every `file.py:line` citation of the reference in the header, the sources and the documents points at a file that exists in the
reference and at lines inside it.  Runs only where /root/reference is mounted (the build container); skipped elsewhere."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("NSB_REFERENCE", "/root/reference")

KNOWN = {}   # basename -> relative paths in the reference


def _index():
    if KNOWN:
        return
    for d, _, files in os.walk(REF):
        if ".git" in d:
            continue
        for f in files:
            if f.endswith(".py"):
                KNOWN.setdefault(f, []).append(os.path.relpath(os.path.join(d, f), REF))


def _files():
    out = [os.path.join(ROOT, "include", "nsb200.h"), os.path.join(ROOT, "DESIGN.md"), os.path.join(ROOT, "INTEGRATION.md"),
           os.path.join(ROOT, "bench.py"), os.path.join(ROOT, "benchlegs.py"), os.path.join(ROOT, "benchlegs_precond.py")]
    for sub in ("navier-stokes-solver_b200", os.path.join("navier-stokes-solver_b200", "csrc"), "oracle", os.path.join("oracle", "c"), "tests"):
        d = os.path.join(ROOT, sub)
        out += [os.path.join(d, f) for f in sorted(os.listdir(d)) if f.endswith((".py", ".cu", ".cuh", ".c", ".h"))]
    return out


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference is mounted in the build container only")
@pytest.mark.parametrize("path", _files(), ids=lambda p: os.path.relpath(p, ROOT))
def test_reference_citations_point_inside_the_reference(path):
    _index()
    text = open(path, errors="replace").read()
    bad = []
    # name.py:12, name.py:12-34, dir/name.py:12,56-78
    for m in re.finditer(r"((?:[\w\-]+/)*[\w\-]+\.py):(\d+(?:-\d+)?(?:,\d+(?:-\d+)?)*)", text):
        rel = m.group(1)
        base = os.path.basename(rel)
        if base not in KNOWN:
            continue                                     # a file of this repo (tests/..., oracle/...), not a reference citation
        cands = [p for p in KNOWN[base] if p.endswith(rel)] or KNOWN[base]
        nlines = max(sum(1 for _ in open(os.path.join(REF, p), errors="replace")) for p in cands)
        for part in m.group(2).split(","):               # 96,126,145-146: every member of a comma list (no spaces) is checked
            a, _, b = part.partition("-")
            a, b = int(a), int(b or a)
            if not (1 <= a <= b <= nlines):
                bad.append((rel + ":" + part, nlines))
    assert bad == []
