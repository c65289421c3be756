"""CPU tests of the drop-in boundary: libag2.so loads, exports every symbol include/ag2.h
declares, its PODs have the documented sizes, and it refuses to run without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from agile_grasp2_b200 import capi
from oracle import binding as ob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ag2.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ag2_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = capi.load().lib
    syms = declared_symbols()
    assert len(syms) >= 30
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_oracle_mirrors_the_stage_api():
    lib = ob.load().lib
    for s in ("create", "destroy", "set_cloud", "set_sample_indices", "set_sample_points", "compute_normals",
              "generate_hypotheses", "prune", "render_images", "classify", "select", "detect", "get_hypotheses"):
        assert hasattr(lib, "ag2o_" + s)


def test_pod_sizes_and_abi_version():
    assert capi.HYPOTHESIS_DTYPE.itemsize == 192 and capi.SWEEP_DTYPE.itemsize == 16 and capi.FRAME_DTYPE.itemsize == 112
    assert C.sizeof(capi.Params) == 432
    lib = capi.load().lib
    assert lib.ag2_abi_version() == 1
    lib.ag2_strerror.restype = C.c_char_p
    assert b"no CPU fallback" in lib.ag2_strerror(capi.AG2_E_NO_DEVICE)


def test_default_params_match_reference_defaults():
    p, q = capi.default_params(), ob.reference_default_params()
    for f, _ in capi.Params._fields_:
        a, b = getattr(p, f), getattr(q, f)
        if hasattr(a, "__len__"):
            assert np.allclose(list(a), list(b), rtol=0, atol=1e-15), f
        else:
            assert a == b, f
    assert p.init_bite == 0.015 and p.min_score_diff == 500.0 and p.antipodal_mode == 1  # grasp_detector.cpp:43,49,53


def test_invalid_params_rejected():
    p = capi.default_params()
    p.num_orientations = 0
    with pytest.raises(capi.Ag2Error) as e:
        capi.Context(p)
    assert e.value.code in (capi.AG2_E_INVALID, capi.AG2_E_NO_DEVICE)


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="GPU present")
def test_no_cpu_fallback():
    with pytest.raises(capi.Ag2Error) as e:
        capi.Context(capi.default_params())
    assert e.value.code == capi.AG2_E_NO_DEVICE


def test_native_merge_of_per_rank_selections_equals_the_python_rule():
    """ag2_merge_selected (host-only C) against sharded.merge_selected on random per-rank blocks with
    padding, score ties and sample ids that interleave across ranks."""
    from agile_grasp2_b200 import sharded
    rng = np.random.default_rng(5)
    for world, cap, k in ((1, 8, 5), (2, 50, 50), (3, 16, 7), (8, 50, 50), (4, 6, 0)):
        blocks = np.zeros((world, cap), capi.HYPOTHESIS_DTYPE)
        parts = []
        for r in range(world):
            n = int(rng.integers(0, cap + 1))
            b = blocks[r, :n]
            b["approach"] = rng.normal(size=(n, 3)) + 3.0
            b["score"] = rng.integers(-3, 4, size=n).astype(np.float64) * 0.5  # many ties
            b["sample_id"] = np.sort(rng.integers(0, 40, size=n))
            b["orientation_id"] = rng.integers(0, 8, size=n)
            b["width"] = rng.random(n)
            loc = b.copy()
            loc["sample_id"] = r + world * loc["sample_id"]
            parts.append(loc)
        want = sharded.merge_selected(parts, k)
        got = capi.merge_selected_blocks(blocks, k)
        assert got.tobytes() == want.tobytes(), (world, cap, k)


def test_headers_are_plain_c(tmp_path):
    """The boundary is a C ABI: include/ag2.h (and the sampler arithmetic shared with the oracle, ag2_detmath.h) must
    compile as C99 with no C++ or CUDA construct, which is what a cgo / JNI / ctypes binding needs."""
    import subprocess

    src = tmp_path / "t.c"
    src.write_text('#include "ag2.h"\n#include "ag2_detmath.h"\n'
                   "int main(void) { ag2_params p; ag2_is_params q; (void)p; (void)q; return ag2_det_normal(0.3, 0.4) > 9.0; }\n")
    r = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                        "-fsyntax-only", str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_every_declaration_block_cites_the_reference_or_says_it_has_no_counterpart():
    """include/ag2.h: each paragraph that declares entry points names the reference interface it replaces
    (file.ext:line) or says that the reference has no counterpart."""
    src = open(os.path.join(ROOT, "include", "ag2.h")).read()
    cite = re.compile(r"\b\w+\.(cpp|h|prototxt|msg|launch):\d+|msg/\w+\.msg|[Nn]o (reference )?counterpart")
    decl = re.compile(r"^\s*(?:int|const char\*)\s+(ag2_\w+)\s*\(", re.M)
    bare = [names for block in re.split(r"\n\s*\n", src) if (names := decl.findall(block)) and not cite.search(block)]
    assert not bare, bare
    # and, one level finer, at least three quarters of the entry points carry a citation in their OWN comment
    own = 0
    decls = list(re.finditer(r"((?:/\*(?:(?!\*/).)*\*/\s*)?)^\s*(?:int|const char\*)\s+(ag2_\w+)\s*\(", src, re.M | re.S))
    for m in decls:
        own += bool(cite.search(re.sub(r"\s*\n\s*\*\s*", " ", m.group(1))))
    assert len(decls) >= 70 and own >= 0.75 * len(decls), (own, len(decls))


def test_reference_citations_point_into_existing_files():
    """Every `file.ext:line[-line]` citation of a reference file in the headers, the sources, the oracle, the tests and
    the design documents names a file of the reference checkout and lines inside it.  Skipped where the checkout is
    absent (the GPU box)."""
    import subprocess

    ref = "/root/reference"
    if not os.path.isdir(ref):
        pytest.skip("reference checkout not present")
    lengths = {}
    for root, _, files in os.walk(ref):
        if ".git" in root:
            continue
        for f in files:
            try:
                n = sum(1 for _ in open(os.path.join(root, f), errors="ignore"))
            except OSError:
                continue
            lengths[f] = max(lengths.get(f, 0), n)
    tracked = subprocess.run(["git", "ls-files"], cwd=ROOT, capture_output=True, text=True).stdout.split()
    if not tracked:
        pytest.skip("not a git checkout")
    ours = [p for p in tracked if p.endswith((".md", ".h", ".py", ".cu", ".cuh", ".cpp"))
            and not p.startswith(("SURVEY", "VERDICT", "ADVICE", "PAPERS", "SNIPPETS"))]
    own_names = {os.path.basename(p) for p in tracked}
    pat = re.compile(r"\b([A-Za-z_0-9]+\.(?:cpp|h|prototxt|launch|msg)):(\d+)(?:-(\d+))?((?:,\s?:?\d+(?:-\d+)?)*)")
    bad, checked = [], 0
    for p in ours:
        text = open(os.path.join(ROOT, p), errors="ignore").read()
        for m in pat.finditer(text):
            name = m.group(1)
            if name not in lengths:
                if name not in own_names:
                    bad.append((p, m.group(0), "no such reference file"))
                continue
            nums = [int(v) for v in re.findall(r"\d+", m.group(0)[len(name):])]
            checked += 1
            if max(nums) > lengths[name] or min(nums) < 1:
                bad.append((p, m.group(0), "file has %d lines" % lengths[name]))
    assert checked > 300 and not bad, bad[:10]
