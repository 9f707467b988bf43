"""INTEGRATION.md section 1 documents the glue a maintainer adds to the reference (`src/cuda/CudaBackend.h`).  This test
keeps it honest: the code block is extracted from the document, compiled against the reference's own `src/gridParams.h`
and linked with libgpusolve_cuda.so.  Needs the reference tree (this container); skipped on the GPU box."""
import os
import re
import subprocess

import pytest

from cases import ROOT

REF = os.environ.get("GPUSOLVE_REF", "/root/reference")


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "src", "gridParams.h")), reason="reference tree not present")
def test_documented_backend_glue_compiles_and_links(gs, tmp_path):
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    m = re.search(r"## 1\. `src/cuda/CudaBackend\.h`.*?```cpp\n(.*?)```", doc, flags=re.S)
    assert m, "section 1 code block not found"
    cuda_dir = tmp_path / "src" / "cuda"
    cuda_dir.mkdir(parents=True)
    (cuda_dir / "CudaBackend.h").write_text(m.group(1))
    os.symlink(os.path.join(REF, "src", "gridParams.h"), tmp_path / "src" / "gridParams.h")
    (tmp_path / "src" / "main_glue.cpp").write_text(
        '#include <iostream>\n#include "cuda/CudaBackend.h"\n'
        "int main() {\n"
        "    GridParams g{};\n"
        "    g.maxiter = 1; g.tol = 1e-5; g.gridDim = {7, 7, 7}; g.mode = GridParams::LINEAR; g.preSmoothing = 2; g.postSmoothing = 2;\n"
        "    g.omega = 0.8; g.gamma = 1.0; g.h = 1.0 / 8;\n"
        "    g.stencil.values = {6, -1, -1, -1, -1, -1, -1};\n"
        "    g.stencil.offsets = {{{0,0,0},{1,0,0},{-1,0,0},{0,1,0},{0,-1,0},{0,0,1},{0,0,-1}}};\n"
        "    gs_params p = CudaBackend::toPod(g);\n"
        "    if (p.dim[2] != 7 || p.stencil_off[6][2] != -1 || p.stencil_val[0] != 6.0 || p.mode != GS_MODE_LINEAR) return 2;\n"
        '    try { CudaBackend::solve(g); } catch (const std::exception& e) { std::cout << "Exception: " << e.what() << "\\n"; return 3; }\n'
        "    return 0;\n}\n")
    exe = tmp_path / "glue"
    libdir = os.path.dirname(gs.LIB)
    cmd = ["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", str(tmp_path / "src"), str(tmp_path / "src" / "main_glue.cpp"),
           "-o", str(exe), "-L", libdir, "-lgpusolve_cuda", f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    # without a GPU the documented error path must fire (exception text from gs_last_error), with one the solve runs
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert r.returncode in (0, 3), (r.returncode, r.stdout, r.stderr)
    if r.returncode == 3:
        assert "Exception:" in r.stdout
